#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_gpu_diagnostics.py tests/test_gpu_chain.py -x -q 2>&1 | tail -5
timeout 600 python bench.py --skip-cpu --skip-e2e > gpurun_out/bench_s.json 2> gpurun_out/bench_s.err; tail -3 gpurun_out/bench_s.err
} > gpurun_out/la_ab.log 2>&1
