#!/bin/bash
# One GPU-box pass: parity tests, bench line, reference arm, ncu launch list + full capture of the dominant kernel.
set -x
mkdir -p gpurun_out
nvidia-smi > gpurun_out/nvidia_smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
SMALL="python bench.py --steps 1 --warmup 3 --skip-cpu --skip-e2e"
$SMALL > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $SMALL > gpurun_out/ncu_list.log 2>&1
SMALL2="python bench.py --steps 1 --warmup 3 --skip-imhk --skip-cpu --skip-e2e"
$SMALL2 > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'klein_chain' -s 3 -c 1 -f -o gpurun_out/prof $SMALL2 > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
