#!/bin/bash
# Short GPU-box pass: parity tests, bench line, IMHK / Klein quick timings, config timings.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/tests.log 2>&1; echo "tests exit $?" >> gpurun_out/tests.log
timeout 600 python bench.py --skip-cpu > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench exit $?" >> gpurun_out/bench.err
timeout 300 python tools/quick_imhk.py 1024 170496 10 3 > gpurun_out/quick_imhk.log 2>&1
timeout 300 python tools/quick_klein.py > gpurun_out/quick_klein.log 2>&1
timeout 600 python tools/config_timings.py > gpurun_out/config_timings.json 2> gpurun_out/config_timings.err
ls -la gpurun_out
