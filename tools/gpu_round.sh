#!/bin/bash
# One GPU-box pass for the round's evidence: parity tests, bench line, reference arm, ncu launch list + full capture of the
# dominant kernel (bench command), dimension sweep, BASELINE config timings, and full captures of three more regimes.
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
timeout 900 python tools/dimension_sweep.py > gpurun_out/dimension_sweep.json 2> gpurun_out/dimension_sweep.err
timeout 600 python tools/config_timings.py > gpurun_out/config_timings.json 2> gpurun_out/config_timings.err
# other regimes: IMHK at n = 1024, C3 (wide windows, n = 512), n = 64 (panel-bound)
C1="python tools/quick_imhk.py 1024 170496 10 1"
$C1 > gpurun_out/plain_imhk.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'klein_chain' -s 2 -c 1 -f -o gpurun_out/prof_imhk $C1 > gpurun_out/ncu_imhk.log 2>&1
C2="python tools/quick_c3.py 1"
$C2 > gpurun_out/plain_c3.log 2>&1 &&
ncu --set full --clock-control none -k regex:'klein_chain' -s 1 -c 1 -f -o gpurun_out/prof_c3 $C2 > gpurun_out/ncu_c3.log 2>&1
C3="python tools/quick_klein.py 64 1363968 1"
$C3 > gpurun_out/plain_n64.log 2>&1 &&
ncu --set full --clock-control none -k regex:'klein_chain' -s 1 -c 1 -f -o gpurun_out/prof_n64 $C3 > gpurun_out/ncu_n64.log 2>&1
ls -la gpurun_out
