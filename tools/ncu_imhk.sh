#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/quick_imhk.py 1024 170496 10 1"
$CMD > gpurun_out/ncu_imhk_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'klein_chain' -s 2 -c 1 -f -o gpurun_out/prof_imhk $CMD > gpurun_out/ncu_imhk.log 2>&1
CMD2="python tools/quick_klein.py 1024 170496 1"
$CMD2 > gpurun_out/ncu_klein_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'klein_chain' -s 1 -c 1 -f -o gpurun_out/prof_klein $CMD2 > gpurun_out/ncu_klein.log 2>&1
ls -la gpurun_out
