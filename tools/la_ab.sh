#!/bin/bash
mkdir -p gpurun_out
{
for ch in 1 0; do
  echo "=== LGS_CHAIN=$ch"
  export LGS_CHAIN=$ch
  for cfg in "10 4091904" "16 4091904" "32 2727936" "64 1363968"; do
    set -- $cfg
    timeout 300 python tools/quick_klein.py $1 $2 5 2>&1 | grep "want_v\|per-batch"
  done
done
} > gpurun_out/la_ab.log 2>&1
