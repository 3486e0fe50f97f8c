#!/bin/bash
mkdir -p gpurun_out
{
export LGS_B200_LIB=$PWD/tools/bin/liblgs_stats2.so
for tpw in 3 2; do
  echo "=== tpw $tpw"
  export LGS_CHAIN_TPW=$tpw
  LGS_CHAIN_STATS=1 timeout 300 python tools/quick_imhk.py 1024 170496 10 1 2>&1 | grep "chain-steps\|chain stats" | tail -2
done
} > gpurun_out/stats_ab.log 2>&1
