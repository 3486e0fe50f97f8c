#!/bin/bash
mkdir -p gpurun_out
{
export LGS_B200_LIB=$PWD/tools/bin/liblgs_stats2.so
for la in 1 0; do
  echo "=== lookahead $la"
  export LGS_CHAIN_LOOKAHEAD=$la
  LGS_CHAIN_STATS=1 timeout 300 python tools/quick_imhk.py 1024 170496 10 1 2>&1 | grep "chain-steps\|chain stats" | tail -2
  LGS_CHAIN_STATS=1 timeout 300 python tools/quick_klein.py 1024 170496 1 2>&1 | grep "want_v\|chain stats" | tail -4
done
} > gpurun_out/stats_ab.log 2>&1
