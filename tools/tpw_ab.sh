#!/bin/bash
mkdir -p gpurun_out
{
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for lib in default; do
  echo "=== $lib"
  timeout 300 python tools/quick_imhk.py 1024 170496 10 3 2>&1 | grep "chain-steps"
  timeout 300 python tools/quick_imhk.py 1024 170496 10 3 exact 2>&1 | grep "chain-steps"
  timeout 300 python tools/quick_klein.py 1024 170496 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 256 681984 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 64 1363968 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 16 4091904 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_c3.py 2>&1 | tail -1
done
export LGS_B200_LIB=$PWD/tools/bin/liblgs_stats2.so
LGS_CHAIN_STATS=1 timeout 300 python tools/quick_imhk.py 1024 170496 10 1 2>&1 | grep "chain stats" | tail -1
LGS_CHAIN_STATS=1 timeout 300 python tools/quick_klein.py 1024 170496 1 2>&1 | grep "chain stats" | tail -3
} > gpurun_out/tpw_ab.log 2>&1
