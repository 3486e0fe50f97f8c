#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_gpu_chain.py tests/test_gpu_blocked.py tests/test_gpu_chain_replay.py tests/test_gpu_configs.py -x -q 2>&1 | tail -3
for cfg in "default 2 1" "default 0 1" "default 2 0"; do
  set -- $cfg
  echo "=== lib=$1 LOOKAHEAD=$2 STAGGER=$3"
  export LGS_CHAIN_LOOKAHEAD=$2 LGS_CHAIN_STAGGER=$3
  timeout 300 python tools/quick_imhk.py 1024 170496 10 3 2>&1 | grep "chain-steps"
  timeout 300 python tools/quick_klein.py 1024 170496 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 512 340992 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 256 681984 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_c3.py 2>&1 | tail -1
done
export LGS_B200_LIB=$PWD/tools/bin/liblgs_stats2.so
export LGS_CHAIN_LOOKAHEAD=2 LGS_CHAIN_STAGGER=1
LGS_CHAIN_STATS=1 timeout 300 python tools/quick_imhk.py 1024 170496 10 1 2>&1 | grep "chain stats" | tail -1
LGS_CHAIN_STATS=1 timeout 300 python tools/quick_klein.py 1024 170496 1 2>&1 | grep "chain stats" | tail -3
} > gpurun_out/la_ab.log 2>&1
