#!/bin/bash
mkdir -p gpurun_out
{
timeout 1200 python -m pytest tests/test_gpu_chain.py tests/test_gpu_chain_replay.py tests/test_gpu_blocked.py tests/test_gpu_replay.py tests/test_gpu_diagnostics.py tests/test_gpu_configs.py -x -q 2>&1 | tail -5
for tpw in 0 1 2 3; do
  echo "=== LGS_CHAIN_TPW=$tpw (0 = default rule)"
  export LGS_CHAIN_TPW=$tpw
  timeout 300 python tools/quick_imhk.py 1024 170496 10 2 reference 4 2>&1 | grep "chain-steps"
done
unset LGS_CHAIN_TPW
for sg in 0 100; do
  echo "=== stagger $sg tpw 1/2"
  LGS_CHAIN_STAGGER=$sg LGS_CHAIN_TPW=1 timeout 300 python tools/quick_imhk.py 1024 170496 10 2 reference 4 2>&1 | grep "chain-steps"
  LGS_CHAIN_STAGGER=$sg LGS_CHAIN_TPW=2 timeout 300 python tools/quick_imhk.py 1024 170496 10 2 reference 4 2>&1 | grep "chain-steps"
done
timeout 300 python tools/quick_imhk.py 1024 170496 10 3 reference 1 2>&1 | grep "chain-steps"
LGS_IMHK_KEEP_IN_LAUNCH=0 timeout 300 python tools/quick_imhk.py 1024 170496 10 2 reference 4 2>&1 | grep "chain-steps"
} > gpurun_out/la_ab.log 2>&1
