#!/bin/bash
# A/B of the global lattice-point queue (LGS_CHAIN_GQ) on the GPU box.
mkdir -p gpurun_out
{
python -m pytest tests/test_gpu_chain.py tests/test_gpu_blocked.py tests/test_gpu_hostpack.py -x -q 2>&1 | tail -2
for gq in 1 0; do
  echo "=== LGS_CHAIN_GQ=$gq"
  LGS_CHAIN_GQ=$gq python tools/quick_klein.py 1024 170496 5 2>&1 | grep -v "^setup\|smoothing\|Poorly"
  LGS_CHAIN_GQ=$gq python tools/quick_klein.py 64 1363968 5 2>&1 | grep -v "^setup\|smoothing\|Poorly"
  LGS_CHAIN_GQ=$gq python tools/quick_c3.py 2>&1 | tail -2
  LGS_CHAIN_GQ=$gq python tools/quick_e2e.py 2>&1 | tail -4
done
} > gpurun_out/gq_ab.log 2>&1
