#!/bin/bash
mkdir -p gpurun_out
{
for tpw in 3 2; do
for sg in 25 50 75 100; do
  echo "=== LGS_CHAIN_TPW=$tpw STAGGER=$sg"
  export LGS_CHAIN_TPW=$tpw LGS_CHAIN_STAGGER=$sg
  timeout 300 python tools/quick_imhk.py 1024 170496 10 3 2>&1 | grep "chain-steps"
done
done
} > gpurun_out/la_ab.log 2>&1
