#!/bin/bash
mkdir -p gpurun_out
{
timeout 900 python -m pytest tests/test_gpu_truncated.py tests/test_gpu_chain.py tests/test_gpu_configs.py tests/test_gpu_native.py -x -q 2>&1 | tail -3
timeout 300 python tools/quick_c3.py 2>&1 | tail -1
LGS_WIDE_TABLE=0 timeout 300 python tools/quick_c3.py 2>&1 | tail -1
LGS_CHAIN_TPW=2 timeout 300 python tools/quick_c3.py 2>&1 | tail -1
} > gpurun_out/la_ab.log 2>&1
