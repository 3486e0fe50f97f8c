#!/bin/bash
mkdir -p gpurun_out
{
timeout 1200 python -m pytest tests/test_gpu_chain.py tests/test_gpu_blocked.py tests/test_gpu_native.py tests/test_gpu_configs.py tests/test_gpu_truncated.py -x -q 2>&1 | tail -5
for lib in default nosplit; do
  echo "=== lib=$lib"
  if [ $lib != default ]; then export LGS_B200_LIB=$PWD/tools/bin/liblgs_$lib.so; else unset LGS_B200_LIB; fi
  timeout 300 python tools/quick_imhk.py 1024 170496 10 2 reference 4 2>&1 | grep "chain-steps"
  timeout 300 python tools/quick_klein.py 1024 170496 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 256 681984 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 64 1363968 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_klein.py 16 4091904 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_c3.py 2>&1 | tail -1
done
} > gpurun_out/la_ab.log 2>&1
