#!/bin/bash
mkdir -p gpurun_out
{
for lib in tpw1 default; do
  echo "=== lib=$lib"
  if [ $lib != default ]; then export LGS_B200_LIB=$PWD/tools/bin/liblgs_$lib.so; else unset LGS_B200_LIB; fi
  timeout 300 python tools/quick_klein.py 1024 170496 5 2>&1 | grep "want_v"
  timeout 300 python tools/quick_imhk.py 1024 170496 10 3 2>&1 | grep "chain-steps"
  timeout 300 python tools/quick_imhk.py 512 340992 10 3 2>&1 | grep "chain-steps"
  timeout 300 python tools/quick_c3.py 2>&1 | tail -1
  timeout 300 python tools/quick_e2e.py 2>&1 | tail -4 | head -1
done
} > gpurun_out/la_ab.log 2>&1
