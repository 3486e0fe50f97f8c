#!/bin/bash
# A/B the development variants in tools/bin on the GPU box: tools/ab.sh name1 name2 ...
mkdir -p gpurun_out
for v in "$@"; do
  echo "=== $v"
  export LGS_B200_LIB=$PWD/tools/bin/liblgs_$v.so
  python -m pytest tests/test_gpu_blocked.py -x -q 2>&1 | tail -1
  python tools/quick_klein.py 1024 17 5 2>&1 | grep -v "^setup\|smoothing\|Poorly"
done
