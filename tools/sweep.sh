#!/bin/bash
# sweep the sub-batch pipelining knobs on the GPU box
python -m pytest tests/test_gpu_blocked.py -x -q 2>&1 | tail -1
for cfg in "1 37888" "2 37888" "3 37888" "4 37888" "4 18944" "6 18944" "8 18944" "3 56832" "2 75776"; do
  set -- $cfg
  echo "=== nstreams=$1 sub=$2"
  LGS_NSTREAMS=$1 LGS_SUBBATCH=$2 python tools/quick_klein.py 1024 151552 5 2>&1 | grep "want_v"
done
