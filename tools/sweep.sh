#!/bin/bash
# sweep the sub-batch pipelining knobs on the GPU box
python -m pytest tests/test_gpu_blocked.py -x -q 2>&1 | tail -1
for S in 17 151552; do
for ns in 1 2 3 4; do
for sub in 18944 37888 75776; do
  echo "=== S=$S nstreams=$ns sub=$sub"
  LGS_NSTREAMS=$ns LGS_SUBBATCH=$sub python tools/quick_klein.py 1024 $S 5 2>&1 | grep "want_v"
done; done; done
