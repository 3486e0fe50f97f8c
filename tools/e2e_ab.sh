mkdir -p gpurun_out
{
python -m pytest tests/test_gpu_chain.py tests/test_gpu_blocked.py -x -q 2>&1 | tail -2
for v in "" notripred; do
  echo "=== variant '$v'"
  if [ -n "$v" ]; then export LGS_B200_LIB=$PWD/tools/bin/liblgs_$v.so; fi
  python tools/quick_klein.py 1024 170496 6 2>&1 | grep -v "^setup\|smoothing\|Poorly"
  python tools/quick_imhk.py 2>&1 | grep -v "^setup\|smoothing\|Poorly" | tail -4
done
} > gpurun_out/ab.log 2>&1
