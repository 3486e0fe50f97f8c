mkdir -p gpurun_out
{
python -m pytest tests/test_gpu_chain.py tests/test_gpu_blocked.py tests/test_gpu_native.py -x -q 2>&1 | tail -2
for v in "" base "" base; do
  echo "=== variant '$v'"
  if [ -n "$v" ]; then export LGS_B200_LIB=$PWD/tools/bin/liblgs_$v.so; else unset LGS_B200_LIB; fi
  python tools/quick_klein.py 1024 170496 6 2>&1 | grep "want_v=True"
  python tools/quick_imhk.py 1024 170496 2>&1 | grep "chain-steps"
done
} > gpurun_out/ab.log 2>&1
