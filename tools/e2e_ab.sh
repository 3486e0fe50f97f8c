mkdir -p gpurun_out
{
timeout 600 python -m pytest tests/test_gpu_hostpack.py -x -q 2>&1 | tail -3
python tools/quick_e2e.py 1024 1
LGS_ZC_DEVMEM=1 python tools/quick_e2e.py 1024 1
LGS_B200_LIB=$PWD/tools/bin/liblgs_nofence.so python tools/quick_e2e.py 1024 1
LGS_B200_LIB=$PWD/tools/bin/liblgs_nofence.so LGS_ZC_DEVMEM=1 python tools/quick_e2e.py 1024 1
} > gpurun_out/quick_e2e.log 2>&1
