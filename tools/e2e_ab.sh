mkdir -p gpurun_out
{
python tools/quick_e2e.py 1024 1
LGS_HOST_PACK=0 python tools/quick_e2e.py 1024 1
python tools/quick_e2e.py 1024 4
} > gpurun_out/quick_e2e.log 2>&1
