mkdir -p gpurun_out
{
for g in 4 9 18 74 148; do LGS_ZC_GROUP=$g python tools/quick_e2e.py 1024 1; done
for g in 9 18; do LGS_ZC_GROUP=$g python tools/quick_e2e.py 1024 3; done
} > gpurun_out/quick_e2e.log 2>&1
