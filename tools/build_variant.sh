#!/bin/bash
# usage: tools/build_variant.sh NAME [extra nvcc -D flags...]   ->  tools/bin/liblgs_NAME.so  (development A/B builds)
set -e
cd "$(dirname "$0")/.."
name=$1; shift
C=lattice-gaussian-mcmc_b200/csrc
mkdir -p tools/bin/obj_$name
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
$NV -fmad=false "$@" -c $C/lgs_ordered.cu -o tools/bin/obj_$name/ordered.o &
$NV "$@" -c $C/lgs_blocked.cu -o tools/bin/obj_$name/blocked.o &
$NV "$@" -c $C/lgs_chain.cu -o tools/bin/obj_$name/chain.o &
$NV "$@" -c $C/lgs_diag.cu -o tools/bin/obj_$name/diag.o &
$NV "$@" -c $C/lgs_capi.cu -o tools/bin/obj_$name/capi.o &
$NV -Xcompiler -mfma -c $C/lgs_reduce.cpp -o tools/bin/obj_$name/reduce.o &
$NV -Xcompiler -mavx2 -c $C/lgs_hostpack.cpp -o tools/bin/obj_$name/hostpack.o &
wait
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o tools/bin/liblgs_$name.so tools/bin/obj_$name/*.o -Xcompiler -fPIC
echo tools/bin/liblgs_$name.so
