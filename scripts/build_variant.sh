#!/bin/bash
# A/B build of the library with extra -D flags for ONE source file (default student_tc.cu; SRC=pool_tc to pick another):
#   scripts/build_variant.sh NAME -DGP_STU_CL=2 ...        SRC=pool_tc scripts/build_variant.sh dense -DGP_POOL_DENSE=1
# -> geopurify_b200/libgp_v_NAME.so (select with GEOPURIFY_B200_LIB=...; delete afterwards, variants travel with gpurun)
set -e
cd "$(dirname "$0")/../geopurify_b200/csrc"
src=${SRC:-student_tc}
name=$1; shift
make -s
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC "$@" -c $src.cu -o build/${src}_v_$name.o
objs=$(ls build/*.o | grep -v "build/${src}.o" | grep -v "_v_")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libgp_v_$name.so $objs build/${src}_v_$name.o -lcudart
rm -f build/${src}_v_$name.o
echo built ../libgp_v_$name.so
