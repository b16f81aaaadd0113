#!/bin/bash
# A/B build of the library with extra -D flags for student_tc.cu:  scripts/build_variant.sh NAME -DGP_STU_RING=2 ...
# -> geopurify_b200/libgp_v_NAME.so (select with GEOPURIFY_B200_LIB=...; delete afterwards, variants travel with gpurun)
set -e
cd "$(dirname "$0")/../geopurify_b200/csrc"
name=$1; shift
make -s
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC "$@" -c student_tc.cu -o build/student_tc_$name.o
objs=$(ls build/*.o | grep -v "student_tc" )
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libgp_v_$name.so $objs build/student_tc_$name.o -lcudart
rm -f build/student_tc_$name.o
echo built ../libgp_v_$name.so
