#!/bin/bash
# Round 2, visit B: pool v2 (2 CTAs/SM, compact plan) parity + timing vs the dense-plan build; more pair-mode student
# variants; ncu full of the student GEMM (default and best pair build).
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -q -s -k "pool or fullsize" > gpurun_out/b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b_pytest.log
timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/b_pool.log 2>&1
GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_pooldense.so timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/b_pool_dense.log 2>&1
for v in default p4_64_3 p4_64_4 p6_64_3 p3_64_4 s4_64_2; do
  if [ "$v" = default ]; then unset GEOPURIFY_B200_LIB; else export GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_$v.so; fi
  timeout 300 python scripts/bench_student.py 150000 >> gpurun_out/b_student_variants.log 2>&1
done
unset GEOPURIFY_B200_LIB
timeout 300 python scripts/bench_student.py 150000 > gpurun_out/b_plain_student.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gather_gemm_tc_kernel -s 12 -c 2 \
   -o gpurun_out/b_prof_student -f python scripts/bench_student.py 150000 > gpurun_out/b_ncu_student.log 2>&1
export GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_p4_64_3.so
timeout 300 python scripts/bench_student.py 150000 > gpurun_out/b_plain_student_pair.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gather_gemm_tc_kernel -s 12 -c 2 \
   -o gpurun_out/b_prof_student_pair -f python scripts/bench_student.py 150000 > gpurun_out/b_ncu_student_pair.log 2>&1
unset GEOPURIFY_B200_LIB
timeout 300 python scripts/bench_pool.py 150000 96 3 > gpurun_out/b_plain_pool.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pool_tc_kernel -s 4 -c 2 \
   -o gpurun_out/b_prof_pool -f python scripts/bench_pool.py 150000 96 3 > gpurun_out/b_ncu_pool.log 2>&1
tail -n 12 gpurun_out/b_pytest.log; grep lib= gpurun_out/b_student_variants.log; cat gpurun_out/b_pool.log gpurun_out/b_pool_dense.log
