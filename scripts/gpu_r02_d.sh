#!/bin/bash
# Round 2, visit D: full GPU suite on the new defaults (pair-mode student with register-direct epilogue, pool v2' with
# async gather completion + pipelined expansion), A/B timings, bench line, ncu of student + pool.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -x > gpurun_out/d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/d_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/d_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/d_smoke.log
for v in default stu_smemepi stu_single stu_single_old; do
  if [ "$v" = default ]; then unset GEOPURIFY_B200_LIB; else export GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_$v.so; fi
  timeout 300 python scripts/bench_student.py 150000 >> gpurun_out/d_student_variants.log 2>&1
done
unset GEOPURIFY_B200_LIB
timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/d_pool.log 2>&1
GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_pooldense.so timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/d_pool_dense.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/d_bench.json 2> gpurun_out/d_bench.err; echo "bench rc=$?" >> gpurun_out/d_bench.err
timeout 300 python scripts/bench_student.py 150000 > gpurun_out/d_plain_student.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gather_gemm_tc_kernel -s 12 -c 2 \
   -o gpurun_out/d_prof_student -f python scripts/bench_student.py 150000 > gpurun_out/d_ncu_student.log 2>&1
timeout 300 python scripts/bench_pool.py 150000 96 3 > gpurun_out/d_plain_pool.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pool_tc_kernel -s 4 -c 2 \
   -o gpurun_out/d_prof_pool -f python scripts/bench_pool.py 150000 96 3 > gpurun_out/d_ncu_pool.log 2>&1
tail -n 8 gpurun_out/d_pytest.log; tail -2 gpurun_out/d_smoke.log; grep lib= gpurun_out/d_student_variants.log; cat gpurun_out/d_pool.log gpurun_out/d_pool_dense.log; cut -c1-600 gpurun_out/d_bench.json; tail -3 gpurun_out/d_bench.err
