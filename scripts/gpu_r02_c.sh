#!/bin/bash
# Round 2, visit C: pool v3 (persistent CTA, TMEM double buffer, split rings, async cp.async completion) parity + timing
# (compact / dense plan), pair-mode student variants, TMEM load-shape probe, ncu of the pool kernel.
set -x
mkdir -p gpurun_out
timeout 120 scripts/probes/tmem_ld_shapes > gpurun_out/c_probe.log 2>&1
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -q -s -k "pool or fullsize" > gpurun_out/c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c_pytest.log
timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/c_pool.log 2>&1
GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_pooldense.so timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/c_pool_dense.log 2>&1
GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_poolc_3_11.so timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/c_pool_3_11.log 2>&1
for v in default p4_64_4 p3_64_5 p5_64_4; do
  if [ "$v" = default ]; then unset GEOPURIFY_B200_LIB; else export GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_$v.so; fi
  timeout 300 python scripts/bench_student.py 150000 >> gpurun_out/c_student_variants.log 2>&1
done
unset GEOPURIFY_B200_LIB
timeout 300 python scripts/bench_pool.py 150000 96 3 > gpurun_out/c_plain_pool.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:pool_tc_kernel -s 4 -c 2 \
   -o gpurun_out/c_prof_pool -f python scripts/bench_pool.py 150000 96 3 > gpurun_out/c_ncu_pool.log 2>&1
tail -n 12 gpurun_out/c_pytest.log; grep lib= gpurun_out/c_student_variants.log; cat gpurun_out/c_pool.log gpurun_out/c_pool_dense.log gpurun_out/c_pool_3_11.log; head -40 gpurun_out/c_probe.log
