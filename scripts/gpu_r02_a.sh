#!/bin/bash
# Round 2, visit A: new full-size oracle parity tests + pair-mode student variants + ncu full with L2 metrics.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.limit --format=csv > gpurun_out/a_smi.txt 2>&1
nproc >> gpurun_out/a_smi.txt; free -g >> gpurun_out/a_smi.txt
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a_pytest.log
for v in default p5_32_3 p5_32_6 p4_64_3 p8_32_4; do
  if [ "$v" = default ]; then unset GEOPURIFY_B200_LIB; else export GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_$v.so; fi
  timeout 300 python scripts/bench_student.py 150000 >> gpurun_out/a_student_variants.log 2>&1
done
unset GEOPURIFY_B200_LIB
timeout 300 python scripts/bench_pool.py 150000 96 19 > gpurun_out/a_pool.log 2>&1
timeout 300 python scripts/bench_student.py 150000 > gpurun_out/a_plain_student.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'gather_gemm_tc_kernel<512>' -s 12 -c 2 \
   -o gpurun_out/a_prof_student -f python scripts/bench_student.py 150000 > gpurun_out/a_ncu_student.log 2>&1
timeout 300 python scripts/bench_pool.py 150000 96 3 > gpurun_out/a_plain_pool.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'pool_tc_kernel' -s 4 -c 2 \
   -o gpurun_out/a_prof_pool -f python scripts/bench_pool.py 150000 96 3 > gpurun_out/a_ncu_pool.log 2>&1
tail -n 15 gpurun_out/a_pytest.log; cat gpurun_out/a_student_variants.log | grep lib=; cat gpurun_out/a_pool.log
