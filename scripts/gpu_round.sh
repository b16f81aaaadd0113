#!/bin/bash
# One GPU visit: parity tests, smoke, bench (ours + reference arm), ncu launch list, ncu full captures of the two
# dominant kernels.  Outputs land in gpurun_out/; profiles/summarize.py turns them into profiles/rNN_*.txt.
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 12 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches.csv \
   python bench.py --steps 1 --warmup 0 --pool 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gather_gemm_tc_kernel' -s 3 -c 3 \
   -o gpurun_out/prof_student -f python bench.py --steps 1 --warmup 0 --pool 1 --no-cpu-baseline > gpurun_out/ncu_student.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pool_tc_kernel' -s 4 -c 2 \
   -o gpurun_out/prof_pool -f python bench.py --steps 1 --warmup 0 --pool 1 --no-cpu-baseline > gpurun_out/ncu_pool.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on \
   -k regex:'lift_mean_kernel|knn_grid_kernel|affinity_softmax_kernel|nn1_query_kernel|scatter_mean_kernel|pool_tc_plan_kernel|project_views_kernel|gather_gemm_tc_kernel<128>|tc_split_input_kernel|logits_split_kernel' \
   -c 14 -o gpurun_out/prof_others -f python bench.py --steps 1 --warmup 0 --pool 1 --no-cpu-baseline > gpurun_out/ncu_others.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'student_reduce_kernel' -s 2 -c 2 \
   -o gpurun_out/prof_reduce -f python bench.py --steps 1 --warmup 0 --pool 1 --no-cpu-baseline > gpurun_out/ncu_reduce.log 2>&1
tail -n 3 gpurun_out/pytest_gpu.log; tail -n 2 gpurun_out/smoke.log; tail -n 2 gpurun_out/bench.log | cut -c1-400
