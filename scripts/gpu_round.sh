#!/bin/bash
# One full GPU visit: parity tests, smoke, bench (ours + reference arm), ncu launch list, ncu full captures of the
# dominant kernels.  Outputs land in gpurun_out/<tag>_*; profiles/summarize.py + profiles/make_traffic.py turn them into
# profiles/<tag>_summary.txt and profiles/traffic.json.      usage: bash scripts/gpu_round.sh [tag]
set -x
T=${1:-round}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/${T}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/${T}_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?" >> gpurun_out/${T}_bench.err
timeout 1200 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/${T}_bench_ref.json 2> gpurun_out/${T}_bench_ref.err
PLAIN="python bench.py --steps 1 --warmup 0 --pool 1 --no-cpu-baseline"
timeout 600 $PLAIN > gpurun_out/${T}_plain.log 2>&1 && {
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${T}_launches.csv \
   $PLAIN > gpurun_out/${T}_ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'gather_gemm_tc_kernel' -s 3 -c 3 \
   -o gpurun_out/${T}_prof_student -f $PLAIN > gpurun_out/${T}_ncu_student.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'pool_tc_kernel' -s 4 -c 2 \
   -o gpurun_out/${T}_prof_pool -f $PLAIN > gpurun_out/${T}_ncu_pool.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on \
   -k regex:'lift_mean_kernel|knn_grid_kernel|affinity_softmax_kernel|nn1_query_kernel|scatter_mean_kernel|pool_tc_plan_kernel|project_views_kernel|tc_split_input_kernel|logits_split_kernel' \
   -c 12 -o gpurun_out/${T}_prof_others -f $PLAIN > gpurun_out/${T}_ncu_others.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'student_reduce_kernel' -s 2 -c 2 \
   -o gpurun_out/${T}_prof_reduce -f $PLAIN > gpurun_out/${T}_ncu_reduce.log 2>&1
}
tail -n 3 gpurun_out/${T}_pytest_gpu.log; tail -n 2 gpurun_out/${T}_smoke.log; cut -c1-400 gpurun_out/${T}_bench.json; cut -c1-300 gpurun_out/${T}_bench_ref.json
