#!/bin/bash
set -x
mkdir -p gpurun_out
for st in 1 2; do
  timeout 600 python bench.py --scenes 312 --pool 8 --streams $st --no-cpu-baseline > gpurun_out/p_list_s${st}.json 2> gpurun_out/p_list_s${st}.err
  python -c "
import json; l=json.load(open('gpurun_out/p_list_s${st}.json')); print('list', $st, l['value'], l['e2e']['value'], l['seconds'], l['e2e']['scenes_per_s'], l['clocks']['sm_mhz'], l['cuda_mallocs_in_timed_region'])"
done
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/p_bench.json 2> gpurun_out/p_bench.err
python -c "
import json; d=json.load(open('gpurun_out/p_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])"
timeout 1200 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/p_bench_ref.json 2> gpurun_out/p_bench_ref.err
cut -c1-200 gpurun_out/p_bench_ref.json
