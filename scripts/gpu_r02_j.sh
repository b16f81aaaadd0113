#!/bin/bash
# small-scene efficiency: points/s and per-stage ms at 30k / 60k points
set -x
mkdir -p gpurun_out
for n in 30000 60000; do
timeout 600 python bench.py --steps 40 --warmup 5 --pool 2 --points $n --no-cpu-baseline > gpurun_out/j_bench_$n.json 2> gpurun_out/j_bench_$n.err
python -c "
import json; d=json.load(open('gpurun_out/j_bench_$n.json')); print($n, d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches']/d['steps']); print({k:v['ms'] for k,v in d['stages'].items()}, sum(v['ms'] for v in d['stages'].values()))"
done
