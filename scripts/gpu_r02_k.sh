#!/bin/bash
# interleaved scenes on 2 / 3 streams vs 1 (device-resident value)
set -x
mkdir -p gpurun_out
for st in 1 2 3; do
timeout 600 python bench.py --steps 40 --warmup 5 --streams $st --no-cpu-baseline > gpurun_out/k_bench_s$st.json 2> gpurun_out/k_bench_s$st.err
python -c "
import json; d=json.load(open('gpurun_out/k_bench_s$st.json')); print('streams', $st, d['value'], d['ms_per_step'], d['clocks'], d['miou_synthetic'])"
done
