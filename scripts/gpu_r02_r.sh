#!/bin/bash
# whole GPU suite with canaries around every workspace (memcheck substitute), then a default bench line (post-reorder check)
set -x
mkdir -p gpurun_out
GEOPURIFY_B200_GUARD=1 timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r_pytest_guard.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r_pytest_guard.log
tail -n 4 gpurun_out/r_pytest_guard.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/r_bench.json 2> gpurun_out/r_bench.err
python -c "
import json; d=json.load(open('gpurun_out/r_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'], d['gpu_launches'], d['clocks'])"
