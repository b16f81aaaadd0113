#!/bin/bash
# two compute streams in the host pipeline + interleaved fixed list: parity tests, bench, 312-scene list
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -q -k "interleaved or host_pipeline or guard" > gpurun_out/l_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/l_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/l_bench.json 2> gpurun_out/l_bench.err; echo "rc=$?" >> gpurun_out/l_bench.err
timeout 900 python bench.py --scenes 312 --pool 8 --no-cpu-baseline > gpurun_out/l_list312.json 2> gpurun_out/l_list312.err; echo "rc=$?" >> gpurun_out/l_list312.err
tail -n 3 gpurun_out/l_pytest.log
python -c "
import json
d=json.load(open('gpurun_out/l_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'], d['clocks'])
l=json.load(open('gpurun_out/l_list312.json')); print('list', l['value'], l['e2e']['value'], l['seconds'], l['scenes_per_s'], l['e2e']['scenes_per_s'])"
tail -n 2 gpurun_out/l_bench.err gpurun_out/l_list312.err
