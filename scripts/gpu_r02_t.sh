#!/bin/bash
# kNN shell table: exactness tests + timing
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py tests/test_configs_gpu.py -m gpu -q -k "knn or fullsize_bf16 or whole_scene or scene_labels" > gpurun_out/t_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t_pytest.log
tail -n 3 gpurun_out/t_pytest.log
timeout 300 python scripts/bench_knn.py 2>&1 | tail -6
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/t_bench.json 2> gpurun_out/t_bench.err
python -c "
import json; d=json.load(open('gpurun_out/t_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'], d['stages']['knn'])"
