#!/bin/bash
# workspace arena: parity subset, then the fixed list twice per stream count (run-to-run spread)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_configs_gpu.py -m gpu -q -x > gpurun_out/n_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/n_pytest.log
tail -n 3 gpurun_out/n_pytest.log
for rep in a b; do for st in 1 2; do
  timeout 600 python bench.py --scenes 312 --pool 8 --streams $st --no-cpu-baseline > gpurun_out/n_list_s${st}_$rep.json 2> gpurun_out/n_list_s${st}_$rep.err
  python -c "
import json; l=json.load(open('gpurun_out/n_list_s${st}_$rep.json')); print('list', $st, '$rep', l['value'], l['e2e']['value'], l['seconds'], l['clocks']['sm_mhz'], l['cuda_mallocs_in_timed_region'])"
done; done
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/n_bench.json 2> gpurun_out/n_bench.err
python -c "
import json; d=json.load(open('gpurun_out/n_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'])"
