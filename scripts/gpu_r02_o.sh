#!/bin/bash
# scratch arena for scene temporaries: whole GPU suite, smoke, the fixed list (1 / 2 streams, twice), default bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/o_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/o_pytest.log
tail -n 3 gpurun_out/o_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for rep in a b; do for st in 1 2; do
  timeout 600 python bench.py --scenes 312 --pool 8 --streams $st --no-cpu-baseline > gpurun_out/o_list_s${st}_$rep.json 2> gpurun_out/o_list_s${st}_$rep.err
  python -c "
import json; l=json.load(open('gpurun_out/o_list_s${st}_$rep.json')); print('list', $st, '$rep', l['value'], l['e2e']['value'], l['seconds'], l['clocks']['sm_mhz'], l['cuda_mallocs_in_timed_region'])"
done; done
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/o_bench.json 2> gpurun_out/o_bench.err
python -c "
import json; d=json.load(open('gpurun_out/o_bench.json')); print('bench', d['value'], d['e2e']['value'], d['ms_per_step'], d['miou_synthetic'])"
