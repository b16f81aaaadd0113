#!/bin/bash
# Round 2, visit F: lift kernel with field-major cameras (bench stage table) + compute-sanitizer memcheck
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "lift or scene_" > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/f_bench.json 2> gpurun_out/f_bench.err; echo "rc=$?" >> gpurun_out/f_bench.err
bash scripts/sanitize.sh memcheck
tail -n 3 gpurun_out/f_pytest.log; python -c "
import json; d=json.load(open('gpurun_out/f_bench.json')); print(d['value'], d['e2e']['value']); print({k:(v['ms'],v['frac']) for k,v in d['stages'].items()})"
