#!/bin/bash
# Round 2, visit I: final code -- whole GPU suite (no -x), smoke, both bench arms.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/i_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/i_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/i_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/i_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?" >> gpurun_out/i_bench.err
timeout 1200 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/i_bench_ref.json 2> gpurun_out/i_bench_ref.err
tail -n 4 gpurun_out/i_pytest_gpu.log; tail -n 2 gpurun_out/i_smoke.log; cut -c1-300 gpurun_out/i_bench.json; cut -c1-200 gpurun_out/i_bench_ref.json
