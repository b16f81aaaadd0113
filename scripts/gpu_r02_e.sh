#!/bin/bash
# Round 2, visit E: suite on the new lift kernel / validation / APE mirror, then every bench mode on 1 GPU:
# default (configs[1]), reference arm on the same scenes, fixed 312-scene list, C=200 (configs[2]), consensus lifting,
# Matterport-region shape (configs[3]).
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s -x > gpurun_out/e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e_pytest.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/e_bench.json 2> gpurun_out/e_bench.err; echo "rc=$?" >> gpurun_out/e_bench.err
timeout 1200 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/e_bench_ref.json 2> gpurun_out/e_bench_ref.err; echo "rc=$?" >> gpurun_out/e_bench_ref.err
timeout 900 python bench.py --scenes 312 --pool 8 --no-cpu-baseline > gpurun_out/e_bench_list312.json 2> gpurun_out/e_bench_list312.err; echo "rc=$?" >> gpurun_out/e_bench_list312.err
timeout 900 python bench.py --steps 20 --warmup 5 --classes 200 --no-cpu-baseline > gpurun_out/e_bench_c200.json 2> gpurun_out/e_bench_c200.err; echo "rc=$?" >> gpurun_out/e_bench_c200.err
timeout 900 python bench.py --steps 20 --warmup 5 --lifting consensus --no-cpu-baseline > gpurun_out/e_bench_consensus.json 2> gpurun_out/e_bench_consensus.err; echo "rc=$?" >> gpurun_out/e_bench_consensus.err
timeout 1200 python bench.py --steps 6 --warmup 3 --pool 1 --points 500000 --views 100 --image 640x512 --classes 160 --vis-thres 0.02 --room 12x10x3 --no-cpu-baseline > gpurun_out/e_bench_matterport.json 2> gpurun_out/e_bench_matterport.err; echo "rc=$?" >> gpurun_out/e_bench_matterport.err
tail -n 6 gpurun_out/e_pytest.log
for f in e_bench e_bench_ref e_bench_list312 e_bench_c200 e_bench_consensus e_bench_matterport; do echo "== $f"; cut -c1-330 gpurun_out/$f.json; tail -n 2 gpurun_out/$f.err; done
