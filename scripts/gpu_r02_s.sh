#!/bin/bash
# pool epilogue: register-direct stores (variant) vs smem transpose (default): parity + timing
set -x
mkdir -p gpurun_out
export GEOPURIFY_B200_LIB=$PWD/geopurify_b200/libgp_v_pooldirect.so
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_fullsize_gpu.py -m gpu -q -k "pool or fullsize_bf16 or fullsize_pooling" > gpurun_out/s_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/s_pytest.log
tail -n 3 gpurun_out/s_pytest.log
for rep in 1 2; do
timeout 300 python scripts/bench_pool.py 150000 96 19 2>&1 | grep rowwise
done
unset GEOPURIFY_B200_LIB
for rep in 1 2; do
timeout 300 python scripts/bench_pool.py 150000 96 19 2>&1 | grep rowwise
done
