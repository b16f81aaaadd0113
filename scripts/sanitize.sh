#!/bin/bash
# compute-sanitizer over the small-scene GPU suite.  ONE tool per gpurun call (B200_PROFILING.md: running
# several tools in one call has left GPUs unusable):
#     gpurun -- 'bash scripts/sanitize.sh memcheck'      then racecheck, then synccheck, each in its own call
# Writes gpurun_out/sanitize_<tool>.log; copy the tail (ERROR SUMMARY line + hazards) to profiles/rNN_sanitize_<tool>.txt.
# The selected tests go through every tcgen05 / TMEM / mbarrier kernel (student gather-GEMM, pooling, logits) and the
# whole-scene path on the 12k-point scene; the full-size tests are left out (the tools slow kernels down 10-100x).
set -x
tool=${1:-memcheck}
mkdir -p gpurun_out
sel="test_scene_ or reference_run or pool_tensor_core or logits_tensor_core or student_tensor_core or nn1_fill or knn_sparse or host_pipeline"
extra=""
[ "$tool" = memcheck ] && extra="--leak-check no"
timeout 1700 compute-sanitizer --tool "$tool" $extra --error-exitcode 99 --print-limit 50 \
    python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "$sel" > gpurun_out/sanitize_$tool.log 2>&1
echo "sanitizer rc=$?" >> gpurun_out/sanitize_$tool.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|passed|failed|sanitizer rc" gpurun_out/sanitize_$tool.log | tail -n 20
