#!/bin/bash
# Round 2 multi-GPU lines: bash scripts/gpu_r02_multi.sh N [matterport]   (run under gpurun --gpus N)
# default weak-scaling bench, the fixed 312-scene list (strong scaling, histogram all-reduce timed), and optionally the
# Matterport-region shape (configs[3]) with one region per rank.
set -x
N=$1
mkdir -p gpurun_out
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $1 bench.py --gpus $N "${@:2}"; }
timeout 900 bash -c "$(declare -f run); N=$N; run 29511 --steps 20 --warmup 5" > gpurun_out/m${N}_bench.json 2> gpurun_out/m${N}_bench.err; echo "rc=$?" >> gpurun_out/m${N}_bench.err
timeout 900 bash -c "$(declare -f run); N=$N; run 29512 --scenes 312 --pool 8 --no-cpu-baseline" > gpurun_out/m${N}_list312.json 2> gpurun_out/m${N}_list312.err; echo "rc=$?" >> gpurun_out/m${N}_list312.err
if [ "$2" = matterport ]; then
timeout 1500 bash -c "$(declare -f run); N=$N; run 29513 --steps 6 --warmup 3 --pool 1 --points 500000 --views 100 --image 640x512 --classes 160 --vis-thres 0.02 --room 12x10x3 --no-cpu-baseline" > gpurun_out/m${N}_matterport.json 2> gpurun_out/m${N}_matterport.err; echo "rc=$?" >> gpurun_out/m${N}_matterport.err
fi
for f in gpurun_out/m${N}_*.json; do echo "== $f"; grep -v "^\[\|W1\|^\*" $f | cut -c1-260; done
tail -n 3 gpurun_out/m${N}_*.err
