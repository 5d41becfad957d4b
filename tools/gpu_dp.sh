#!/bin/bash
# Multi-GPU visit (gpurun --gpus N): numeric proof of the gradient exchange for the three recipes, then bench lines.
# usage: tools/gpu_dp.sh N tag [full]
N=${1:-2}; tag=${2:-r02_dp}; full=${3:-}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; timeout 900 $TR bench.py --gpus $N "$@" > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err; echo "== $name rc=$?"; tail -1 gpurun_out/${tag}_${name}.json | cut -c1-1200; tail -2 gpurun_out/${tag}_${name}.err | cut -c1-300; }
run verify_default --verify-dp
run verify_unfreeze_lm --verify-dp --recipe unfreeze_lm
run verify_unfreeze_pc_base --verify-dp --encoder base --recipe unfreeze_pc
run bench_default --steps 10 --warmup 3
if [ -n "$full" ]; then
  run bench_configE --steps 8 --warmup 3 --encoder base --recipe unfreeze_pc
  run bench_unfreeze_lm --steps 6 --warmup 3 --recipe unfreeze_lm --batch 4 --no-decode
fi
