#!/bin/bash
# N GPUs: numeric proof of the exchange for the three recipes, then the default bench line (with decode) and, if asked, config E
N=${1:-2}; tag=${2:-r02zc}; extra=${3:-}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
run() { name=$1; shift; timeout 900 $TR bench.py --gpus $N "$@" > gpurun_out/${tag}_${name}.json 2> gpurun_out/${tag}_${name}.err; echo "== $name rc=$?"; tail -1 gpurun_out/${tag}_${name}.json | cut -c1-900; tail -2 gpurun_out/${tag}_${name}.err | cut -c1-200; }
run verify_default --verify-dp
[ "$N" -le 2 ] && run verify_unfreeze_lm --verify-dp --recipe unfreeze_lm
run verify_unfreeze_pc_base --verify-dp --encoder base --recipe unfreeze_pc
if [ "$extra" = "bench" ]; then
  run bench_default --steps 10 --warmup 3 --no-gpu-baseline --no-cpu-baseline --no-traj-tokens
  run bench_configE --steps 8 --warmup 3 --encoder base --recipe unfreeze_pc --no-decode --no-gpu-baseline --no-cpu-baseline --no-traj-tokens
fi
