#!/bin/bash
# N-GPU bench of the default recipe with the dp_timeline diagnostic, optionally under an environment (A/B)
N=${1:-2}; tag=${2:-r02x}; shift; shift
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
i=0
for envs in "$@"; do
  i=$((i+1))
  env $envs timeout 600 $TR bench.py --gpus $N --steps 10 --warmup 3 --no-decode --no-gpu-baseline --no-cpu-baseline --no-traj-tokens > gpurun_out/${tag}_n${N}_$i.json 2> gpurun_out/${tag}_n${N}_$i.err
  echo "== [$envs] rc=$?"; python - <<PY
import json
d=json.loads(open("gpurun_out/${tag}_n${N}_$i.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], json.dumps(d.get("dp_timeline")), "dp_check", (d.get("dp_check") or {}).get("ok"))
PY
  tail -2 gpurun_out/${tag}_n${N}_$i.err | cut -c1-300
done
