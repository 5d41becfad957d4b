#!/bin/bash
# same-box A/B of the training step under two environments: tools/gpu_ab_env.sh "VAR=a" "VAR=b" (each run twice, alternating)
for envs in "$1" "$2" "$1" "$2"; do
  env $envs timeout 300 python bench.py --steps 12 --warmup 3 --no-decode --no-gpu-baseline --no-cpu-baseline --no-traj-tokens 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$envs', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], 'gemm', d['roofline']['gemm_ms_per_step'])"
done
