#!/bin/bash
tag=${1:-r02j}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_gputests.log 2>&1; tail -4 gpurun_out/${tag}_gputests.log | cut -c1-300
timeout 200 python tools/attn_bench.py > gpurun_out/${tag}_attn_bench.txt 2>&1; cat gpurun_out/${tag}_attn_bench.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench.json 2>gpurun_out/${tag}_bench.err; cut -c1-200 gpurun_out/${tag}_bench.json
