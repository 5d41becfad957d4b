#!/bin/bash
# second GPU visit: re-run the previously failing tests, the fused-kernel micro-benchmark, and A/B bench lines (fusions off / on)
tag=${1:-r02b}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/${tag}_gputests.log 2>&1; tail -12 gpurun_out/${tag}_gputests.log
timeout 300 python tools/fused_bench.py > gpurun_out/${tag}_fused_bench.txt 2>&1; cat gpurun_out/${tag}_fused_bench.txt
ESB_FUSE_SWIGLU=0 ESB_FUSE_DSWIGLU=0 ESB_FUSE_ROPE=0 ESB_FUSE_ROPE_BWD=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_unfused.json 2>gpurun_out/${tag}_bench_unfused.err; cut -c1-400 gpurun_out/${tag}_bench_unfused.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline > gpurun_out/${tag}_bench_fused.json 2>gpurun_out/${tag}_bench_fused.err; cut -c1-400 gpurun_out/${tag}_bench_fused.json; grep -o '"traj_tokens".*' gpurun_out/${tag}_bench_fused.json | cut -c1-900
