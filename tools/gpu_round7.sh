#!/bin/bash
tag=${1:-r02h}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${tag}_gputests.log 2>&1; tail -8 gpurun_out/${tag}_gputests.log | cut -c1-300
ESB_TOP_LAYER_ROWS=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_full_top.json 2>gpurun_out/${tag}_bench_full_top.err; cut -c1-200 gpurun_out/${tag}_bench_full_top.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_rows_top.json 2>gpurun_out/${tag}_bench_rows_top.err; cut -c1-200 gpurun_out/${tag}_bench_rows_top.json
ESB_TOP_LAYER_ROWS=0 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_full_top_b.json 2>gpurun_out/${tag}_bench_full_top_b.err; cut -c1-200 gpurun_out/${tag}_bench_full_top_b.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_rows_top_b.json 2>gpurun_out/${tag}_bench_rows_top_b.err; cut -c1-200 gpurun_out/${tag}_bench_rows_top_b.json
