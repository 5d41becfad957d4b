#!/bin/bash
tag=${1:-r02i}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${tag}_gputests.log 2>&1; tail -6 gpurun_out/${tag}_gputests.log | cut -c1-300
timeout 300 python tools/fused_bench.py > gpurun_out/${tag}_fused_bench.txt 2>&1; tail -6 gpurun_out/${tag}_fused_bench.txt
timeout 600 python bench.py --steps 8 --warmup 3 --recipe unfreeze_lm --batch 4 --no-decode --no-cpu-baseline > gpurun_out/${tag}_bench_unfreeze_lm_b4.json 2>gpurun_out/${tag}_bench_unfreeze_lm_b4.err; cut -c1-200 gpurun_out/${tag}_bench_unfreeze_lm_b4.json
timeout 600 python bench.py --steps 6 --warmup 3 --recipe unfreeze_lm --batch 8 --no-decode --no-cpu-baseline > gpurun_out/${tag}_bench_unfreeze_lm_b8.json 2>gpurun_out/${tag}_bench_unfreeze_lm_b8.err; cut -c1-200 gpurun_out/${tag}_bench_unfreeze_lm_b8.json; tail -2 gpurun_out/${tag}_bench_unfreeze_lm_b8.err | cut -c1-300
timeout 600 python bench.py --steps 8 --warmup 3 --encoder base --recipe unfreeze_pc --no-cpu-baseline > gpurun_out/${tag}_bench_configE.json 2>gpurun_out/${tag}_bench_configE.err; cut -c1-200 gpurun_out/${tag}_bench_configE.json
