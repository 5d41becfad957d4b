#!/bin/bash
tag=${1:-r02f}
mkdir -p gpurun_out
timeout 200 python tools/gemm2_check.py --full > gpurun_out/${tag}_gemm2_check.json 2>gpurun_out/${tag}_gemm2_check.err; tail -1 gpurun_out/${tag}_gemm2_check.json | cut -c1-120; tail -1 gpurun_out/${tag}_gemm2_check.json | grep -o '"us".*'
ESB_GEMM_2CTA=1 timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${tag}_gputests_2cta.log 2>&1; tail -4 gpurun_out/${tag}_gputests_2cta.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_1cta.json 2>gpurun_out/${tag}_bench_1cta.err; cut -c1-200 gpurun_out/${tag}_bench_1cta.json
ESB_GEMM_2CTA=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_2cta.json 2>gpurun_out/${tag}_bench_2cta.err; cut -c1-200 gpurun_out/${tag}_bench_2cta.json
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_1cta_b.json 2>gpurun_out/${tag}_bench_1cta_b.err; cut -c1-200 gpurun_out/${tag}_bench_1cta_b.json
ESB_GEMM_2CTA=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_2cta_b.json 2>gpurun_out/${tag}_bench_2cta_b.err; cut -c1-200 gpurun_out/${tag}_bench_2cta_b.json
ESB_GEMM_2CTA=1 ESB_FUSE_DSWIGLU=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_2cta_dsw.json 2>gpurun_out/${tag}_bench_2cta_dsw.err; cut -c1-200 gpurun_out/${tag}_bench_2cta_dsw.json
timeout 300 python tools/fused_bench.py > gpurun_out/${tag}_fused_bench.txt 2>&1; sed -n 5,8p gpurun_out/${tag}_fused_bench.txt
