#!/bin/bash
tag=${1:-r02e}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_gpu.py tests/test_variants_gpu.py tests/test_hd128_gpu.py -m gpu -q > gpurun_out/${tag}_gputests.log 2>&1; tail -4 gpurun_out/${tag}_gputests.log
ESB_FUSE_DSWIGLU=1 timeout 300 python -m pytest tests/test_gemm_gpu.py tests/test_hd128_gpu.py -m gpu -q -k "dswiglu or trunk_fused or hd128_forward" > gpurun_out/${tag}_gputests_dswiglu.log 2>&1; tail -3 gpurun_out/${tag}_gputests_dswiglu.log
timeout 300 python tools/fused_bench.py > gpurun_out/${tag}_fused_bench.txt 2>&1; sed -n 5,9p gpurun_out/${tag}_fused_bench.txt
for d in 0 4; do echo "== ESB_GEMM2_DBG=$d"; ESB_GEMM2_DBG=$d timeout 120 python tools/gemm2_check.py --time 2>&1 | tail -1 | tee gpurun_out/${tag}_gemm2_dbg$d.json; done
timeout 200 python tools/gemm2_check.py > gpurun_out/${tag}_gemm2_check.json 2>&1; tail -1 gpurun_out/${tag}_gemm2_check.json | cut -c1-200
timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline > gpurun_out/${tag}_bench_a.json 2>gpurun_out/${tag}_bench_a.err; cut -c1-200 gpurun_out/${tag}_bench_a.json
ESB_FUSE_DSWIGLU=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_dswiglu.json 2>gpurun_out/${tag}_bench_dswiglu.err; cut -c1-200 gpurun_out/${tag}_bench_dswiglu.json
ESB_GEMM_2CTA=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-decode --no-cpu-baseline --no-gpu-baseline --no-traj-tokens > gpurun_out/${tag}_bench_2cta.json 2>gpurun_out/${tag}_bench_2cta.err; cut -c1-200 gpurun_out/${tag}_bench_2cta.json
grep -o '"traj_tokens".*' gpurun_out/${tag}_bench_a.json | cut -c1-420
