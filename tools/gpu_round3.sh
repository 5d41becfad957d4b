#!/bin/bash
# GPU visit: whole GPU suite, fused-kernel micro-benchmark, the default bench line, then (last, isolated) the CTA-pair GEMM check
tag=${1:-r02c}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${tag}_gputests.log 2>&1; tail -15 gpurun_out/${tag}_gputests.log
timeout 300 python tools/fused_bench.py > gpurun_out/${tag}_fused_bench.txt 2>&1; cat gpurun_out/${tag}_fused_bench.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2>gpurun_out/${tag}_bench.err; cut -c1-300 gpurun_out/${tag}_bench.json; tail -3 gpurun_out/${tag}_bench.err
timeout 200 python tools/gemm2_check.py --full > gpurun_out/${tag}_gemm2.json 2>gpurun_out/${tag}_gemm2.err; echo "gemm2 rc=$?"; cat gpurun_out/${tag}_gemm2.json; tail -5 gpurun_out/${tag}_gemm2.err
nvidia-smi --query-gpu=name,clocks.sm --format=csv,noheader; dmesg 2>/dev/null | grep -i xid | tail -3
