#!/bin/bash
tag=${1:-r02d}
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${tag}_gputests.log 2>&1; tail -6 gpurun_out/${tag}_gputests.log
timeout 300 python tools/fused_bench.py > gpurun_out/${tag}_fused_bench.txt 2>&1; cat gpurun_out/${tag}_fused_bench.txt
for d in 0 1 2 3; do echo "== ESB_GEMM2_DBG=$d"; ESB_GEMM2_DBG=$d timeout 120 python tools/gemm2_check.py --time 2>&1 | tail -1 | tee gpurun_out/${tag}_gemm2_dbg$d.json; done
