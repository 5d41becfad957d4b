#!/bin/bash
tag=${1:-r02k}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_ops_gpu.py tests/test_hd128_gpu.py tests/test_gemm_gpu.py -m gpu -q -x > gpurun_out/${tag}_gputests.log 2>&1; tail -3 gpurun_out/${tag}_gputests.log | cut -c1-300
timeout 200 python tools/attn_bench.py > gpurun_out/${tag}_attn_bench.txt 2>&1; cat gpurun_out/${tag}_attn_bench.txt
