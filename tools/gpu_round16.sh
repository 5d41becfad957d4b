#!/bin/bash
tag=${1:-r02zb}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_ops_gpu.py tests/test_trainer_dp_gpu.py tests/test_hd128_gpu.py -m gpu -q -x > gpurun_out/${tag}_tests.log 2>&1; tail -5 gpurun_out/${tag}_tests.log | cut -c1-300
timeout 200 python tools/attn_bench.py > gpurun_out/${tag}_attn_bench.txt 2>&1; cat gpurun_out/${tag}_attn_bench.txt
