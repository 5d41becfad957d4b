#!/bin/bash
tag=${1:-r02t}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gemm_chain_gpu.py tests/test_model_gpu.py tests/test_hd128_gpu.py tests/test_sampling_gpu.py -m gpu -q -x > gpurun_out/${tag}_decode_tests.log 2>&1
tail -4 gpurun_out/${tag}_decode_tests.log | cut -c1-300
DECODE_VARIANTS=chain,plain timeout 600 python tools/decode_variants.py 60 > gpurun_out/${tag}_decode_variants.txt 2>&1; tail -9 gpurun_out/${tag}_decode_variants.txt | cut -c1-400
