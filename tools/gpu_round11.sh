#!/bin/bash
# chain kernel: unit tests (bounded), decode tests, then the decode A/B
tag=${1:-r02n}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_chain_gpu.py -m gpu -q -x -s > gpurun_out/${tag}_chain_tests.log 2>&1; rc=$?
tail -15 gpurun_out/${tag}_chain_tests.log | cut -c1-300
if [ $rc -ne 0 ]; then echo "chain tests rc=$rc"; exit 0; fi
timeout 600 python -m pytest tests/test_model_gpu.py tests/test_hd128_gpu.py tests/test_sampling_gpu.py -m gpu -q -x > gpurun_out/${tag}_decode_tests.log 2>&1
tail -8 gpurun_out/${tag}_decode_tests.log | cut -c1-300
DECODE_VARIANTS=chain,plain timeout 600 python tools/decode_variants.py 60 > gpurun_out/${tag}_decode_variants.txt 2>&1; tail -12 gpurun_out/${tag}_decode_variants.txt | cut -c1-400
