#!/bin/bash
tag=${1:-r02q}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gemm_chain_gpu.py -m gpu -q -x > gpurun_out/${tag}_chain_tests.log 2>&1; rc=$?
tail -5 gpurun_out/${tag}_chain_tests.log | cut -c1-300
if [ $rc -ne 0 ]; then exit 0; fi
ESB_CHAIN_PREFETCH=0 timeout 200 python tools/chain_trace.py > gpurun_out/${tag}_chain_trace.txt 2>&1; cat gpurun_out/${tag}_chain_trace.txt
SWEEP="${SWEEP:-0 4 12}" tools/gpu_round12.sh ${tag} | grep -v passed
