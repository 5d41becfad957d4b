#!/bin/bash
# chain kernel: L2 prefetch distance sweep, then the unit tests
tag=${1:-r02o}
mkdir -p gpurun_out
for d in ${SWEEP:-0 8 16 24 32 48 64}; do
  echo "== ESB_CHAIN_PREFETCH=$d" >> gpurun_out/${tag}_chain_sweep.txt
  ESB_CHAIN_PREFETCH=$d timeout 200 python tools/chain_bench.py >> gpurun_out/${tag}_chain_sweep.txt 2>&1
done
grep "==\|chain, one\|4 skinny" gpurun_out/${tag}_chain_sweep.txt
timeout 300 python -m pytest tests/test_gemm_chain_gpu.py -m gpu -q -x > gpurun_out/${tag}_chain_tests.log 2>&1; tail -3 gpurun_out/${tag}_chain_tests.log | cut -c1-300
