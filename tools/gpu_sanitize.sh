#!/bin/bash
# compute-sanitizer on the tiny smoke() run (forward, training step, greedy decode, tokenizer kernels): ONE tool per gpurun call.
# usage: tools/gpu_sanitize.sh memcheck|racecheck
tool=${1:-memcheck}
mkdir -p gpurun_out
export ESB_SMOKE_FULL=0
python __graft_entry__.py smoke > gpurun_out/r02_sanitize_plain.log 2>&1 && \
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python __graft_entry__.py smoke > gpurun_out/r02_sanitizer_${tool}.log 2>&1
echo "rc=$?"; tail -12 gpurun_out/r02_sanitizer_${tool}.log | cut -c1-300
