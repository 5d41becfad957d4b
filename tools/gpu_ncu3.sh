#!/bin/bash
mkdir -p gpurun_out
A="python tools/attn_bench.py"
$A > gpurun_out/p_attn.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:attn_bwd_d -s 2 -c 2 -o gpurun_out/r02_attn_bwd $A > gpurun_out/n_attn.log 2>&1; echo "rc=$?"
ls -la gpurun_out/r02_attn_bwd.ncu-rep
