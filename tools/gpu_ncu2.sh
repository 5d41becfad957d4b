#!/bin/bash
# ncu visit 2: full-set captures of the dominant GEMM launches (gate|up with the SwiGLU epilogue = the largest launch; qkv 1-CTA vs CTA pair)
mkdir -p gpurun_out
for d in 0 1; do echo "== ESB_GEMM2_DBG=$d"; ESB_GEMM2_DBG=$d timeout 120 python tools/gemm2_check.py --time 2>&1 | tail -1 | tee gpurun_out/r02_gemm2_v6_dbg$d.json; done
timeout 600 python bench.py --steps 8 --warmup 3 --recipe unfreeze_lm --batch 4 --no-decode --no-cpu-baseline > gpurun_out/r02_bench_unfreeze_lm_n1.json 2>gpurun_out/r02_bench_unfreeze_lm_n1.err; cut -c1-200 gpurun_out/r02_bench_unfreeze_lm_n1.json
A="python tools/gemm_one.py 5568 11008 4096 swiglu"; B="python tools/gemm_one.py 5568 12288 4096"; C="python tools/gemm_one.py 5568 12288 4096 2cta"; D="python tools/gemm_one.py 5568 11008 4096 dswiglu"
NCU="ncu --set full --clock-control none --import-source on -k regex:gemm -s 2 -c 1"
$A > gpurun_out/p_a.log 2>&1 && $NCU -o gpurun_out/r02_gemm_swiglu $A > gpurun_out/n_a.log 2>&1; echo "a rc=$?"
$B > gpurun_out/p_b.log 2>&1 && $NCU -o gpurun_out/r02_gemm_qkv_1cta $B > gpurun_out/n_b.log 2>&1; echo "b rc=$?"
$C > gpurun_out/p_c.log 2>&1 && $NCU -o gpurun_out/r02_gemm_qkv_2cta $C > gpurun_out/n_c.log 2>&1; echo "c rc=$?"
ESB_FUSE_DSWIGLU=1 $D > gpurun_out/p_d.log 2>&1 && $NCU -o gpurun_out/r02_gemm_dswiglu $D > gpurun_out/n_d.log 2>&1; echo "d rc=$?"
ls -la gpurun_out/*.ncu-rep
