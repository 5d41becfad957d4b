#!/usr/bin/env python
"""One GEMM shape, a few launches, for `ncu --set full` captures.  Usage: gemm_one.py M N K [2cta] [swiglu|dswiglu|rope]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops
from egoscaler_b200.llm import rope_tables

M, N, K = (int(x) for x in sys.argv[1:4])
flags = sys.argv[4:]
ops.set_gemm_2cta("2cta" in flags)
g = torch.Generator(device="cuda").manual_seed(0)
a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
for it in range(4):
    if "swiglu" in flags:
        w = (torch.randn(2 * N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
        out = ops.gemm_swiglu(a, w, True)
    elif "dswiglu" in flags:
        w = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
        gu = torch.randn(M, 2 * N, device="cuda", generator=g).bfloat16()
        out = ops.gemm_dswiglu(a, w, gu)
    elif "rope" in flags:
        cos, sin = rope_tables(128, 10000.0, 2048, "cuda")
        w = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
        out = ops.gemm_rope(a, w, cos, sin, 696, 0, N // 3 * 2, 128)
    else:
        w = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
        out = ops.gemm(a, w)
torch.cuda.synchronize()
print("ok")
