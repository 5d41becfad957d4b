#!/usr/bin/env python
"""One tcgen05 GEMM launch per shape (run under ncu --metrics dram__bytes_read.sum,... -k regex:gemm_bf16): which operand's
re-reads make up the DRAM traffic above A + B + C?  Shapes vary one dimension of the gate|up projection at a time."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops  # noqa: E402

SHAPES = [(5568, 22016, 4096), (128, 22016, 4096), (1024, 22016, 4096), (2816, 22016, 4096), (5568, 256, 4096), (5568, 2048, 4096),
          (5568, 22016, 1024), (5632, 22016, 4096), (5568, 4096, 4096)]
g = torch.Generator(device="cuda").manual_seed(0)
for M, N, K in SHAPES:
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = torch.randn(N, K, device="cuda", generator=g).bfloat16()
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    ops.gemm(a, w, out=out)
    torch.cuda.synchronize()
    print(M, N, K, "A+B MB", round((M * K + N * K) * 2 / 1e6, 1), "C MB", round(M * N * 2 / 1e6, 1), flush=True)
