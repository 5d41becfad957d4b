#!/usr/bin/env python
"""Micro-benchmark of the decode-step GEMMs (M = 32 tokens): general tcgen05 kernel vs the split-K weight-streaming
kernel, per LLaMA-7B weight shape, rotating over enough distinct weight copies that nothing stays in the 126 MB L2.
Prints us/launch and achieved HBM GB/s (weight bytes only).  Usage: python tools/skinny_bench.py [M]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops  # noqa: E402

M = int(sys.argv[1]) if len(sys.argv) > 1 else 32
SHAPES = [("qkv", 12288, 4096), ("o", 4096, 4096), ("gate_up", 22016, 4096), ("down", 4096, 11008), ("lm_head", 32262, 4096)]


def timeit(fn, n_w, reps=5):
    """CUDA-graph replay of n_w launches (one per weight copy): device time only, no host launch overhead - this is how
    the decode step runs (decode.py captures the token step)."""
    for i in range(n_w):
        fn(i)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for i in range(n_w):
            fn(i)
    g.replay()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / n_w)
    return best


def main():
    dev = torch.device("cuda:0")
    for name, N, K in SHAPES:
        n_w = max(4, int(600e6 // (N * K * 2)) + 1)
        ws = [(torch.randn(N, K, device=dev) / K ** 0.5).bfloat16() for _ in range(n_w)]
        x = torch.randn(M, K, device=dev).bfloat16()
        res = torch.randn(M, N, device=dev).bfloat16()
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        gb = N * K * 2 / 1e9
        ops.skinny_enabled = False
        t = timeit(lambda i: ops.gemm(x, ws[i], residual=res, out=out), n_w)
        ops.skinny_enabled = True
        line = f"{name:8s} N={N:6d} K={K:6d} floor@6.55TB/s={gb / 6548.2 * 1e6:6.1f}us  general={t:6.1f}us ({gb / t * 1e6:5.0f} GB/s)"
        res_s = []
        for s in (0, 1, 2, 3, 4, 6, 8):
            t = timeit(lambda i: ops.gemm_skinny(x, ws[i], residual=res, out=out, force_splits=s), n_w)
            res_s.append(f"S{s}={t:.1f}")
        print(line + "  skinny: " + " ".join(res_s), flush=True)
        wt = [ops.retile_weight(w) for w in ws]
        res_s = []
        for s in (0, 1, 2, 3, 4, 6, 8):
            t = timeit(lambda i: ops.gemm_skinny(x, wt[i], residual=res, out=out, force_splits=s), n_w)
            res_s.append(f"S{s}={t:.1f}")
        print(" " * 40 + "tiled W: " + " ".join(res_s), flush=True)
        del wt
        del ws


if __name__ == "__main__":
    main()
