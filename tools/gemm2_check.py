#!/usr/bin/env python
"""Correctness + timing of the CTA-pair GEMM (tcgen05.mma.cta_group::2, esb_set_gemm_2cta) against the single-CTA kernel.
Run as a SUBPROCESS with a timeout (tests/test_gemm_gpu.py): a protocol bug in a two-CTA mbarrier pipeline shows up as a hang.
Prints one JSON line: {"ok": bool, "checks": {...}, "us": {...}}."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops
from egoscaler_b200.llm import rope_tables


def rnd(*s, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*s, device="cuda", generator=g) * scale).bfloat16()


def timeit(fn, reps=10):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def both(fn):
    ops.set_gemm_2cta(False)
    a = fn()
    ops.set_gemm_2cta(True)
    try:
        b = fn()
        torch.cuda.synchronize()
    finally:
        ops.set_gemm_2cta(False)
    return a, b


def main():
    full = "--full" in sys.argv
    checks, us = {}, {}
    time_only = "--time" in sys.argv
    shapes = [] if time_only else [(1392, 768, 256), (1100, 600, 320), (2048, 1024, 1024)] + ([(5568, 4096, 4096), (5568, 12288, 4096), (5568, 4096, 11008)] if full else [])
    for (M, N, K) in shapes:
        a, w = rnd(M, K, seed=M + N), rnd(N, K, seed=K + 1, scale=K ** -0.5)
        res = rnd(M, N, seed=3)
        one, two = both(lambda: ops.gemm(a, w, residual=res))
        ref = a.float() @ w.float().t() + res.float()
        checks[f"plain {M}x{N}x{K} equal"] = bool(torch.equal(one, two))
        checks[f"plain {M}x{N}x{K} vs fp32"] = bool((two.float() - ref).abs().max() <= 1.2e-2 * ref.abs().max())
        o32, t32 = both(lambda: ops.gemm(a, w, out_dtype=torch.float32))
        checks[f"fp32-out {M}x{N}x{K} equal"] = bool(torch.equal(o32, t32))
    # fused epilogues
    for (M, f, K) in [] if time_only else [(1392, 688, 256)] + ([(5568, 11008, 4096)] if full else []):
        a, wgu = rnd(M, K, seed=11), rnd(2 * f, K, seed=12, scale=K ** -0.5)
        (a1, g1), (a2, g2) = both(lambda: ops.gemm_swiglu(a, wgu, True))
        checks[f"swiglu {M}x{f}x{K} equal"] = bool(torch.equal(a1, a2) and torch.equal(g1, g2))
        dout, wdT, gu = rnd(M, K, seed=13), rnd(f, K, seed=14, scale=K ** -0.5), rnd(M, 2 * f, seed=15)
        d1, d2 = both(lambda: ops.gemm_dswiglu(dout, wdT, gu))
        checks[f"dswiglu {M}x{f}x{K} equal"] = bool(torch.equal(d1, d2))
    cos, sin = rope_tables(128, 10000.0, 1024, "cuda")
    for (B, L, H, K) in [] if time_only else [(2, 696, 2, 256)] + ([(8, 696, 32, 4096)] if full else []):
        a, w = rnd(B * L, K, seed=16), rnd(3 * H * 128, K, seed=17, scale=K ** -0.5)
        r1, r2 = both(lambda: ops.gemm_rope(a, w, cos, sin, L, 0, 2 * H * 128, 128))
        checks[f"rope {B * L}x{3 * H * 128}x{K} equal"] = bool(torch.equal(r1, r2))
    if full or time_only:                                          # timing on the benchmark's launch shapes
        for name, (M, N, K) in {"qkv": (5568, 12288, 4096), "o": (5568, 4096, 4096), "gate_up": (5568, 22016, 4096), "down": (5568, 4096, 11008)}.items():
            a, w = rnd(M, K, seed=1), rnd(N, K, seed=2, scale=K ** -0.5)
            out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
            for on in (False, True):
                ops.set_gemm_2cta(on)
                t = timeit(lambda: ops.gemm(a, w, out=out))
                us[f"{name} {'2cta' if on else '1cta'}"] = round(t, 1)
                us[f"{name} {'2cta' if on else '1cta'} TF/s"] = round(2.0 * M * N * K / t / 1e6, 1)
            ops.set_gemm_2cta(False)
    print(json.dumps({"ok": all(checks.values()), "checks": checks, "us": us}))


if __name__ == "__main__":
    main()
