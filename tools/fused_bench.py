#!/usr/bin/env python
"""Micro-benchmark of the fused GEMM epilogues / TN / NN forms against the launches they replace, at the full-size shapes of one
LLaMA-7B layer at B=8 (T = 5568 tokens).  CUDA-event timing, rotating operand sets larger than L2.  Prints one line per kernel."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops
from egoscaler_b200.llm import rope_tables

T, d, f, H, L, B = 5568, 4096, 11008, 32, 696, 8
NS = 3


def rnd(*s, scale=1.0):
    return (torch.randn(*s, device="cuda") * scale).bfloat16()


def timeit(fn, reps=12):
    for i in range(3):
        fn(i % NS)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i % NS)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps * 1e3


def main():
    out = {}
    x = [rnd(T, d) for _ in range(NS)]
    wgu = [rnd(2 * f, d, scale=d ** -0.5) for _ in range(NS)]
    out["gate_up gemm + silu_mul (us)"] = timeit(lambda i: ops.silu_mul(ops.gemm(x[i], wgu[i])))
    out["gate_up gemm alone (us)"] = timeit(lambda i: ops.gemm(x[i], wgu[i]))
    out["gemm_swiglu, gu kept (us)"] = timeit(lambda i: ops.gemm_swiglu(x[i], wgu[i], True))
    out["gemm_swiglu, act only (us)"] = timeit(lambda i: ops.gemm_swiglu(x[i], wgu[i], False))
    wdT = [rnd(f, d, scale=d ** -0.5) for _ in range(NS)]
    wd = [w.t().contiguous() for w in wdT]
    gu = [rnd(T, 2 * f) for _ in range(NS)]
    out["d(act) gemm + silu_mul_bwd (us)"] = timeit(lambda i: ops.silu_mul_bwd(gu[i], ops.gemm(x[i], wdT[i])))
    out["d(act) gemm alone (us)"] = timeit(lambda i: ops.gemm(x[i], wdT[i]))
    out["gemm_dswiglu (us)"] = timeit(lambda i: ops.gemm_dswiglu(x[i], wdT[i], gu[i]))
    out["gemm_dswiglu on down_proj [d,f] (NN) (us)"] = timeit(lambda i: ops.gemm_dswiglu(x[i], wd[i], gu[i], w_is_down_proj=True))
    del wdT, wd, gu
    cos, sin = rope_tables(128, 10000.0, 2048, "cuda")
    wqkv = [rnd(3 * d, d, scale=d ** -0.5) for _ in range(NS)]
    out["qkv gemm + rope (us)"] = timeit(lambda i: ops.rope_(ops.gemm(x[i], wqkv[i]), L, 2 * H, 128, cos, sin, 0))
    out["qkv gemm alone (us)"] = timeit(lambda i: ops.gemm(x[i], wqkv[i]))
    out["gemm_rope (us)"] = timeit(lambda i: ops.gemm_rope(x[i], wqkv[i], cos, sin, L, 0, 2 * d, 128))
    # attention backward +- fused inverse rope
    qkv = rnd(B, L, 3, H, 128)
    do = rnd(B, L, d)
    km = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    o, lse = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True, kmask=km, want_lse=True)
    out["attention fwd (us)"] = timeit(lambda i: ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True, kmask=km, want_lse=True))
    out["attention bwd + inverse rope (us)"] = timeit(lambda i: ops.rope_(ops.attention_bwd(qkv, o, lse, do, True, km).view(B * L, 3 * d), L, 2 * H, 128, cos, sin, 0, inverse=True))
    out["attention bwd, fused inverse rope (us)"] = timeit(lambda i: ops.attention_bwd(qkv, o, lse, do, True, km, rope=(cos, sin)))
    # weight gradient: dW[3d, d] = dqkv^T . xn
    dy = [rnd(T, 3 * d) for _ in range(NS)]
    acc = torch.zeros(3 * d, d, device="cuda")
    out["wgrad qkv: 2 transposes + NT gemm (us)"] = timeit(lambda i: ops.gemm(ops.transpose(dy[i]), ops.transpose(x[i]), out=acc, accumulate=True))
    out["wgrad qkv: TN gemm (us)"] = timeit(lambda i: ops.gemm_tn(dy[i], x[i], out=acc, accumulate=True))
    # data gradient: dxn[T, d] = dqkv . Wqkv
    wT = [w.t().contiguous() for w in wqkv]
    out["dgrad qkv: NT gemm on transposed copy (us)"] = timeit(lambda i: ops.gemm(dy[i], wT[i]))
    out["dgrad qkv: NN gemm on the parameter (us)"] = timeit(lambda i: ops.gemm_nn(dy[i], wqkv[i]))
    flops = {"gate_up": 2 * T * 2 * f * d, "qkv": 2 * T * 3 * d * d, "dact": 2 * T * f * d}
    for k, v in out.items():
        print(f"{k:55s} {v:9.1f}")
    print(json.dumps({"tflops_gate_up_alone": flops["gate_up"] / out["gate_up gemm alone (us)"] / 1e6,
                      "tflops_gemm_swiglu_kept": flops["gate_up"] / out["gemm_swiglu, gu kept (us)"] / 1e6,
                      "tflops_dswiglu": flops["dact"] / out["gemm_dswiglu (us)"] / 1e6,
                      "tflops_dact_alone": flops["dact"] / out["d(act) gemm alone (us)"] / 1e6,
                      "tflops_qkv_alone": flops["qkv"] / out["qkv gemm alone (us)"] / 1e6,
                      "tflops_gemm_rope": flops["qkv"] / out["gemm_rope (us)"] / 1e6}))


if __name__ == "__main__":
    main()
