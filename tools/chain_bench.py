#!/usr/bin/env python
"""Times the decode-step Linears of ONE layer at full size (B=32): four esb_gemm_skinny launches (+ silu_mul) against the
persistent chain (esb_gemm_chain_bf16), and every Linear alone through both kernels.  Weights rotate over NL layer sets so that
every launch streams from HBM (NL x 404 MB >> L2)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops  # noqa: E402

M, d, f, NL = 32, 4096, 11008, 6
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(0)


def rnd(*s, scale=1.0):
    return (torch.randn(*s, device=dev, generator=g) * scale).bfloat16()


L = []
for _ in range(NL):
    wo, wgu, wd, wq = rnd(d, d, scale=d ** -0.5), rnd(2 * f, d, scale=d ** -0.5), rnd(d, f, scale=f ** -0.5), rnd(3 * d, d, scale=d ** -0.5)
    ln = torch.ones(d, device=dev, dtype=torch.bfloat16)
    L.append(dict(to=ops.retile_weight(wo), tgu=ops.retile_weight(wgu, ln), tgu_il=ops.retile_weight_gateup(wgu, ln),
                  td=ops.retile_weight(wd), tq=ops.retile_weight(wq, ln)))
    del wo, wgu, wd, wq
o, x = rnd(M, d), rnd(M, d)
nt = d // 128
ssq = torch.zeros(2, nt, 64, device=dev)
h, a, xn, q = (torch.empty(M, d, device=dev, dtype=torch.bfloat16), torch.empty(M, f, device=dev, dtype=torch.bfloat16),
               torch.empty(M, d, device=dev, dtype=torch.bfloat16), torch.empty(M, 3 * d, device=dev, dtype=torch.bfloat16))
gu = torch.empty(M, 2 * f, device=dev, dtype=torch.bfloat16)
eps = 1e-6
MB = {"o": d * d * 2e-6, "gu": 2 * f * d * 2e-6, "down": d * f * 2e-6, "qkv": 3 * d * d * 2e-6}


def separate(w):
    ops.gemm_skinny(o, w["to"], residual=x, ssq_out=ssq[0], out=h, weights_const=True)
    ops.gemm_skinny(h, w["tgu"], row_scale=(ssq[0], nt), eps=eps, out=gu, weights_const=True)
    ops._l.call("esb_silu_mul_fwd", gu.data_ptr(), a.data_ptr(), M, f, ops._stream())
    ops.gemm_skinny(a, w["td"], residual=h, ssq_out=ssq[1], out=xn, weights_const=True)
    ops.gemm_skinny(xn, w["tq"], row_scale=(ssq[1], nt), eps=eps, out=q, weights_const=True)


def phases(w):
    return [ops.ChainPhase(o, w["to"], h, residual=x, ssq_out=ssq[0]),
            ops.ChainPhase(h, w["tgu_il"], a, row_scale=(ssq[0], nt), eps=eps, swiglu=True),
            ops.ChainPhase(a, w["td"], xn, residual=h, ssq_out=ssq[1]),
            ops.ChainPhase(xn, w["tq"], q, row_scale=(ssq[1], nt), eps=eps)]


def chain(w):
    ops.gemm_chain(phases(w))


def timeit(fn, reps=5):
    """us per call, graph replay of NL calls (one per weight set), best of reps."""
    for w in L:
        fn(w)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr):
        for w in L:
            fn(w)
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        gr.replay()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e3 / NL)
    return best


tot = sum(MB.values())
for name, fn in (("4 skinny launches + silu_mul", separate), ("chain, one launch", chain)):
    t = timeit(fn)
    print(f"{name:32s} {t:8.1f} us/layer  {tot / t * 1e3:7.0f} GB/s of weights")
single = {"o": (lambda w: ops.gemm_skinny(o, w["to"], residual=x, ssq_out=ssq[0], out=h, weights_const=True), lambda w: ops.gemm_chain(phases(w)[0:1])),
          "gu": (lambda w: ops.gemm_skinny(h, w["tgu"], row_scale=(ssq[0], nt), eps=eps, out=gu, weights_const=True), lambda w: ops.gemm_chain(phases(w)[1:2])),
          "down": (lambda w: ops.gemm_skinny(a, w["td"], residual=h, ssq_out=ssq[1], out=xn, weights_const=True), lambda w: ops.gemm_chain(phases(w)[2:3])),
          "qkv": (lambda w: ops.gemm_skinny(xn, w["tq"], row_scale=(ssq[1], nt), eps=eps, out=q, weights_const=True), lambda w: ops.gemm_chain(phases(w)[3:4]))}
for k, (fs, fc) in single.items():
    ts, tc = timeit(fs), timeit(fc)
    print(f"{k:5s} {MB[k]:6.1f} MB   skinny {ts:7.1f} us {MB[k] / ts * 1e3:6.0f} GB/s   chain(1 phase) {tc:7.1f} us {MB[k] / tc * 1e3:6.0f} GB/s")
for n in (2, 3):
    t = timeit(lambda w: ops.gemm_chain(phases(w)[:n]))
    mb = sum(list(MB.values())[:n])
    print(f"chain, first {n} phases: {t:7.1f} us  {mb / t * 1e3:6.0f} GB/s")
