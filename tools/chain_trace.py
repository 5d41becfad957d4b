#!/usr/bin/env python
"""Where a chain launch (esb_gemm_chain_bf16, full-size decode layer, B=32) spends its time: %globaltimer stamps per CTA and
phase (esb_gemm_chain_debug_buffer), reduced to min / median / max over CTAs, in us from the first stamp of the launch."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import lib, ops  # noqa: E402

M, d, f, NL = 32, 4096, 11008, 3
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(0)


def rnd(*s, scale=1.0):
    return (torch.randn(*s, device=dev, generator=g) * scale).bfloat16()


ln = torch.ones(d, device=dev, dtype=torch.bfloat16)
L = []
for _ in range(NL):
    wgu = rnd(2 * f, d, scale=d ** -0.5)
    L.append(dict(to=ops.retile_weight(rnd(d, d, scale=d ** -0.5)), tgu_il=ops.retile_weight_gateup(wgu, ln),
                  td=ops.retile_weight(rnd(d, f, scale=f ** -0.5)), tq=ops.retile_weight(rnd(3 * d, d, scale=d ** -0.5), ln)))
    del wgu
o, x = rnd(M, d), rnd(M, d)
nt = d // 128
ssq = torch.zeros(2, nt, 64, device=dev)
h, a, xn, q = (torch.empty(M, d, device=dev, dtype=torch.bfloat16), torch.empty(M, f, device=dev, dtype=torch.bfloat16),
               torch.empty(M, d, device=dev, dtype=torch.bfloat16), torch.empty(M, 3 * d, device=dev, dtype=torch.bfloat16))


def phases(w):
    return [ops.ChainPhase(o, w["to"], h, residual=x, ssq_out=ssq[0]),
            ops.ChainPhase(h, w["tgu_il"], a, row_scale=(ssq[0], nt), eps=1e-6, swiglu=True),
            ops.ChainPhase(a, w["td"], xn, residual=h, ssq_out=ssq[1]),
            ops.ChainPhase(xn, w["tq"], q, row_scale=(ssq[1], nt), eps=1e-6)]


sms = torch.cuda.get_device_properties(dev).multi_processor_count
buf = torch.zeros(2 * sms, 4, 8, device=dev, dtype=torch.int64)
for w in L:
    ops.gemm_chain(phases(w))
torch.cuda.synchronize()
lib.lib().raw("esb_gemm_chain_debug_buffer")(buf.data_ptr())
for rep in range(2):
    for w in L:
        buf.zero_()
        ops.gemm_chain(phases(w))
        torch.cuda.synchronize()
t = buf.cpu().double()
lib.lib().raw("esb_gemm_chain_debug_buffer")(None)
used = t[:, 0, 0] > 0
t = t[used]
t0 = t[t > 0].min()
names = ["producer enters phase", "producer: input ready", "MMA: first block", "MMA: last commit", "epilogue: phase start",
         "epilogue: segments done", "epilogue: counter bumped"]
print(f"{int(used.sum())} CTAs; us since the first stamp (min / median / max over CTAs)")
for pi in range(4):
    print(f"-- phase {pi}")
    for ev, nm in enumerate(names):
        v = t[:, pi, ev]
        v = (v[v > 0] - t0) / 1e3
        if v.numel():
            print(f"   {nm:28s} {v.min():8.2f} {v.median():8.2f} {v.max():8.2f}")
