#!/usr/bin/env python
"""Micro-benchmark of the attention kernels at the training shape (B=8, H=32, L=696, hd=128, causal + key mask)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops

def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n

B, H, L, hd = 8, 32, 696, 128
d = H * hd
g = torch.Generator(device="cuda").manual_seed(0)
qkv = (torch.randn(B, L, 3, H, hd, device="cuda", generator=g) * 0.5).to(torch.bfloat16)
kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8); kmask[:, 530:536] = 0
q, k, v = qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2]
o, lse = ops.attention(q, k, v, causal=True, kmask=kmask, want_lse=True)
# reference check (fp32 torch)
qf, kf, vf = (t.float().permute(0, 2, 1, 3) for t in (q, k, v))
s = qf @ kf.transpose(-1, -2) * hd ** -0.5
mask = torch.tril(torch.ones(L, L, device="cuda", dtype=torch.bool))[None, None] & kmask.bool()[:, None, None, :]
s = s.masked_fill(~mask, float("-inf"))
ref = (torch.softmax(s, -1).to(torch.bfloat16).float() @ vf).permute(0, 2, 1, 3).reshape(B, L, d)
err = (o.float() - ref).abs().max().item() / ref.abs().max().item()
lerr = (lse - torch.logsumexp(s, -1)).abs().max().item()
print(f"fwd max rel err {err:.3e}  lse abs err {lerr:.3e}")
flops_fwd = 4 * B * H * L * L * hd / 2
ms = timeit(lambda: ops.attention(q, k, v, causal=True, kmask=kmask, want_lse=True))
print(f"attn fwd: {ms*1e3:.1f} us  {flops_fwd/ms/1e9:.1f} TFLOP/s (causal half counted)")
do = torch.randn(B, L, d, device="cuda", generator=g).to(torch.bfloat16)
ms = timeit(lambda: ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask))
print(f"attn bwd: {ms*1e3:.1f} us  {2.5*flops_fwd/ms/1e9:.1f} TFLOP/s")
