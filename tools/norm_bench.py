#!/usr/bin/env python
"""rmsnorm forward / backward at the trunk shape [5568, 4096]: register-resident kernels vs the general two-pass ones."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import ops  # noqa: E402

rows, d, NS = 5568, 4096, 8
xs = [torch.randn(rows, d, device="cuda").bfloat16() for _ in range(NS)]
dys = [torch.randn(rows, d, device="cuda").bfloat16() for _ in range(NS)]
w = torch.ones(d, device="cuda", dtype=torch.bfloat16)
_, rstd = ops.rmsnorm(xs[0], w, 1e-6, want_rstd=True)


def timeit(fn, reps=40):
    for i in range(4):
        fn(i % NS)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(reps):
        fn(i % NS)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


for mode in ("1", "0", "1", "0"):
    os.environ["ESB_RMSNORM_GENERIC"] = mode
    f = timeit(lambda i: ops.rmsnorm(xs[i], w, 1e-6))
    b = timeit(lambda i: ops.rmsnorm_bwd(dys[i], xs[i], w, rstd))
    br = timeit(lambda i: ops.rmsnorm_bwd(dys[i], xs[i], w, rstd, dres=xs[(i + 1) % NS]))
    print(f"{'general' if mode == '1' else 'register'}: fwd {f:.1f} us ({rows * d * 4 / f / 1e6:.2f} TB/s)  bwd {b:.1f} us ({rows * d * 6 / b / 1e6:.2f} TB/s)"
          f"  bwd+residual {br:.1f} us ({rows * d * 8 / br / 1e6:.2f} TB/s)")
