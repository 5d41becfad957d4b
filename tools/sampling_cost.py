#!/usr/bin/env python
"""ms per cached token step of generate(): greedy vs the reference's default sampling mode (top_k=50, top_p=0.95), B=32, P=544."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from egoscaler_b200 import synth  # noqa: E402
from egoscaler_b200.config import TrajVLMConfig  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 80
dev = torch.device("cuda", 0)
cfg = TrajVLMConfig.full()
model = bench.build_model(torch, cfg, dev)
bt = synth.make_batch(cfg, 32, seed=77)
model.eval()
args = dict(input_ids=bt["prompts"].to(dev), attention_mask=bt["prompt_masks"].to(dev), point_clouds=bt["pcrgbs"].to(dev),
            suppress_eos=True, fps_start=torch.zeros(32, dtype=torch.long, device=dev))


def run(t_new, **kw):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    model.generate(max_length=t_new, **args, **kw)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


for name, kw in (("greedy", dict(do_sample=False)), ("sample k50 p0.95", dict(do_sample=True, seed=1)),
                 ("sample + repetition_penalty 1.2", dict(do_sample=True, seed=1, repetition_penalty=1.2)),
                 ("greedy", dict(do_sample=False)), ("sample k50 p0.95", dict(do_sample=True, seed=1))):
    run(4, **kw)
    t1 = min(run(1, **kw) for _ in range(2))
    tT = min(run(T, **kw) for _ in range(2))
    print(f"{name}: {(tT - t1) / (T - 1):.3f} ms/token", flush=True)
