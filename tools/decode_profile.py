#!/usr/bin/env python
"""Profiling aid: full-size model, B=32 prompts of P=544, a few EAGER cached token steps (no CUDA graph) so that
`ncu --metrics gpu__time_duration.sum` lists the kernels of one decode step.  Usage: decode_profile.py [new_tokens]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from egoscaler_b200 import synth  # noqa: E402
from egoscaler_b200.config import TrajVLMConfig  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 4
dev = torch.device("cuda", 0)
cfg = TrajVLMConfig.full()
model = bench.build_model(torch, cfg, dev)
bt = synth.make_batch(cfg, 32, seed=77)
model.eval()
out = model.generate(input_ids=bt["prompts"].to(dev), attention_mask=bt["prompt_masks"].to(dev), point_clouds=bt["pcrgbs"].to(dev),
                     max_length=T, do_sample=False, suppress_eos=True, use_cuda_graph=False,
                     fps_start=torch.zeros(32, dtype=torch.long, device=dev))
torch.cuda.synchronize()
print("ok", tuple(out.sequences.shape))
