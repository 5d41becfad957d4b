#!/usr/bin/env python
"""Times the cached token step of generate() (B=32, P=544) in four variants: {CUDA graph, eager} x {skinny split-K
GEMMs on tiled weights, general tcgen05 GEMM}.  ms/token = (t(T) - t(1)) / (T - 1)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from egoscaler_b200 import ops, synth  # noqa: E402
from egoscaler_b200.config import TrajVLMConfig  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 60
dev = torch.device("cuda", 0)
cfg = TrajVLMConfig.full()
model = bench.build_model(torch, cfg, dev)
bt = synth.make_batch(cfg, 32, seed=77)
model.eval()
args = dict(input_ids=bt["prompts"].to(dev), attention_mask=bt["prompt_masks"].to(dev), point_clouds=bt["pcrgbs"].to(dev),
            do_sample=False, suppress_eos=True, fps_start=torch.zeros(32, dtype=torch.long, device=dev))


def run(t_new, graph):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    model.generate(max_length=t_new, use_cuda_graph=graph, **args)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


seqs = {}
for skinny, pdl, chain in ((True, True, True), (True, True, False), (True, True, True), (True, True, False), (True, False, False),
                           (False, False, False)):
    ops.SKINNY_MAX_M = 64 if skinny else 0
    ops.skinny_enabled = skinny
    ops.pdl_enabled = pdl
    ops.pdl_chain_enabled = chain
    print("pdl", pdl, "chain", chain, end=" ")
    for graph in (True,):
        run(4, graph)
        t1 = min(run(1, graph) for _ in range(2))
        tT = min(run(T, graph) for _ in range(2))
        print(f"skinny={skinny} graph={graph}: prefill+1 {t1:.1f} ms, {T} tokens {tT:.1f} ms -> {(tT - t1) / (T - 1):.3f} ms/token", flush=True)
    out = model.generate(max_length=12, use_cuda_graph=True, **args)
    seqs[(skinny, pdl, chain)] = out.sequences.cpu()
ref = seqs[(True, False, False)]
for k, v in seqs.items():
    print(k, "ids equal to the no-PDL run:", bool(torch.equal(v, ref)))
