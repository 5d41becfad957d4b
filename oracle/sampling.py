"""ORACLE (test infrastructure only -- never imported by egoscaler_b200/): CPU restatement of the sampling step of the
reference's default decode mode.

Reference: ``TrajPointLLMForCausalLM.generate`` (model_arch.py:77-108) forwards ``do_sample=True, temperature=1.0,
top_k=50, top_p=0.95, repetition_penalty=1.0`` to HF ``GenerationMixin._sample``; the arithmetic lives in the
un-vendored dependency transformers (5.5.0 installed): generation/logits_process.py ``RepetitionPenaltyLogitsProcessor``
(:302), ``TemperatureLogitsWarper`` (:236), ``TopKLogitsWarper`` (:536), ``TopPLogitsWarper`` (:469).

Pinned: tests/test_oracle_golden.py::test_sampling_filter_matches_hf_processors runs the real HF classes (the
dependency is importable in this image) against :func:`hf_filter` on the same scores.
"""
from __future__ import annotations

import torch


def hf_filter(scores: torch.Tensor, input_ids: torch.Tensor | None = None, temperature: float = 1.0, top_k: int = 50,
              top_p: float = 0.95, penalty: float = 1.0) -> torch.Tensor:
    """scores [B,V] fp32 -> processed scores (``-inf`` where filtered), in the order HF applies the processors."""
    s = scores.clone().float()
    if penalty != 1.0 and input_ids is not None:                       # logits_process.py:302
        g = torch.gather(s, 1, input_ids)
        g = torch.where(g < 0, g * penalty, g / penalty)
        s = s.scatter(1, input_ids, g)
    if temperature != 1.0:                                             # :236
        s = s / temperature
    if top_k and top_k > 0:                                            # :536  ties with the k-th value survive
        k = min(top_k, s.shape[-1])
        kth = torch.topk(s, k)[0][..., -1, None]
        s = s.masked_fill(s < kth, float("-inf"))
    if top_p < 1.0:                                                    # :469  stable ascending order, largest always kept
        sl, si = torch.sort(s, descending=False, stable=True)
        cp = sl.softmax(dim=-1).cumsum(dim=-1)
        rm = cp <= (1 - top_p)
        rm[..., -1:] = False
        s = s.masked_fill(rm.scatter(1, si, rm), float("-inf"))
    return s


def inverse_cdf_draw(processed: torch.Tensor, u: torch.Tensor) -> torch.Tensor:
    """The draw the CUDA kernel makes: survivors in (score desc, id desc) order, first j with cumsum(exp(s-max)) > u*Z."""
    B, V = processed.shape
    out = torch.zeros(B, dtype=torch.int64)
    for b in range(B):
        idx = torch.nonzero(processed[b] > float("-inf")).flatten()
        if idx.numel() == 0:
            continue
        val = processed[b, idx]
        order = sorted(range(idx.numel()), key=lambda j: (-float(val[j]), -int(idx[j])))
        e = torch.exp(val[order] - val.max()).float()
        target = float(u[b]) * float(e.sum())
        c = torch.cumsum(e, 0)
        j = int(torch.nonzero(c > target).flatten()[0]) if bool((c > target).any()) else len(order) - 1
        out[b] = idx[order[j]]
    return out
