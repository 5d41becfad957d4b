"""Autoregressive trajectory-token generation with a static KV cache.

Reference: ``TrajPointLLMForCausalLM.generate`` (model_arch.py:77-108) -> HF ``GenerationMixin`` with
``prepare_inputs_for_generation`` (pointllm.py:255-275): the first step runs the full prompt including the point
encoder, later steps feed one token against the cache; finished rows emit ``pad``; the attention mask grows by ones
(HF generation/utils.py:2762-2806).  Callers: train.py:221-231, evaluate.py:114-124.

Greedy decoding is the mode with a bit-exact contract (ids from identical logits).  ``do_sample=True`` (the reference's
default) runs HF's processor chain repetition-penalty -> temperature -> top-k -> top-p on device
(``esb_sample_rows``, csrc/sampling.cu): the surviving ids and their fp32 ``scores`` are HF's on identical logits, the
draw uses uniforms from a torch generator (``seed=`` / ``generator=``) and is distribution-equal to
``torch.multinomial``.  ``num_return_sequences`` repeats every prompt like HF ``_expand_inputs_for_generation``.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import ops
from .llm import KVCache
from .model import GenerateOutput, TrajPointLLMForCausalLM


@torch.no_grad()
def generate(model: TrajPointLLMForCausalLM, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None,
             point_clouds: Optional[torch.Tensor] = None, max_length: int = 20, temperature: float = 1.0, top_k: int = 50,
             top_p: float = 0.95, repetition_penalty: float = 1.0, do_sample: bool = True, num_return_sequences: int = 1,
             fps_start: Optional[torch.Tensor] = None, suppress_eos: bool = False, output_scores: bool = True,
             eos_check_every: int = 16, use_cuda_graph: bool = True, **kwargs) -> GenerateOutput:
    if do_sample and not (1 <= int(top_k) <= 1024):
        raise NotImplementedError("sampling decode supports top_k in 1..1024 (the reference uses 50)")
    cfg, dev = model.cfg, model.dev
    was_training = model.training
    model.eval()
    input_ids = input_ids.to(dev)
    n_ret = int(num_return_sequences)
    if n_ret > 1:                                         # HF _expand_inputs_for_generation: repeat_interleave on dim 0
        input_ids = input_ids.repeat_interleave(n_ret, dim=0)
        attention_mask = None if attention_mask is None else attention_mask.repeat_interleave(n_ret, dim=0)
        point_clouds = None if point_clouds is None else point_clouds.repeat_interleave(n_ret, dim=0)
        fps_start = None if fps_start is None else fps_start.repeat_interleave(n_ret, dim=0)
    B, P = input_ids.shape
    T = int(max_length)                                   # model_arch.py:96: max_new_tokens = max_length
    V, d = cfg.vocab_size, cfg.hidden_size
    cache = KVCache.allocate(cfg, B, P + T, dev)
    mask = torch.zeros(B, P + T, device=dev, dtype=torch.uint8)
    mask[:, :P] = 1 if attention_mask is None else attention_mask.to(dev).to(torch.uint8)
    seq = torch.full((B, P + T), cfg.pad_token_id, device=dev, dtype=torch.int64)
    seq[:, :P] = input_ids
    unfinished = torch.ones(B, device=dev, dtype=torch.int32)
    nxt = torch.empty(B, device=dev, dtype=torch.int64)
    tok = torch.empty(B, 1, device=dev, dtype=torch.int64)
    scores = torch.empty(T, B, V, device=dev, dtype=torch.float32) if output_scores else None
    last = torch.empty(B, 1, d, device=dev, dtype=torch.bfloat16)
    sup = cfg.eos_token_id if suppress_eos else -1
    u = None
    if do_sample:
        gen = kwargs.get("generator")
        if gen is None and kwargs.get("seed") is not None:
            gen = torch.Generator(device=dev).manual_seed(int(kwargs["seed"]))
        u = torch.rand(T, B, device=dev, dtype=torch.float32, generator=gen)
    pen = float(repetition_penalty)

    def pick(lg, step_off, step_ptr, sc, hist_len, hist_ptr):
        """next-token choice for logits ``lg``: greedy arg-max, or the HF sampling chain."""
        if not do_sample:
            ops.argmax_rows(lg, V, scores=sc, suppress_col=sup, out=nxt, step_ptr=step_ptr, step_off=step_off,
                            scores_step_stride=B * V)
        else:
            ops.sample_rows(lg, V, u, temperature, top_k, top_p, pen, hist=seq if pen != 1.0 else None, hist_len=hist_len,
                            pos_ptr=hist_ptr, scores=sc, suppress_col=sup, out=nxt, step_ptr=step_ptr, step_off=step_off,
                            scores_step_stride=B * V, u_step_stride=B)

    # Device-side cache-length counter: lets ONE captured CUDA graph of the single-token step (32 layers x 9 kernels +
    # head + argmax + bookkeeping, ~300 launches) be replayed for every generated token, so the step is bound by HBM
    # and not by launch latency.  Step 0 (prefill incl. the point encoder) and step 1 run eagerly.
    pos_dev = torch.full((1,), P, device=dev, dtype=torch.int32)

    def token_step():
        with ops.pdl_chain():
            hl_, _ = model.hidden_states(tok, mask, None, None, cache, pos_dev=pos_dev)
            lg = model.logits_from_hidden(hl_, decode=True)
        pick(lg, 1 - P, pos_dev, scores, 1, pos_dev)
        ops.greedy_update(nxt, tok.view(B), unfinished, seq, 1, mask, cfg.eos_token_id, cfg.pad_token_id, pos_ptr=pos_dev)

    graph = None
    steps_done = 0
    for s in range(T):
        if s == 0:
            h, _ = model.hidden_states(input_ids, mask[:, :P], point_clouds, fps_start, cache)
            ops.copy_rows(h.view(B, P, d)[:, P - 1:P, :], last)
            logits = model.logits_from_hidden(last.view(B, d))
            pick(logits, 0, None, scores, P, None)
            ops.greedy_update(nxt, tok.view(B), unfinished, seq, P, mask, cfg.eos_token_id, cfg.pad_token_id)
        elif s == 1 or not use_cuda_graph:
            token_step()
        else:
            if graph is None:
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    token_step()
            graph.replay()
        steps_done = s + 1
        if not suppress_eos and eos_check_every > 0 and (s + 1) % eos_check_every == 0:
            if int(unfinished.max().item()) == 0:        # the reference checks every step (one sync per token)
                break
    if not suppress_eos and steps_done == T:
        pass
    # HF stops as soon as every row has finished; trim the columns generated after that point
    if not suppress_eos:
        gen = seq[:, P:P + steps_done]
        fin = (gen == cfg.eos_token_id)
        if bool(fin.any(dim=1).all().item()):
            first = torch.where(fin.any(dim=1), fin.float().argmax(dim=1), torch.full((B,), steps_done - 1, device=dev))
            steps_done = int(first.max().item()) + 1
    model.train(was_training)
    out_scores = tuple(scores[i] for i in range(steps_done)) if scores is not None else None
    return GenerateOutput(sequences=seq[:, :P + steps_done], scores=out_scores)


TrajPointLLMForCausalLM.generate = lambda self, *a, **k: generate(self, *a, **k)
