"""The training forward / backward of the trajectory VLM as reusable pieces, and the ``torch.autograd.Function`` that makes
the module a drop-in for the reference's own training loop (train.py:165-184):

    outputs = model(input_ids=tokens, attention_mask=attention_masks, point_clouds=pcrgbs, return_dict=True)
    logits = outputs.logits[:, prompts.shape[1] - 1:-1, :]
    loss = F.cross_entropy(logits.reshape(-1, logits.shape[-1]), tokens[:, prompts.shape[1]:].flatten(), ignore_index=pad)
    loss.backward()                     # model_engine.backward(loss)
    optimizer.step()                    # torch.optim.AdamW / DeepSpeed over model.parameters()

``forward_hidden_train`` / ``backward_from_hidden`` are shared with ``trainer.Trainer`` (the fast path: head on the T loss
rows only, fused cross-entropy, fused AdamW, overlapped gradient exchange).  ``TrajVLMFunction`` wraps the same hand-written
backward for arbitrary ``d loss / d logits``: the head runs on all B*L rows (as pointllm.py:228 does), the backward fills
``.grad`` of exactly the parameters the freezing recipe leaves trainable (model_arch.py:33-51).  Every kernel is reached
through the C ABI (ops.py); nothing here computes with ATen beyond dtype casts / copies of the gradient buffers.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional

import torch

from . import ops
from .encoder import projector_backward


@dataclass
class TrainCtx:
    """Activations of one training forward (what autograd would have saved)."""
    tokens: torch.Tensor
    kmask: Optional[torch.Tensor]
    B: int
    L: int
    splice_pos: Optional[torch.Tensor]
    saved: list = field(default_factory=list)           # per-layer llm.LayerSaved
    proj_saved: list = field(default_factory=list)      # projector (input, pre-activation) pairs
    enc_saved: Optional[dict] = None                    # train-mode point encoder activations (--unfreeze_pc_encoder)


def forward_hidden_train(model, tokens, attention_masks, pcrgbs, fps_start=None, drop_u=None, train_encoder: bool = True,
                         out_rows=None):
    """Point encoder -> projector -> splice -> N LLaMA layers with everything the backward needs kept.  Returns
    (hidden [B*L, d] bf16 before the final norm, TrainCtx).  pointllm.py:106-178.  ``out_rows = (r0, r1)``: the caller only reads
    these rows of every sample; the hidden states may then come back as [B*(r1-r0), d] (llm.LlamaTrunk._top_layer_rows) and the
    gradient handed to ``backward_from_hidden`` must have the same shape."""
    dev = model.dev
    tokens = tokens.to(dev)
    B, L = tokens.shape
    kmask = model._kmask(attention_masks, dev)
    ctx = TrainCtx(tokens=tokens, kmask=kmask, B=B, L=L, splice_pos=None)
    enc_train = model.enc_train() if train_encoder else None
    if enc_train is not None:
        enc_train["drop_u"] = drop_u                    # injected DropPath uniforms [depth, 2, B] (None: drawn on device)
    h, ctx.splice_pos = model.hidden_states(tokens, attention_masks, pcrgbs, fps_start, None, ctx.saved, ctx.proj_saved,
                                            enc_train=enc_train, out_rows=out_rows)
    if enc_train is not None:
        ctx.enc_saved = enc_train.pop("saved", None)
    return h, ctx


def backward_from_hidden(model, ctx: TrainCtx, dh: torch.Tensor, grad: Dict[str, torch.Tensor],
                         wgrad: Optional[Dict[str, List[torch.Tensor]]] = None, on_layer_done: Optional[Callable] = None,
                         on_ready: Optional[Callable[[List[str]], None]] = None, embed_exchange: Optional[Callable] = None,
                         wgrad_accumulate: bool = True):
    """dh [B*L, d] bf16 = d loss / d hidden (before the final norm).  Accumulates fp32 gradients into ``grad`` (by parameter
    name: embed_tokens, point_proj.*, point_backbone.* when it trains) and ``wgrad`` (packed LLM-layer units, only with
    --unfreeze_language_model).  ``on_ready(names)`` is called as soon as a group of gradients is final (the trainer starts
    its all-reduce there).  ``embed_exchange(tokens, dx, splice_pos, n_point)``: when given, the embed_tokens gradient is NOT
    scattered here - the callback owns it (the trainer's sparse exchange between data-parallel ranks) and the embedding is left
    out of ``on_ready``."""
    cfg = model.cfg
    B, L, d = ctx.B, ctx.L, cfg.hidden_size
    n_point = model.point_backbone_config["point_token_len"]
    dx = model.trunk.backward(dh, B, L, ctx.kmask, ctx.saved, wgrad=wgrad, on_layer_done=on_layer_done,
                              wgrad_accumulate=wgrad_accumulate)
    has_points = ctx.splice_pos is not None and len(ctx.proj_saved) > 0
    dfeats = torch.empty(B, n_point, d, device=dx.device, dtype=torch.bfloat16) if has_points else None
    ops.embed_splice_bwd(ctx.tokens, dx, ctx.splice_pos if has_points else None, n_point,
                         None if embed_exchange is not None else grad["model.embed_tokens.weight"], dfeats, cfg.pad_token_id,
                         vocab=cfg.vocab_size)
    if embed_exchange is not None:
        embed_exchange(ctx.tokens, dx, ctx.splice_pos if has_points else None, n_point)
    pw = [w for w, _ in model.projector_params()]
    pb = [b for _, b in model.projector_params()]
    if has_points:
        dtok = projector_backward(model._packed_proj, ctx.proj_saved, dfeats.view(B * n_point, d), [grad[n] for n in pw],
                                  [grad[n] for n in pb], need_input_grad=ctx.enc_saved is not None)
    if on_ready is not None:
        on_ready(([] if embed_exchange is not None else ["model.embed_tokens.weight"]) + pw + pb)
    if ctx.enc_saved is not None and has_points:
        from .encoder_train import train_backward
        train_backward(model.enc_train()["pk"], ctx.enc_saved, dtok, grad)
        ctx.enc_saved = None
        if on_ready is not None:
            on_ready([n for n in grad if n.startswith("model.point_backbone.")])


class TrajVLMFunction(torch.autograd.Function):
    """logits = TrajVLM(input_ids, mask, clouds; trainable parameters) with the hand-written CUDA backward."""

    @staticmethod
    def forward(ctx, model, input_ids, attention_mask, point_clouds, fps_start, drop_u, names, *params):
        cfg = model.cfg
        h, tctx = forward_hidden_train(model, input_ids, attention_mask, point_clouds, fps_start, drop_u,
                                       train_encoder=model.training)
        B, L, V = tctx.B, tctx.L, cfg.vocab_size
        norm_w = model.model.norm.weight.detach()
        hn, rstd = ops.rmsnorm(h, norm_w, cfg.rms_norm_eps, want_rstd=True)
        lbuf = torch.empty(B * L, model.v_pad, device=h.device, dtype=torch.bfloat16)
        ops.gemm(hn, model.lm_head.weight.detach(), out=lbuf[:, :V])
        ctx.model, ctx.tctx, ctx.names = model, tctx, names
        ctx.h, ctx.hn, ctx.rstd = h, hn, rstd
        return lbuf.view(B, L, model.v_pad)[:, :, :V]

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dlogits):
        model, tctx, names = ctx.model, ctx.tctx, ctx.names
        if tctx is None:
            raise RuntimeError("egoscaler_b200: the activations of this forward were already consumed by a backward pass")
        cfg = model.cfg
        B, L, V, d = tctx.B, tctx.L, cfg.vocab_size, cfg.hidden_size
        dev = ctx.h.device
        named = dict(model.named_parameters())
        grad = {n: torch.zeros(named[n].shape, device=dev, dtype=torch.float32) for n in names if not n.startswith("model.layers.")}
        wgrad = None
        if model.unfreeze_language_model:
            wgrad = {k: [torch.zeros(getattr(lw, k).shape, device=dev, dtype=torch.float32) for lw in model.trunk.layers]
                     for k in ("wqkv", "wo", "wgu", "wdown", "ln1", "ln2")}
        # ---- head on all B*L rows (pointllm.py:228): dW = dlogits^T . hn, dhn = dlogits . W
        dl = torch.zeros(B * L, model.v_pad, device=dev, dtype=torch.bfloat16)
        dl.view(B, L, model.v_pad)[:, :, :V].copy_(dlogits)
        dl2 = dl[:, :V]
        ops.gemm_tn(dl2, ctx.hn, out=grad["lm_head.weight"], accumulate=True)
        dhn = ops.gemm_nn(dl2, model.lm_head.weight.detach()) if ops.gemm_nn_enabled else ops.gemm(dl2, model.lm_head_T())
        dh = ops.rmsnorm_bwd(dhn, ctx.h, model.model.norm.weight.detach(), ctx.rstd, dw=grad["model.norm.weight"])
        del dl, dl2, dhn
        backward_from_hidden(model, tctx, dh, grad, wgrad)
        ctx.tctx = None
        out = []
        f = cfg.intermediate_size
        for n in names:
            p = named[n]
            if not n.startswith("model.layers."):
                g = grad[n]
            else:                                           # slices of the packed q|k|v and gate|up units
                li = int(n.split(".")[2])
                leaf = n.split(".", 3)[3]
                g = {"self_attn.q_proj.weight": lambda: wgrad["wqkv"][li][:d], "self_attn.k_proj.weight": lambda: wgrad["wqkv"][li][d:2 * d],
                     "self_attn.v_proj.weight": lambda: wgrad["wqkv"][li][2 * d:], "self_attn.o_proj.weight": lambda: wgrad["wo"][li],
                     "mlp.gate_proj.weight": lambda: wgrad["wgu"][li][:f], "mlp.up_proj.weight": lambda: wgrad["wgu"][li][f:],
                     "mlp.down_proj.weight": lambda: wgrad["wdown"][li], "input_layernorm.weight": lambda: wgrad["ln1"][li],
                     "post_attention_layernorm.weight": lambda: wgrad["ln2"][li]}[leaf]()
            out.append(g.to(p.dtype).view(p.shape))
        return (None, None, None, None, None, None, None, *out)
