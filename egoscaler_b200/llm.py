"""LLaMA trunk on the CUDA ops: forward (training / prefill / cached decode) and the hand-written backward.

Reference arithmetic: HF ``LlamaDecoderLayer`` x N (transformers 5.5.0 modeling_llama.py:292-340: RMSNorm :53-67,
attention :253-290 with RoPE :138-168, MLP :171-184) called from pointllm/model/pointllm.py:173-178, final
``LlamaModel.norm`` and ``lm_head`` (pointllm.py:228).

Per layer (T = B*L tokens, d hidden, f MLP width):
    xn  = rmsnorm(x)                                  [T,d]
    qkv = xn . Wqkv^T                                 [T,3d]   one tcgen05 GEMM for q|k|v
    rope(q,k) in place (+ KV-cache write)
    o   = attention(q,k,v, causal & key-padding)      [T,d]
    h   = o . Wo^T + x                                (residual fused in the GEMM epilogue)
    hn  = rmsnorm(h)
    gu  = hn . Wgu^T                                  [T,2f]   one GEMM for gate|up
    a   = silu(gate) * up                             [T,f]
    out = a . Wdown^T + h
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch

from . import ops
from .config import TrajVLMConfig


def rope_tables(head_dim: int, theta: float, max_pos: int, device) -> "tuple[torch.Tensor, torch.Tensor]":
    """cos/sin [max_pos, head_dim/2] bf16, computed exactly like LlamaRotaryEmbedding (modeling_llama.py:117-135):
    fp32 inv_freq and outer product on the CPU, cast to the activation dtype.  Setup-time table, not hot path."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32) / head_dim))
    fr = torch.arange(max_pos, dtype=torch.float32)[:, None] * inv[None, :]
    return fr.cos().to(torch.bfloat16).to(device).contiguous(), fr.sin().to(torch.bfloat16).to(device).contiguous()


@dataclass
class LayerWeights:
    wqkv: torch.Tensor
    wo: torch.Tensor
    wgu: torch.Tensor
    wdown: torch.Tensor
    ln1: torch.Tensor
    ln2: torch.Tensor
    # contraction-major copies for the data-gradient GEMMs (built on first backward)
    wqkvT: Optional[torch.Tensor] = None
    woT: Optional[torch.Tensor] = None
    wguT: Optional[torch.Tensor] = None
    wdownT: Optional[torch.Tensor] = None

    # decode-layout copies (128x64 tiles, pre-swizzled; ops.retile_weight) streamed by the single-token step
    tqkv: Optional["ops.TiledWeight"] = None
    to: Optional["ops.TiledWeight"] = None
    tgu: Optional["ops.TiledWeight"] = None
    tdown: Optional["ops.TiledWeight"] = None
    tiled_folded: bool = False          # tqkv / tgu carry ln1 / ln2 folded into their columns (ops.fold_norm_enabled)
    tgu_il: Optional["ops.TiledWeight"] = None    # gate|up rows interleaved per tile + ln2 folded (ops.gemm_chain SwiGLU phase)

    def ensure_tiled(self, folded: bool = False):
        if self.tqkv is None or self.tgu is None or self.tiled_folded != folded:
            self.tqkv = ops.retile_weight(self.wqkv, self.ln1 if folded else None)
            self.tgu = ops.retile_weight(self.wgu, self.ln2 if folded else None)
            self.tiled_folded = folded
        if self.to is None:
            self.to = ops.retile_weight(self.wo)
            self.tdown = ops.retile_weight(self.wdown)

    def ensure_chain(self):
        """decode-layout copies of ``LlamaTrunk._decode_step_chain``: q|k|v with ln1 folded, gate|up interleaved with ln2 folded."""
        if self.tqkv is None or not self.tiled_folded:
            self.tqkv = ops.retile_weight(self.wqkv, self.ln1)
            self.tgu = None                # the plain-layout gate|up copy is not needed by the chain (405 MB per layer)
            self.tiled_folded = True
        if self.tgu_il is None:
            self.tgu_il = ops.retile_weight_gateup(self.wgu, self.ln2)
        if self.to is None:
            self.to = ops.retile_weight(self.wo)
            self.tdown = ops.retile_weight(self.wdown)

    def refresh_derived(self):
        """After an optimizer step on the trunk weights (--unfreeze_language_model): rewrite the contraction-major
        copies in place and drop the decode-layout copies."""
        if self.wqkvT is not None:
            ops.transpose(self.wqkv, out=self.wqkvT)
            ops.transpose(self.wo, out=self.woT)
            ops.transpose(self.wgu, out=self.wguT)
            ops.transpose(self.wdown, out=self.wdownT)
        self.tqkv = self.to = self.tgu = self.tdown = self.tgu_il = None

    def ensure_transposed(self):
        if self.wqkvT is None:
            self.wqkvT = ops.transpose(self.wqkv)
            self.woT = ops.transpose(self.wo)
            self.wguT = ops.transpose(self.wgu)
            self.wdownT = ops.transpose(self.wdown)


@dataclass
class KVCache:
    """Static cache [layer] -> K, V [B, max_len, d] bf16 (keys post-RoPE like DynamicCache)."""
    k: List[torch.Tensor]
    v: List[torch.Tensor]
    max_len: int
    length: int = 0

    @staticmethod
    def allocate(cfg: TrajVLMConfig, batch: int, max_len: int, device) -> "KVCache":
        d = cfg.hidden_size
        k = [torch.empty(batch, max_len, d, device=device, dtype=torch.bfloat16) for _ in range(cfg.num_hidden_layers)]
        v = [torch.empty(batch, max_len, d, device=device, dtype=torch.bfloat16) for _ in range(cfg.num_hidden_layers)]
        return KVCache(k, v, max_len, 0)


@dataclass
class LayerSaved:
    x: torch.Tensor
    rstd1: torch.Tensor
    qkv: torch.Tensor
    o: torch.Tensor
    lse: torch.Tensor
    h: torch.Tensor
    rstd2: torch.Tensor
    gu: torch.Tensor
    rows: Optional[tuple] = None       # (r0, r1): o / lse / h / rstd2 / gu cover only these rows of every sample (top layer)


class LlamaTrunk:
    def __init__(self, cfg: TrajVLMConfig, layers: List[LayerWeights], device):
        self.cfg = cfg
        self.layers = layers
        self.cos, self.sin = rope_tables(cfg.head_dim, cfg.rope_theta, cfg.max_position_embeddings, device)
        self.decode_stat = None      # (fp32 statistics, parts) of the last decode step's output rows (_decode_step_folded)

    # ------------------------------------------------------------------ forward
    def forward(self, x: torch.Tensor, B: int, L: int, kmask: Optional[torch.Tensor], pos0: int = 0,
                cache: Optional[KVCache] = None, saved: Optional[List[LayerSaved]] = None,
                pos_dev: Optional[torch.Tensor] = None, out_rows: Optional[tuple] = None) -> torch.Tensor:
        """x [B*L, d] bf16 -> hidden [B*L, d] (before the final norm).  kmask uint8 [B, >= pos0+L] or None.
        With ``cache``: K/V of these L tokens are written at pos0.. and attention runs over cache[:pos0+L].
        ``out_rows = (r0, r1)``: the caller reads only rows r0..r1-1 of every sample (the loss rows of train.py:174) - the TOP
        layer then computes its attention output, o-proj and MLP for those rows only (its K / V still cover all positions) and
        the result is [B*(r1-r0), d].  Same arithmetic per row; ~1.6 % of a full-size training step."""
        c = self.cfg
        d, H, hd, f = c.hidden_size, c.num_attention_heads, c.head_dim, c.intermediate_size
        attn_decode = cache is not None and L == 1 and (pos0 > 0 or pos_dev is not None)
        decode = attn_decode and B <= ops.SKINNY_MAX_M      # weight-streaming GEMMs on the tiled copies
        if pos_dev is not None:          # graph-replayed decode step: the position lives in a device counter
            assert attn_decode and pos0 == 0
        self.decode_stat = None
        if decode and ops.decode_chain_enabled and saved is None:
            return self._decode_step_chain(x, B, kmask, pos0, cache, pos_dev)
        if decode and ops.fold_norm_enabled and saved is None:
            return self._decode_step_folded(x, B, kmask, pos0, cache, pos_dev)
        if decode:
            for w in self.layers:          # decode-layout copies exist before the first kernel of the step is queued, so no
                w.ensure_tiled(False)      # skinny GEMM is ever launched right behind the kernel that wrote its weights
        sub_ok = (out_rows is not None and cache is None and not decode and hd == 128 and ops.top_layer_rows_enabled
                  and 0 <= out_rows[0] < out_rows[1] <= L and (out_rows[1] - out_rows[0]) < L)
        for li, w in enumerate(self.layers):
            if sub_ok and li == len(self.layers) - 1:
                return self._top_layer_rows(x, w, B, L, kmask, pos0, saved, out_rows)
            if saved is not None:
                xn, rstd1 = ops.rmsnorm(x, w.ln1, c.rms_norm_eps, want_rstd=True)
            else:
                xn, rstd1 = ops.rmsnorm(x, w.ln1, c.rms_norm_eps), None
            roped = False
            if decode:
                qkv = ops.gemm_skinny(xn, w.tqkv, weights_const=True)
            elif ops.fuse_rope and cache is None and hd == 128 and x.shape[0] > ops.SKINNY_MAX_M:
                qkv = ops.gemm_rope(xn, w.wqkv, self.cos, self.sin, L, pos0, 2 * d, hd)     # projection + RoPE in one launch
                roped = True
            else:
                qkv = ops.gemm(xn, w.wqkv)                                        # [T, 3d]
            if roped:
                pass
            elif cache is not None:
                ops.rope_(qkv, L, 2 * H, hd, self.cos, self.sin, pos0, kcache=cache.k[li], vcache=cache.v[li],
                          kv_head0=H, kv_heads=H, v_col0=2 * d, cache_len=cache.max_len, pos_ptr=pos_dev)
            else:
                ops.rope_(qkv, L, 2 * H, hd, self.cos, self.sin, pos0)
            lse = None
            if attn_decode:
                o = ops.attention_decode(qkv, cache.k[li], cache.v[li], H, hd, kmask=kmask, len_ptr=pos_dev, len_static=pos0 + 1)
            else:
                q4 = qkv.view(B, L, 3, H, hd)
                if cache is not None:
                    k4 = cache.k[li][:, :pos0 + L].view(B, pos0 + L, H, hd)
                    v4 = cache.v[li][:, :pos0 + L].view(B, pos0 + L, H, hd)
                else:
                    k4, v4 = q4[:, :, 1], q4[:, :, 2]
                r = ops.attention(q4[:, :, 0], k4, v4, causal=True, kmask=kmask, q_pos0=pos0, want_lse=saved is not None)
                o, lse = r if saved is not None else (r, None)
                o = o.view(B * L, d)
            h = ops.gemm_skinny(o, w.to, residual=x, weights_const=True) if decode else ops.gemm(o, w.wo, residual=x)
            if saved is not None:
                hn, rstd2 = ops.rmsnorm(h, w.ln2, c.rms_norm_eps, want_rstd=True)
            else:
                hn, rstd2 = ops.rmsnorm(h, w.ln2, c.rms_norm_eps), None
            if not decode and ops.fuse_swiglu and x.shape[0] > ops.SKINNY_MAX_M:
                a, gu = ops.gemm_swiglu(hn, w.wgu, want_gu=saved is not None)     # gate|up projection + SwiGLU in one launch
            else:
                gu = ops.gemm_skinny(hn, w.tgu, weights_const=True) if decode else ops.gemm(hn, w.wgu)
                a = ops.silu_mul(gu)
            out = ops.gemm_skinny(a, w.tdown, residual=h, weights_const=True) if decode else ops.gemm(a, w.wdown, residual=h)
            if saved is not None:
                saved.append(LayerSaved(x, rstd1, qkv, o, lse, h, rstd2, gu))
            x = out
        return x

    def _top_layer_rows(self, x, w: LayerWeights, B: int, L: int, kmask, pos0: int, saved, rows) -> torch.Tensor:
        """The last decoder layer when only rows r0..r1-1 of every sample are read afterwards: queries, o-proj and MLP on
        those rows; keys / values (and therefore the q|k|v projection) on all L positions."""
        c = self.cfg
        d, H, hd = c.hidden_size, c.num_attention_heads, c.head_dim
        r0, r1 = rows
        Tn = r1 - r0
        keep = saved is not None
        xn, rstd1 = ops.rmsnorm(x, w.ln1, c.rms_norm_eps, want_rstd=True) if keep else (ops.rmsnorm(x, w.ln1, c.rms_norm_eps), None)
        if ops.fuse_rope:
            qkv = ops.gemm_rope(xn, w.wqkv, self.cos, self.sin, L, pos0, 2 * d, hd)
        else:
            qkv = ops.gemm(xn, w.wqkv)
            ops.rope_(qkv, L, 2 * H, hd, self.cos, self.sin, pos0)
        q4 = qkv.view(B, L, 3, H, hd)
        r = ops.attention(q4[:, r0:r1, 0], q4[:, :, 1], q4[:, :, 2], causal=True, kmask=kmask, q_pos0=pos0 + r0, want_lse=keep)
        o, lse = r if keep else (r, None)
        o = o.view(B * Tn, d)
        xs = torch.empty(B, Tn, d, device=x.device, dtype=torch.bfloat16)
        ops.copy_rows(x.view(B, L, d)[:, r0:r1, :], xs)
        h = ops.gemm(o, w.wo, residual=xs.view(B * Tn, d))
        hn, rstd2 = ops.rmsnorm(h, w.ln2, c.rms_norm_eps, want_rstd=True) if keep else (ops.rmsnorm(h, w.ln2, c.rms_norm_eps), None)
        if ops.fuse_swiglu and B * Tn > ops.SKINNY_MAX_M:
            a, gu = ops.gemm_swiglu(hn, w.wgu, want_gu=keep)
        else:
            gu = ops.gemm(hn, w.wgu)
            a = ops.silu_mul(gu)
        out = ops.gemm(a, w.wdown, residual=h)
        if keep:
            saved.append(LayerSaved(x, rstd1, qkv, o, lse, h, rstd2, gu, rows=(r0, r1)))
        return out

    def _decode_step_folded(self, x: torch.Tensor, B: int, kmask, pos0: int, cache: KVCache, pos_dev) -> torch.Tensor:
        """The cached single-token step with every RMSNorm folded around the Linear that consumes it (7 kernels per layer
        instead of 9):  Linear(norm(x)) = rstd(x) * (x . (W * g)^T) - g lives in the decode-layout weight copy, rstd is applied
        in the GEMM epilogue from the sums of squares the PRODUCING GEMM (o-proj / down-proj, which already hold the finished
        row tile) left behind.  ``self.decode_stat`` hands the last layer's statistics to the final norm + lm_head."""
        c = self.cfg
        d, H, hd, eps = c.hidden_size, c.num_attention_heads, c.head_dim, c.rms_norm_eps
        for w in self.layers:
            w.ensure_tiled(True)
        nt = (d + 127) // 128
        _, rstd0 = ops.rmsnorm(x, self.layers[0].ln1, eps, want_rstd=True)      # statistics of the embedding rows
        stat = (rstd0, 0)
        ssq = torch.empty(2, nt, 64, device=x.device, dtype=torch.float32)
        for li, w in enumerate(self.layers):
            qkv = ops.gemm_skinny(x, w.tqkv, weights_const=True, row_scale=stat, eps=eps)
            ops.rope_(qkv, 1, 2 * H, hd, self.cos, self.sin, pos0, kcache=cache.k[li], vcache=cache.v[li],
                      kv_head0=H, kv_heads=H, v_col0=2 * d, cache_len=cache.max_len, pos_ptr=pos_dev)
            o = ops.attention_decode(qkv, cache.k[li], cache.v[li], H, hd, kmask=kmask, len_ptr=pos_dev, len_static=pos0 + 1)
            h = ops.gemm_skinny(o, w.to, residual=x, weights_const=True, ssq_out=ssq[0])
            gu = ops.gemm_skinny(h, w.tgu, weights_const=True, row_scale=(ssq[0], nt), eps=eps)
            a = ops.silu_mul(gu)
            x = ops.gemm_skinny(a, w.tdown, residual=h, weights_const=True, ssq_out=ssq[1])
            stat = (ssq[1], nt)
        self.decode_stat = stat
        return x

    def _decode_step_chain(self, x: torch.Tensor, B: int, kmask, pos0: int, cache: KVCache, pos_dev) -> torch.Tensor:
        """The cached single-token step with the Linears between two attention calls in ONE persistent launch
        (ops.gemm_chain, csrc/gemm_chain.cu): per layer  rope + cache write -> attention -> [o-proj + residual -> gate|up + SwiGLU
        -> down-proj + residual -> next layer's q|k|v], norms folded as in ``_decode_step_folded``.  3 kernels per layer."""
        c = self.cfg
        d, H, hd, eps, f = c.hidden_size, c.num_attention_heads, c.head_dim, c.rms_norm_eps, c.intermediate_size
        for w in self.layers:
            w.ensure_chain()
        nt = (d + 127) // 128
        dev = x.device
        _, rstd0 = ops.rmsnorm(x, self.layers[0].ln1, eps, want_rstd=True)      # statistics of the embedding rows
        qkv = ops.gemm_skinny(x, self.layers[0].tqkv, weights_const=True, row_scale=(rstd0, 0), eps=eps)
        ssq = torch.empty(2, nt, 64, device=dev, dtype=torch.float32)
        n_layers = len(self.layers)
        for li, w in enumerate(self.layers):
            ops.rope_(qkv, 1, 2 * H, hd, self.cos, self.sin, pos0, kcache=cache.k[li], vcache=cache.v[li],
                      kv_head0=H, kv_heads=H, v_col0=2 * d, cache_len=cache.max_len, pos_ptr=pos_dev)
            o = ops.attention_decode(qkv, cache.k[li], cache.v[li], H, hd, kmask=kmask, len_ptr=pos_dev, len_static=pos0 + 1)
            h = torch.empty(B, d, device=dev, dtype=torch.bfloat16)
            a = torch.empty(B, f, device=dev, dtype=torch.bfloat16)
            xn = torch.empty(B, d, device=dev, dtype=torch.bfloat16)
            phases = [ops.ChainPhase(o, w.to, h, residual=x, ssq_out=ssq[0]),
                      ops.ChainPhase(h, w.tgu_il, a, row_scale=(ssq[0], nt), eps=eps, swiglu=True),
                      ops.ChainPhase(a, w.tdown, xn, residual=h, ssq_out=ssq[1])]
            if li + 1 < n_layers:
                qkv = torch.empty(B, 3 * d, device=dev, dtype=torch.bfloat16)
                phases.append(ops.ChainPhase(xn, self.layers[li + 1].tqkv, qkv, row_scale=(ssq[1], nt), eps=eps))
            ops.gemm_chain(phases, weights_const=True)
            x = xn
        self.decode_stat = (ssq[1], nt)
        return x

    # ------------------------------------------------------------------ backward (data gradients; frozen weights)
    def backward(self, dout: torch.Tensor, B: int, L: int, kmask: Optional[torch.Tensor], saved: List[LayerSaved],
                 wgrad: Optional[Dict[str, List[torch.Tensor]]] = None, on_layer_done=None, wgrad_accumulate: bool = True) -> torch.Tensor:
        """dout [B*L, d] bf16 -> dx [B*L, d].  Default recipe freezes the trunk (model_arch.py:33-51): only data
        gradients flow.  ``wgrad`` (fp32 accumulators per layer) enables the --unfreeze_language_model path; with
        ``wgrad_accumulate=False`` the four weight-gradient GEMMs of a layer OVERWRITE their accumulators (first micro-batch after
        ``zero_grad``: saves zero-filling and re-reading 27 GB of fp32 per step; the norm-weight gradients always accumulate)."""
        c = self.cfg
        d, H, hd, f = c.hidden_size, c.num_attention_heads, c.head_dim, c.intermediate_size
        # Trainable trunk (--unfreeze_language_model): the weights change every step, so the data gradients read them in their
        # [out, in] layout through the NN GEMM instead of maintaining 13.5 GB of contraction-major copies (and re-transposing
        # them after every optimizer step); the weight gradients dY^T . X read the token-major activations through the TN GEMM.
        nn_dgrad = wgrad is not None and ops.gemm_nn_enabled
        for li in reversed(range(len(self.layers))):
            w, s = self.layers[li], saved[li]
            if not nn_dgrad:
                w.ensure_transposed()
            rows = s.rows                     # top layer computed on a row range only: dout is [B*(r1-r0), d] until its attention
            Lq = L if rows is None else rows[1] - rows[0]
            if ops.fuse_dswiglu and dout.shape[0] > ops.SKINNY_MAX_M:
                # [T, 2f]: d(act) GEMM + dSwiGLU, one launch
                dgu = ops.gemm_dswiglu(dout, w.wdown, s.gu, w_is_down_proj=True) if nn_dgrad else ops.gemm_dswiglu(dout, w.wdownT, s.gu)
            else:
                da = ops.gemm_nn(dout, w.wdown) if nn_dgrad else ops.gemm(dout, w.wdownT)     # [T, f]
                dgu = ops.silu_mul_bwd(s.gu, da)                                  # [T, 2f]
            dhn = ops.gemm_nn(dgu, w.wgu) if nn_dgrad else ops.gemm(dgu, w.wguT)   # [T, d]
            if wgrad is not None:
                hn = ops.rmsnorm(s.h, w.ln2, c.rms_norm_eps)
                a = ops.silu_mul(s.gu)
                ops.gemm_tn(dout, a, out=wgrad["wdown"][li], accumulate=wgrad_accumulate)
                ops.gemm_tn(dgu, hn, out=wgrad["wgu"][li], accumulate=wgrad_accumulate)
                dh = ops.rmsnorm_bwd(dhn, s.h, w.ln2, s.rstd2, dres=dout, dw=wgrad["ln2"][li])
            else:
                dh = ops.rmsnorm_bwd(dhn, s.h, w.ln2, s.rstd2, dres=dout)
            do = ops.gemm_nn(dh, w.wo) if nn_dgrad else ops.gemm(dh, w.woT)       # [T, d]
            fused_bwd = ops.fuse_rope_bwd and hd == 128
            dqkv = ops.attention_bwd(s.qkv.view(B, L, 3, H, hd), s.o.view(B, Lq, d), s.lse, do.view(B, Lq, d), causal=True,
                                     kmask=kmask, rope=(self.cos, self.sin) if fused_bwd else None, q_rows=rows)      # [T, 3d]
            if not fused_bwd:
                ops.rope_(dqkv.view(B * L, 3 * d), L, 2 * H, hd, self.cos, self.sin, 0, inverse=True)
            dqkv2 = dqkv.view(B * L, 3 * d)
            dxn = ops.gemm_nn(dqkv2, w.wqkv) if nn_dgrad else ops.gemm(dqkv2, w.wqkvT)
            if wgrad is not None:
                xn = ops.rmsnorm(s.x, w.ln1, c.rms_norm_eps)
                ops.gemm_tn(dh, s.o, out=wgrad["wo"][li], accumulate=wgrad_accumulate)
                ops.gemm_tn(dqkv2, xn, out=wgrad["wqkv"][li], accumulate=wgrad_accumulate)
            if rows is not None:              # the residual gradient exists on the row range only: scatter it into a full tensor
                dres = torch.zeros(B, L, d, device=dh.device, dtype=torch.bfloat16)
                ops.copy_rows(dh.view(B, Lq, d), dres[:, rows[0]:rows[1], :])
                dres = dres.view(B * L, d)
            else:
                dres = dh
            if wgrad is not None:
                dout = ops.rmsnorm_bwd(dxn, s.x, w.ln1, s.rstd1, dres=dres, dw=wgrad["ln1"][li])
            else:
                dout = ops.rmsnorm_bwd(dxn, s.x, w.ln1, s.rstd1, dres=dres)
            saved[li] = None                                                      # free activations as we go
            if on_layer_done is not None:
                on_layer_done(li)                                                 # e.g. start this layer's gradient reduction
        return dout
