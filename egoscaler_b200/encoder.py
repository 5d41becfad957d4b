"""PointBERT point-cloud encoder + projector on the CUDA ops (forward; projector also backward).

Reference: pointbert/point_encoder.py:101-189 (PointTransformer.forward), dvae.py:150-221 (Group, Encoder),
misc.py:40-60 (fps), pointllm/model/pointllm.py:67-81,124-129 (point_proj).  The encoder is frozen and in eval mode
in the default training recipe (model_arch.py:33-51,110-124), so BatchNorm uses running statistics and is folded
into the preceding 1x1 conv when the operands are packed.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from . import ops
from .config import TrajVLMConfig

KPAD = 64          # K of the 6->128 conv and the 3->128 pos-embed linear is zero-padded to one TMA box


def _bf(t: torch.Tensor) -> torch.Tensor:
    return t.detach().to(torch.bfloat16).contiguous()


def _f32(t: torch.Tensor) -> torch.Tensor:
    return t.detach().float().contiguous()


def _pad_k(w: torch.Tensor, k: int) -> torch.Tensor:
    out = torch.zeros(w.shape[0], k, device=w.device, dtype=torch.bfloat16)
    out[:, :w.shape[1]] = w.to(torch.bfloat16)
    return out


class PackedPointEncoder:
    """bf16 GEMM operands / fp32 biases of ``model.point_backbone`` (setup-time packing, not on the hot path)."""

    def __init__(self, cfg: TrajVLMConfig, named: Dict[str, torch.Tensor]):
        p = cfg.point
        self.cfg = cfg
        g = lambda n: named["model.point_backbone." + n]

        def fold(conv, bn):
            w = _f32(g(conv + ".weight")).squeeze(-1)
            b = _f32(g(conv + ".bias"))
            s = _f32(g(bn + ".weight")) / torch.sqrt(_f32(g(bn + ".running_var")) + 1e-5)
            return w * s[:, None], (b - _f32(g(bn + ".running_mean"))) * s + _f32(g(bn + ".bias"))

        w1, b1 = fold("encoder.first_conv.0", "encoder.first_conv.1")
        self.c1_w, self.c1_b = _pad_k(w1, KPAD), b1.contiguous()
        self.c2_w, self.c2_b = _bf(g("encoder.first_conv.3.weight").squeeze(-1)), _f32(g("encoder.first_conv.3.bias"))
        w3, b3 = fold("encoder.second_conv.0", "encoder.second_conv.1")
        self.c3_w, self.c3_b = _bf(w3), b3.contiguous()
        self.c4_w, self.c4_b = _bf(g("encoder.second_conv.3.weight").squeeze(-1)), _f32(g("encoder.second_conv.3.bias"))
        self.red_w, self.red_b = _bf(g("reduce_dim.weight")), _f32(g("reduce_dim.bias"))
        self.pos1_w, self.pos1_b = _pad_k(g("pos_embed.0.weight"), KPAD), _f32(g("pos_embed.0.bias"))
        self.pos2_w, self.pos2_b = _bf(g("pos_embed.2.weight")), _f32(g("pos_embed.2.bias"))
        self.cls_token, self.cls_pos = _bf(g("cls_token")).view(1, 1, -1), _bf(g("cls_pos")).view(1, 1, -1)
        self.blocks = []
        for j in range(p.depth):
            b = f"blocks.blocks.{j}."
            self.blocks.append(dict(
                n1_g=_f32(g(b + "norm1.weight")), n1_b=_f32(g(b + "norm1.bias")),
                n2_g=_f32(g(b + "norm2.weight")), n2_b=_f32(g(b + "norm2.bias")),
                qkv_w=_bf(g(b + "attn.qkv.weight")), proj_w=_bf(g(b + "attn.proj.weight")), proj_b=_f32(g(b + "attn.proj.bias")),
                fc1_w=_bf(g(b + "mlp.fc1.weight")), fc1_b=_f32(g(b + "mlp.fc1.bias")),
                fc2_w=_bf(g(b + "mlp.fc2.weight")), fc2_b=_f32(g(b + "mlp.fc2.bias"))))
        self.norm_g, self.norm_b = _f32(g("norm.weight")), _f32(g("norm.bias"))


def point_encoder_forward(pk: PackedPointEncoder, pts: torch.Tensor, fps_start: torch.Tensor) -> torch.Tensor:
    """pts [B, N, C] fp32 cuda, fps_start [B] int64 cuda -> point tokens [B, Np, trans_dim] bf16."""
    cfg = pk.cfg.point
    B, N, C = pts.shape
    G, K, T, H = cfg.num_group, cfg.group_size, cfg.trans_dim, cfg.num_heads
    hd = T // H
    centers, _, cpad = ops.fps(pts, G, fps_start, pad_ld=KPAD)                       # misc.py:40-60
    nb = ops.knn_group(pts, centers, K, out_ld=KPAD)                                 # dvae.py:150-187
    # mini-PointNet (dvae.py:207-221)
    h = ops.gemm(nb, pk.c1_w, bias=pk.c1_b, act=ops.ACT_RELU)
    f = ops.gemm(h, pk.c2_w, bias=pk.c2_b)
    cat = ops.group_maxpool(f, K, concat=True)
    h = ops.gemm(cat, pk.c3_w, bias=pk.c3_b, act=ops.ACT_RELU)
    f = ops.gemm(h, pk.c4_w, bias=pk.c4_b)
    tok = ops.group_maxpool(f, K, concat=False)                                       # [B*G, enc]
    tok = ops.gemm(tok, pk.red_w, bias=pk.red_b)                                      # point_encoder.py:174
    Np = G + 1
    x = torch.empty(B, Np, T, device=pts.device, dtype=torch.bfloat16)
    pos = torch.empty(B, Np, T, device=pts.device, dtype=torch.bfloat16)
    ops.copy_rows(tok.view(B, G, T), x[:, 1:, :])                                     # :181 cat(cls, tokens)
    ops.copy_rows(pk.cls_token.expand(B, 1, T), x[:, :1, :])
    pe = ops.gemm(ops.gemm(cpad, pk.pos1_w, bias=pk.pos1_b, act=ops.ACT_GELU), pk.pos2_w, bias=pk.pos2_b)
    ops.copy_rows(pe.view(B, G, T), pos[:, 1:, :])                                    # :182 cat(cls_pos, pos)
    ops.copy_rows(pk.cls_pos.expand(B, 1, T), pos[:, :1, :])
    x2 = x.view(B * Np, T)
    pos2 = pos.view(B * Np, T)
    for blk in pk.blocks:                                                             # point_encoder.py:95-98, 58-76
        y = ops.layernorm(x2, blk["n1_g"], blk["n1_b"], add=pos2, sum_out=x2)         # x <- x + pos ; y = LN1(x)
        qkv = ops.gemm(y, blk["qkv_w"]).view(B, Np, 3, H, hd)
        o = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=False)
        x1 = ops.gemm(o.view(B * Np, T), blk["proj_w"], bias=blk["proj_b"], residual=x2)
        y = ops.layernorm(x1, blk["n2_g"], blk["n2_b"])
        hmid = ops.gemm(y, blk["fc1_w"], bias=blk["fc1_b"], act=ops.ACT_GELU)
        x2 = ops.gemm(hmid, blk["fc2_w"], bias=blk["fc2_b"], residual=x1)
    out = ops.layernorm(x2, pk.norm_g, pk.norm_b).view(B, Np, T)
    if not cfg.use_max_pool:
        return out
    # point_encoder.py:186-189: one token per cloud = cat(cls feature, max over the G group tokens)  [B, 1, 2T]
    rest = torch.empty(B, G, T, device=pts.device, dtype=torch.bfloat16)
    ops.copy_rows(out[:, 1:, :], rest)
    mx = ops.group_maxpool(rest.view(B * G, T), G, concat=False)
    res = torch.empty(B, 1, 2 * T, device=pts.device, dtype=torch.bfloat16)
    ops.copy_rows(out[:, :1, :], res[:, :, :T])
    ops.copy_rows(mx.view(B, 1, T), res[:, :, T:])
    return res


class PackedProjector:
    """``model.point_proj`` operands: bf16 weights (the parameters themselves), fp32 biases, lazily transposed copies."""

    def __init__(self, weights: List[torch.Tensor], biases: List[torch.Tensor]):
        self.w = weights
        self.b_param = biases
        self.refresh()

    def refresh(self):
        self.b = [_f32(b) for b in self.b_param]
        self.wT: Optional[List[torch.Tensor]] = None

    def transposed(self) -> List[torch.Tensor]:
        if self.wT is None:
            self.wT = [ops.transpose(w.detach()) for w in self.w]
        return self.wT


def projector_forward(pp: PackedProjector, feats: torch.Tensor, save: Optional[list] = None) -> torch.Tensor:
    """feats [R, trans] bf16 -> [R, d].  With ``save`` the pre-activations are kept for the backward."""
    x = feats
    n = len(pp.w)
    for i in range(n):
        last = i == n - 1
        if save is None:
            x = ops.gemm(x, pp.w[i].detach(), bias=pp.b[i], act=ops.ACT_NONE if last else ops.ACT_GELU)
        else:
            z = ops.gemm(x, pp.w[i].detach(), bias=pp.b[i])
            save.append((x, z))
            x = z if last else ops.gelu(z)
    return x


def projector_backward(pp: PackedProjector, saved: list, dout: torch.Tensor, grads_w: List[torch.Tensor],
                       grads_b: List[torch.Tensor], need_input_grad: bool = False) -> Optional[torch.Tensor]:
    """dout [R, d] bf16.  Accumulates fp32 dW / db into ``grads_w`` / ``grads_b``.  The gradient w.r.t. the encoder
    output is only computed (and returned) when the encoder trains (--unfreeze_pc_encoder)."""
    wT = None if ops.gemm_nn_enabled else pp.transposed()
    n = len(pp.w)
    dz = dout
    for i in reversed(range(n)):
        x, z = saved[i]
        if i != n - 1:
            dz = ops.gelu_bwd(z, dz)
        ops.colsum(dz, grads_b[i])
        # dW[N,K] = dz^T [N,R] . x [R,K]  ->  A = dz^T (rows N, contraction R), B = x^T (rows K, contraction R)
        ops.gemm_tn(dz, x, out=grads_w[i], accumulate=True)
        if i > 0 or need_input_grad:          # dx[R,K] = dz[R,N] . W[N,K]
            dz = ops.gemm_nn(dz, pp.w[i].detach()) if wT is None else ops.gemm(dz, wT[i])
    return dz if need_input_grad else None
