"""PointBERT point encoder in TRAIN mode on the CUDA ops: forward with saved activations + hand-written backward.

Only used with ``--unfreeze_pc_encoder`` (model_arch.py:36-38: the encoder's parameters get gradients; :121-122: it is
no longer forced into eval).  Train mode changes the arithmetic of the reference:
  * ``nn.BatchNorm1d`` in the mini-PointNet (pointbert/dvae.py:196,202) normalises with BATCH statistics over all
    B*G*K neighbourhood rows and updates its running statistics (momentum 0.1, unbiased variance);
  * every transformer block applies timm ``DropPath`` to both residual branches (point_encoder.py:62,74-75) with rates
    ``linspace(0, drop_path_rate, depth)`` (:134): branch * floor(keep + U) / keep, one draw per sample.
FPS / kNN carry no gradient (index selection; the cloud itself is an input).

Forward = same GEMMs as encoder.py but with the BatchNorm unfolded; backward = the chain rule written out, every
contraction on the tcgen05 GEMM (data gradient: dz . W via a transposed copy of W; weight gradient: dz^T . x with fp32
accumulation into the gradient buffer), everything else on the HBM-bound kernels of csrc/encoder_train.cu.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch

from . import ops
from .config import TrajVLMConfig
from .encoder import KPAD, _bf, _f32, _pad_k

PRE = "model.point_backbone."


class TrainPointEncoder:
    """bf16 operands of the trainable encoder, rebuilt from the fp32 parameters after every optimizer step."""

    def __init__(self, cfg: TrajVLMConfig, named: Dict[str, torch.Tensor]):
        self.cfg = cfg
        self.named = named
        self.refresh()

    def g(self, n: str) -> torch.Tensor:
        return self.named[PRE + n]

    def refresh(self):
        g = self.g
        p = self.cfg.point
        self.c1_w, self.c1_b = _pad_k(g("encoder.first_conv.0.weight").detach().squeeze(-1), KPAD), _f32(g("encoder.first_conv.0.bias"))
        self.c2_w, self.c2_b = _bf(g("encoder.first_conv.3.weight").squeeze(-1)), _f32(g("encoder.first_conv.3.bias"))
        self.c3_w, self.c3_b = _bf(g("encoder.second_conv.0.weight").squeeze(-1)), _f32(g("encoder.second_conv.0.bias"))
        self.c4_w, self.c4_b = _bf(g("encoder.second_conv.3.weight").squeeze(-1)), _f32(g("encoder.second_conv.3.bias"))
        self.bn = {}
        for key, name in (("bn1", "encoder.first_conv.1"), ("bn3", "encoder.second_conv.1")):
            self.bn[key] = dict(gamma=_f32(g(name + ".weight")), beta=_f32(g(name + ".bias")),
                                rm=g(name + ".running_mean"), rv=g(name + ".running_var"),
                                nbt=g(name + ".num_batches_tracked"), name=name)
        self.red_w, self.red_b = _bf(g("reduce_dim.weight")), _f32(g("reduce_dim.bias"))
        self.pos1_w, self.pos1_b = _pad_k(g("pos_embed.0.weight").detach(), KPAD), _f32(g("pos_embed.0.bias"))
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
        self.drop_rates: List[float] = [v.item() for v in torch.linspace(0, 0.1, p.depth)]      # yaml drop_path_rate 0.1


def _bn_train_fwd(z: torch.Tensor, bn: dict, save: dict, key: str) -> torch.Tensor:
    """relu(BatchNorm1d(z)) with batch statistics over the rows of z [R, C]; updates the running statistics in place."""
    R, C = z.shape
    f = lambda: torch.zeros(C, device=z.device, dtype=torch.float32)
    s0, s1, mean, rstd, A, Cc = f(), f(), f(), f(), f(), f()
    ops.colsum2(z, None, s0, s1)
    ops.bn_coeffs(0, s0, s1, R, bn["gamma"], bn["beta"], mean, rstd, A, None, Cc, running_mean=bn["rm"], running_var=bn["rv"])
    bn["nbt"].add_(1)                   # nn.BatchNorm1d bookkeeping (state_dict parity; momentum is fixed, the count is unused)
    y = ops.affine_cols(z, A, Cc, relu=True)
    save[key] = (z, y, mean, rstd)
    return y


def _bn_train_bwd(dy: torch.Tensor, bn: dict, saved, g_gamma: torch.Tensor, g_beta: torch.Tensor) -> torch.Tensor:
    z, y, mean, rstd = saved
    R, C = z.shape
    f = lambda: torch.zeros(C, device=z.device, dtype=torch.float32)
    t0, t1, A, B, Cc = f(), f(), f(), f(), f()
    dym = ops.relu_mask(dy, y)
    ops.colsum2(dym, z, t0, t1)
    ops.bn_coeffs(1, t0, t1, R, bn["gamma"], None, mean, rstd, A, B, Cc, dgamma=g_gamma, dbeta=g_beta)
    return ops.affine_cols(dym, A, Cc, y=z, B=B)


def drop_scales(pk: TrainPointEncoder, B: int, device, drop_u: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Per-sample DropPath factors [depth, 2, B] fp32: floor(keep + u) / keep (1 where the rate is 0).  ``drop_u`` injects
    the uniforms (tests: the reference's own draws); otherwise they come from torch's device generator."""
    depth = len(pk.blocks)
    if drop_u is None:
        drop_u = torch.rand(depth, 2, B, device=device, dtype=torch.float32)
    u = drop_u.to(device=device, dtype=torch.float32)
    keep = torch.tensor([1.0 - r for r in pk.drop_rates], device=device, dtype=torch.float32).view(depth, 1, 1)
    return ((keep + u).floor() / keep).contiguous()


def train_forward(pk: TrainPointEncoder, pts: torch.Tensor, fps_start: torch.Tensor, drop_u: Optional[torch.Tensor] = None):
    """pts [B, N, C] fp32, fps_start [B] int64 -> (point tokens [B, Np, T] bf16, saved activations for the backward)."""
    cfg = pk.cfg.point
    if cfg.use_max_pool:
        raise NotImplementedError("use_max_pool variant (point_encoder.py:186-189) is not built")
    B, N, C = pts.shape
    G, K, T, H = cfg.num_group, cfg.group_size, cfg.trans_dim, cfg.num_heads
    hd = T // H
    Np = G + 1
    sv: dict = dict(B=B)
    centers, _, cpad = ops.fps(pts, G, fps_start, pad_ld=KPAD)
    nb = ops.knn_group(pts, centers, K, out_ld=KPAD)
    sv["nb"], sv["cpad"] = nb, cpad
    # ---- mini-PointNet, BatchNorm on batch statistics
    z1 = ops.gemm(nb, pk.c1_w, bias=pk.c1_b)
    y1 = _bn_train_fwd(z1, pk.bn["bn1"], sv, "bn1")
    f1 = ops.gemm(y1, pk.c2_w, bias=pk.c2_b)
    cat = ops.group_maxpool(f1, K, concat=True)
    z3 = ops.gemm(cat, pk.c3_w, bias=pk.c3_b)
    y3 = _bn_train_fwd(z3, pk.bn["bn3"], sv, "bn3")
    f2 = ops.gemm(y3, pk.c4_w, bias=pk.c4_b)
    tok = ops.group_maxpool(f2, K, concat=False)
    sv["f1"], sv["cat"], sv["f2"], sv["tok"] = f1, cat, f2, tok
    t = ops.gemm(tok, pk.red_w, bias=pk.red_b)
    x = torch.empty(B, Np, T, device=pts.device, dtype=torch.bfloat16)
    pos = torch.empty(B, Np, T, device=pts.device, dtype=torch.bfloat16)
    ops.copy_rows(t.view(B, G, T), x[:, 1:, :])
    ops.copy_rows(pk.cls_token.expand(B, 1, T), x[:, :1, :])
    zpe = ops.gemm(cpad, pk.pos1_w, bias=pk.pos1_b)
    gpe = ops.gelu(zpe)
    pe = ops.gemm(gpe, pk.pos2_w, bias=pk.pos2_b)
    sv["zpe"], sv["gpe"] = zpe, gpe
    ops.copy_rows(pe.view(B, G, T), pos[:, 1:, :])
    ops.copy_rows(pk.cls_pos.expand(B, 1, T), pos[:, :1, :])
    x2, pos2 = x.view(B * Np, T), pos.view(B * Np, T)
    scales = drop_scales(pk, B, pts.device, drop_u)
    sv["scales"] = scales
    sv["blocks"] = []
    for j, blk in enumerate(pk.blocks):
        drop = pk.drop_rates[j] > 0.0
        xin = torch.empty_like(x2)
        y = ops.layernorm(x2, blk["n1_g"], blk["n1_b"], add=pos2, sum_out=xin)        # xin = x + pos ; y = LN1(xin)
        qkv = ops.gemm(y, blk["qkv_w"]).view(B, Np, 3, H, hd)
        o, lse = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=False, want_lse=True)
        o2 = o.view(B * Np, T)
        if drop:
            br = ops.gemm(o2, blk["proj_w"], bias=blk["proj_b"])
            x1 = ops.scale_rows_add(br, scales[j, 0], Np, x=xin)
        else:
            x1 = ops.gemm(o2, blk["proj_w"], bias=blk["proj_b"], residual=xin)
        y2 = ops.layernorm(x1, blk["n2_g"], blk["n2_b"])
        z = ops.gemm(y2, blk["fc1_w"], bias=blk["fc1_b"])
        hmid = ops.gelu(z)
        if drop:
            br2 = ops.gemm(hmid, blk["fc2_w"], bias=blk["fc2_b"])
            xo = ops.scale_rows_add(br2, scales[j, 1], Np, x=x1)
        else:
            xo = ops.gemm(hmid, blk["fc2_w"], bias=blk["fc2_b"], residual=x1)
        sv["blocks"].append(dict(xin=xin, y=y, qkv=qkv, o=o, lse=lse, x1=x1, y2=y2, z=z, hmid=hmid))
        x2 = xo
    out = ops.layernorm(x2, pk.norm_g, pk.norm_b)
    sv["x_final"] = x2
    return out.view(B, Np, T), sv


def _wgrad(dz: torch.Tensor, x: torch.Tensor, gw: torch.Tensor, gb: Optional[torch.Tensor]):
    """Linear / Conv1d(k=1) parameter gradients, fp32-accumulated: gW[N,K] += dz^T . x ; gb[N] += column sums of dz."""
    if gb is not None:
        ops.colsum(dz, gb)
    ops.gemm_tn(dz, x, out=gw, accumulate=True)


def _wgrad_padded(dz: torch.Tensor, xpad: torch.Tensor, gw: torch.Tensor, gb: torch.Tensor):
    """Same for the layers whose input was zero-padded to one TMA box (6 -> 64, 3 -> 64 columns): the first ``in`` columns
    of the padded gradient are added to the [out, in] parameter gradient."""
    out_f, in_f = gw.shape[0], gw.shape[1]
    tmp = torch.zeros(out_f, xpad.shape[1], device=dz.device, dtype=torch.float32)
    _wgrad(dz, xpad, tmp, gb)
    gw.add_(tmp[:, :in_f])


def train_backward(pk: TrainPointEncoder, sv: dict, dout: torch.Tensor, grads: Dict[str, torch.Tensor]) -> None:
    """dout [B*Np, T] bf16 = gradient w.r.t. the encoder output.  Accumulates fp32 gradients of EVERY encoder parameter
    into ``grads`` (keyed by the reference's parameter names, shapes of the parameters)."""
    cfg = pk.cfg.point
    B = sv["B"]
    G, K, T, H = cfg.num_group, cfg.group_size, cfg.trans_dim, cfg.num_heads
    hd = T // H
    Np = G + 1
    gr = lambda n: grads[PRE + n]
    gr2 = lambda n: grads[PRE + n].view(grads[PRE + n].shape[0], -1)          # conv [out, in, 1] -> [out, in]
    scales = sv["scales"]
    dx = ops.layernorm_bwd(dout, sv["x_final"], pk.norm_g, dgamma=gr("norm.weight"), dbeta=gr("norm.bias"))
    for j in reversed(range(len(pk.blocks))):
        blk, s = pk.blocks[j], sv["blocks"][j]
        b = f"blocks.blocks.{j}."
        drop = pk.drop_rates[j] > 0.0
        # ---- MLP branch: x = x1 + s_m * fc2(gelu(fc1(LN2(x1))))
        dbr2 = ops.scale_rows_add(dx, scales[j, 1], Np) if drop else dx
        _wgrad(dbr2, s["hmid"], gr(b + "mlp.fc2.weight"), gr(b + "mlp.fc2.bias"))
        dh = ops.gemm_nn(dbr2, blk["fc2_w"])
        dz = ops.gelu_bwd(s["z"], dh)
        _wgrad(dz, s["y2"], gr(b + "mlp.fc1.weight"), gr(b + "mlp.fc1.bias"))
        dy2 = ops.gemm_nn(dz, blk["fc1_w"])
        dx1 = ops.layernorm_bwd(dy2, s["x1"], blk["n2_g"], dgamma=gr(b + "norm2.weight"), dbeta=gr(b + "norm2.bias"), dres=dx)
        # ---- attention branch: x1 = xin + s_a * proj(attn(LN1(xin)))
        dbr = ops.scale_rows_add(dx1, scales[j, 0], Np) if drop else dx1
        o2 = s["o"].view(B * Np, T)
        _wgrad(dbr, o2, gr(b + "attn.proj.weight"), gr(b + "attn.proj.bias"))
        do = ops.gemm_nn(dbr, blk["proj_w"])
        dqkv = ops.attention_bwd(s["qkv"], s["o"], s["lse"], do.view(B, Np, T), causal=False).view(B * Np, 3 * T)
        _wgrad(dqkv, s["y"], gr(b + "attn.qkv.weight"), None)
        dy = ops.gemm_nn(dqkv, blk["qkv_w"])
        dxin = ops.layernorm_bwd(dy, s["xin"], blk["n1_g"], dgamma=gr(b + "norm1.weight"), dbeta=gr(b + "norm1.bias"), dres=dx1)
        # ---- xin = x + pos: the same gradient flows to the previous block and (summed over blocks) to pos.  The pos path is
        # linear in its output gradient, so each block's share is pushed through it right away (no fp32 staging buffer).
        d3 = dxin.view(B, Np, T)
        ops.colsum(d3[:, 0, :], gr("cls_pos").view(-1))                                    # rows b of stride Np*T
        dpe = torch.empty(B, G, T, device=dxin.device, dtype=torch.bfloat16)
        ops.copy_rows(d3[:, 1:, :], dpe)
        dpe2 = dpe.view(B * G, T)
        _wgrad(dpe2, sv["gpe"], gr("pos_embed.2.weight"), gr("pos_embed.2.bias"))
        dzpe = ops.gelu_bwd(sv["zpe"], ops.gemm_nn(dpe2, pk.pos2_w))
        _wgrad_padded(dzpe, sv["cpad"], gr("pos_embed.0.weight"), gr("pos_embed.0.bias"))
        dx = dxin
        sv["blocks"][j] = None
    # ---- x0 = cat(cls_token, reduce_dim(tokens))
    d3 = dx.view(B, Np, T)
    ops.colsum(d3[:, 0, :], gr("cls_token").view(-1))
    dt = torch.empty(B, G, T, device=dx.device, dtype=torch.bfloat16)
    ops.copy_rows(d3[:, 1:, :], dt)
    dt2 = dt.view(B * G, T)
    _wgrad(dt2, sv["tok"], gr("reduce_dim.weight"), gr("reduce_dim.bias"))
    dtok = ops.gemm_nn(dt2, pk.red_w)
    # ---- mini-PointNet backward (dvae.py:207-221)
    df2 = ops.group_maxpool_bwd(sv["f2"], K, dpool=dtok)
    y3 = sv["bn3"][1]
    _wgrad(df2, y3, gr2("encoder.second_conv.3.weight"), gr("encoder.second_conv.3.bias"))
    dy3 = ops.gemm_nn(df2, pk.c4_w)
    dz3 = _bn_train_bwd(dy3, pk.bn["bn3"], sv["bn3"], gr("encoder.second_conv.1.weight"), gr("encoder.second_conv.1.bias"))
    _wgrad(dz3, sv["cat"], gr2("encoder.second_conv.0.weight"), gr("encoder.second_conv.0.bias"))
    dcat = ops.gemm_nn(dz3, pk.c3_w)
    df1 = ops.group_maxpool_bwd(sv["f1"], K, dcat=dcat)
    y1 = sv["bn1"][1]
    _wgrad(df1, y1, gr2("encoder.first_conv.3.weight"), gr("encoder.first_conv.3.bias"))
    dy1 = ops.gemm_nn(df1, pk.c2_w)
    dz1 = _bn_train_bwd(dy1, pk.bn["bn1"], sv["bn1"], gr("encoder.first_conv.1.weight"), gr("encoder.first_conv.1.bias"))
    _wgrad_padded(dz1, sv["nb"], gr2("encoder.first_conv.0.weight"), gr("encoder.first_conv.0.bias"))
