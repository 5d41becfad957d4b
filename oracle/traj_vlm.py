"""Plain-PyTorch fp32 restatement of the reference trajectory model (TEST INFRASTRUCTURE).

This is the CPU oracle for the hot path, not product code (see ``oracle/__init__``).
Every function cites the reference file:line it follows (paths relative to
``/root/reference/egoscaler/models/pointllm/``).  The dense LLaMA arithmetic of the
reference lives in the un-vendored third-party dependency ``transformers``
(installed/pinned here: 5.5.0, ``models/llama/modeling_llama.py``); its published
algorithm is restated below and the call sites that anchor it are
``pointllm/model/pointllm.py:173-178`` (``super().forward``) and ``:228`` (lm_head).

Parity pinning: pinned against the imported reference itself through
``tests/golden/*.npz`` (made by ``oracle/make_golden.py``); the reference ships no
tests of its own for this path.

Module/parameter names mirror the reference ``state_dict`` exactly so a reference
checkpoint loads with ``load_state_dict(strict=True)``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F


# ----------------------------------------------------------------------------- configs
@dataclass
class PointCfg:
    """PointBERT yaml (pointllm/model/pointbert/PointTransformer_8192point_2layer.yaml:1-16)."""
    trans_dim: int = 384
    depth: int = 12
    num_heads: int = 6
    group_size: int = 32
    num_group: int = 512
    encoder_dims: int = 256
    point_dims: int = 6
    proj_hidden: Tuple[int, ...] = (1024, 2048)
    use_max_pool: bool = False
    drop_path_rate: float = 0.1         # both yaml files; active only when the encoder trains (--unfreeze_pc_encoder)

    @staticmethod
    def base() -> "PointCfg":
        """PointTransformer_base_8192point.yaml:1-13 (v1.1: single-Linear projector)."""
        return PointCfg(trans_dim=1152, depth=12, num_heads=12, group_size=48, num_group=512,
                        encoder_dims=512, point_dims=6, proj_hidden=())


@dataclass
class LlmCfg:
    hidden_size: int = 256
    intermediate_size: int = 688
    num_layers: int = 2
    num_heads: int = 4
    vocab_size: int = 32262
    rms_eps: float = 1e-6
    rope_theta: float = 10000.0
    pad_token_id: int = 0
    bos_token_id: int = 1
    eos_token_id: int = 2
    point_patch_token: int = 32000
    point_start_token: int = 32001
    point_end_token: int = 32002
    mm_use_point_start_end: bool = True

    @staticmethod
    def tiny() -> "LlmCfg":
        return LlmCfg()

    @staticmethod
    def tiny_hd128() -> "LlmCfg":
        """Tiny trunk with the FULL-size head width (2 heads x 128): pins the head_dim-128 attention / decode kernels."""
        return LlmCfg(num_heads=2)

    @staticmethod
    def full() -> "LlmCfg":
        return LlmCfg(hidden_size=4096, intermediate_size=11008, num_layers=32, num_heads=32)


# ----------------------------------------------------------------------------- point encoder
def fps(xyz: torch.Tensor, npoint: int, start: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Farthest point sampling -- pointbert/misc.py:40-60.

    The reference draws the start index with ``torch.randint`` on the CPU generator
    (misc.py:52); ``start`` lets a caller inject the same draw.  Returns centre xyz
    [B, npoint, 3] (the reference returns ``index_points(xyz, centroids)``).
    """
    B, N, _ = xyz.shape
    dev = xyz.device                     # device-agnostic: bench.py's gpu_baseline runs this same code on cuda
    centroids = torch.zeros(B, npoint, dtype=torch.long, device=dev)
    distance = torch.ones(B, N, device=dev) * 1e10
    farthest = (torch.randint(0, N, (B,), dtype=torch.long) if start is None else start.clone()).to(dev)
    batch = torch.arange(B, device=dev)
    for i in range(npoint):
        centroids[:, i] = farthest
        c = xyz[batch, farthest, :].view(B, 1, 3)
        dist = torch.sum((xyz - c) ** 2, -1)
        distance = torch.min(distance, dist)
        farthest = torch.max(distance, -1)[1]
    return xyz[batch[:, None], centroids, :]


def knn_group(pts: torch.Tensor, center: torch.Tensor, k: int) -> torch.Tensor:
    """kNN grouping -- pointbert/dvae.py:107-140 (square_distance + topk) and :150-187 (Group).

    Returns neighbourhood [B, G, k, C] with xyz re-centred on the group centre (dvae.py:182)
    and the extra channels (rgb) gathered unchanged (dvae.py:178-186).
    """
    xyz = pts[..., :3]
    B, N, _ = xyz.shape
    d = -2 * torch.matmul(center, xyz.permute(0, 2, 1))
    d += torch.sum(center ** 2, -1).view(B, -1, 1)
    d += torch.sum(xyz ** 2, -1).view(B, 1, N)
    idx = torch.topk(d, k, dim=-1, largest=False, sorted=False)[1]           # [B,G,k]
    bidx = torch.arange(B, device=pts.device).view(B, 1, 1)
    nb = pts[bidx, idx, :]                                                    # [B,G,k,C]
    nb = torch.cat((nb[..., :3] - center.unsqueeze(2), nb[..., 3:]), dim=-1)
    return nb


class _Encoder(nn.Module):
    """mini-PointNet -- pointbert/dvae.py:189-221."""

    def __init__(self, channel: int, in_dims: int):
        super().__init__()
        self.channel = channel
        self.first_conv = nn.Sequential(nn.Conv1d(in_dims, 128, 1), nn.BatchNorm1d(128),
                                        nn.ReLU(inplace=True), nn.Conv1d(128, 256, 1))
        self.second_conv = nn.Sequential(nn.Conv1d(512, 512, 1), nn.BatchNorm1d(512),
                                         nn.ReLU(inplace=True), nn.Conv1d(512, channel, 1))

    def forward(self, g: torch.Tensor) -> torch.Tensor:
        bs, G, n, c = g.shape
        x = g.reshape(bs * G, n, c).transpose(2, 1)
        f = self.first_conv(x)
        fg = torch.max(f, dim=2, keepdim=True)[0]
        f = torch.cat([fg.expand(-1, -1, n), f], dim=1)
        f = self.second_conv(f)
        return torch.max(f, dim=2)[0].reshape(bs, G, self.channel)


class _Attn(nn.Module):
    """pointbert/point_encoder.py:30-55 (qkv without bias, explicit softmax(QK^T*scale)V)."""

    def __init__(self, dim: int, heads: int):
        super().__init__()
        self.heads = heads
        self.scale = (dim // heads) ** -0.5
        self.qkv = nn.Linear(dim, dim * 3, bias=False)
        self.proj = nn.Linear(dim, dim)

    def forward(self, x):
        B, N, C = x.shape
        qkv = self.qkv(x).reshape(B, N, 3, self.heads, C // self.heads).permute(2, 0, 3, 1, 4)
        q, k, v = qkv[0], qkv[1], qkv[2]
        a = ((q @ k.transpose(-2, -1)) * self.scale).softmax(dim=-1)
        return self.proj((a @ v).transpose(1, 2).reshape(B, N, C))


class _Mlp(nn.Module):
    """pointbert/point_encoder.py:11-27."""

    def __init__(self, dim: int, hidden: int):
        super().__init__()
        self.fc1 = nn.Linear(dim, hidden)
        self.fc2 = nn.Linear(hidden, dim)

    def forward(self, x):
        return self.fc2(F.gelu(self.fc1(x)))


class _Block(nn.Module):
    """pointbert/point_encoder.py:58-76 (pre-LN).  DropPath = timm 0.4.12 ``drop_path``: identity in eval or at rate 0
    (nn.Identity, :62); in train mode ``x / keep * floor(keep + U[0,1))`` with one draw per sample.  ``u`` injects the
    two uniform draws [2, B] of this block (attention branch, MLP branch); None draws them like timm does."""

    def __init__(self, dim: int, heads: int, drop_path: float = 0.0):
        super().__init__()
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.mlp = _Mlp(dim, int(dim * 4.0))
        self.attn = _Attn(dim, heads)
        self.drop_path = float(drop_path)

    def _dp(self, y, u):
        if self.drop_path == 0.0 or not self.training:
            return y
        keep = 1.0 - self.drop_path
        r = torch.rand((y.shape[0], 1, 1), dtype=y.dtype, device=y.device) if u is None else u.reshape(-1, 1, 1).to(y.dtype)
        return y.div(keep) * (keep + r).floor()

    def forward(self, x, u=None):
        x = x + self._dp(self.attn(self.norm1(x)), None if u is None else u[0])
        return x + self._dp(self.mlp(self.norm2(x)), None if u is None else u[1])


class _Blocks(nn.Module):
    """pointbert/point_encoder.py:79-98: ``x = block(x + pos)`` at every depth."""

    def __init__(self, dim: int, depth: int, heads: int, drop_path_rate: float = 0.0):
        super().__init__()
        dpr = [v.item() for v in torch.linspace(0, drop_path_rate, depth, device="cpu")]            # point_encoder.py:134
        self.blocks = nn.ModuleList([_Block(dim, heads, dpr[i]) for i in range(depth)])

    def forward(self, x, pos, drop_u=None):
        for i, b in enumerate(self.blocks):
            x = b(x + pos, None if drop_u is None else drop_u[i])
        return x


class PointTransformer(nn.Module):
    """pointbert/point_encoder.py:101-189."""

    def __init__(self, cfg: PointCfg):
        super().__init__()
        self.cfg = cfg
        self.encoder = _Encoder(cfg.encoder_dims, cfg.point_dims)
        self.reduce_dim = nn.Linear(cfg.encoder_dims, cfg.trans_dim)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, cfg.trans_dim))
        self.cls_pos = nn.Parameter(torch.randn(1, 1, cfg.trans_dim))
        self.pos_embed = nn.Sequential(nn.Linear(3, 128), nn.GELU(), nn.Linear(128, cfg.trans_dim))
        self.blocks = _Blocks(cfg.trans_dim, cfg.depth, cfg.num_heads, cfg.drop_path_rate)
        self.drop_u: Optional[torch.Tensor] = None      # [depth, 2, B] injected DropPath uniforms (tests)
        self.norm = nn.LayerNorm(cfg.trans_dim)

    def forward(self, pts: torch.Tensor, fps_start: Optional[torch.Tensor] = None) -> torch.Tensor:
        center = fps(pts[..., :3].contiguous(), self.cfg.num_group, fps_start)      # :171 -> dvae.py:165
        nb = knn_group(pts, center, self.cfg.group_size)
        tok = self.reduce_dim(self.encoder(nb))                                     # :173-174
        B = tok.size(0)
        x = torch.cat((self.cls_token.expand(B, -1, -1), tok), dim=1)               # :181
        pos = torch.cat((self.cls_pos.expand(B, -1, -1), self.pos_embed(center)), dim=1)
        x = self.norm(self.blocks(x, pos, self.drop_u))                             # :184-185
        if not self.cfg.use_max_pool:
            return x
        return torch.cat([x[:, 0], x[:, 1:].max(1)[0]], dim=-1).unsqueeze(1)        # :188


# ----------------------------------------------------------------------------- LLaMA (transformers 5.5.0)
class _RMSNorm(nn.Module):
    """HF LlamaRMSNorm (modeling_llama.py:53-67): fp32 stats, cast, then * weight."""

    def __init__(self, dim: int, eps: float):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(dim))
        self.eps = eps

    def forward(self, x):
        dt = x.dtype
        x32 = x.float()
        x32 = x32 * torch.rsqrt(x32.pow(2).mean(-1, keepdim=True) + self.eps)
        return self.weight * x32.to(dt)


def rope_tables(head_dim: int, theta: float, positions: torch.Tensor):
    """HF LlamaRotaryEmbedding (modeling_llama.py:117-135): fp32 freqs, cat(freqs, freqs)."""
    inv = 1.0 / (theta ** (torch.arange(0, head_dim, 2, dtype=torch.float32, device=positions.device) / head_dim))
    fr = positions.float()[:, None] * inv[None, :]
    emb = torch.cat((fr, fr), dim=-1)
    return emb.cos(), emb.sin()


def _rot_half(x):
    h = x.shape[-1] // 2
    return torch.cat((-x[..., h:], x[..., :h]), dim=-1)


class _SelfAttn(nn.Module):
    """HF LlamaAttention (modeling_llama.py:253-290) with sdpa semantics (bool mask, scale hd^-0.5)."""

    def __init__(self, d: int, heads: int):
        super().__init__()
        self.heads, self.hd = heads, d // heads
        self.q_proj = nn.Linear(d, d, bias=False)
        self.k_proj = nn.Linear(d, d, bias=False)
        self.v_proj = nn.Linear(d, d, bias=False)
        self.o_proj = nn.Linear(d, d, bias=False)

    def forward(self, x, cos, sin, mask, cache):
        B, Lq, _ = x.shape
        q = self.q_proj(x).view(B, Lq, self.heads, self.hd).transpose(1, 2)
        k = self.k_proj(x).view(B, Lq, self.heads, self.hd).transpose(1, 2)
        v = self.v_proj(x).view(B, Lq, self.heads, self.hd).transpose(1, 2)
        q = q * cos + _rot_half(q) * sin                       # apply_rotary_pos_emb :138-168
        k = k * cos + _rot_half(k) * sin
        if cache is not None:                                  # DynamicCache.update: keys stored post-RoPE
            if cache[0] is not None:
                k = torch.cat((cache[0], k), dim=2)
                v = torch.cat((cache[1], v), dim=2)
            cache[0], cache[1] = k, v
        s = (q @ k.transpose(-1, -2)) * (self.hd ** -0.5)
        s = s.masked_fill(~mask, float("-inf"))
        o = torch.softmax(s, dim=-1, dtype=torch.float32).to(q.dtype) @ v
        return self.o_proj(o.transpose(1, 2).reshape(B, Lq, -1))


class _LlamaMlp(nn.Module):
    """HF LlamaMLP (modeling_llama.py:171-184): down(silu(gate(x)) * up(x))."""

    def __init__(self, d: int, f: int):
        super().__init__()
        self.gate_proj = nn.Linear(d, f, bias=False)
        self.up_proj = nn.Linear(d, f, bias=False)
        self.down_proj = nn.Linear(f, d, bias=False)

    def forward(self, x):
        return self.down_proj(F.silu(self.gate_proj(x)) * self.up_proj(x))


class _Layer(nn.Module):
    """HF LlamaDecoderLayer (modeling_llama.py:292-340)."""

    def __init__(self, c: LlmCfg):
        super().__init__()
        self.self_attn = _SelfAttn(c.hidden_size, c.num_heads)
        self.mlp = _LlamaMlp(c.hidden_size, c.intermediate_size)
        self.input_layernorm = _RMSNorm(c.hidden_size, c.rms_eps)
        self.post_attention_layernorm = _RMSNorm(c.hidden_size, c.rms_eps)

    def forward(self, x, cos, sin, mask, cache):
        x = x + self.self_attn(self.input_layernorm(x), cos, sin, mask, cache)
        return x + self.mlp(self.post_attention_layernorm(x))


class _Model(nn.Module):
    """PointLLMLlamaModel -- pointllm/model/pointllm.py:26-178."""

    def __init__(self, c: LlmCfg, p: PointCfg):
        super().__init__()
        self.c, self.p = c, p
        self.embed_tokens = nn.Embedding(c.vocab_size, c.hidden_size, padding_idx=c.pad_token_id)
        self.layers = nn.ModuleList([_Layer(c) for _ in range(c.num_layers)])
        self.norm = _RMSNorm(c.hidden_size, c.rms_eps)
        self.point_backbone = PointTransformer(p)
        out_dim = p.trans_dim * (2 if p.use_max_pool else 1)
        if len(p.proj_hidden) > 0:                                            # pointllm.py:67-81
            layers, last = [], out_dim
            for h in p.proj_hidden:
                layers += [nn.Linear(last, h), nn.GELU()]
                last = h
            layers.append(nn.Linear(last, c.hidden_size))
            self.point_proj = nn.Sequential(*layers)
        else:
            self.point_proj = nn.Linear(out_dim, c.hidden_size)
        self.point_token_len = 1 if p.use_max_pool else p.num_group + 1

    def splice(self, input_ids, embeds, feats):
        """pointllm.py:134-171, mm_use_point_start_end branch and the plain-patch branch."""
        c, out = self.c, []
        for ids, e, f in zip(input_ids, embeds, feats):
            if (ids == c.point_patch_token).sum() == 0:
                out.append(e)
                continue
            n = f.shape[0]
            if c.mm_use_point_start_end:
                if (ids == c.point_start_token).sum() != (ids == c.point_end_token).sum():
                    raise ValueError("The number of point start tokens and point end tokens should be the same.")
                for s in torch.where(ids == c.point_start_token)[0]:
                    if ids[s + n + 1] != c.point_end_token:
                        raise ValueError("The point end token should follow the point start token.")
                    e = torch.cat((e[:s + 1], f, e[s + n + 1:]), dim=0)
            else:
                if (ids == c.point_patch_token).sum() != n:
                    raise ValueError("The number of point patch tokens should be the same as the number of point patches.")
                m = torch.where(ids == c.point_patch_token)[0]
                if (m != torch.arange(int(m[0]), int(m[0]) + n, device=m.device)).any():
                    raise ValueError("The point patch tokens should be consecutive.")
                e = torch.cat((e[:m[0]], f, e[m[0] + n:]), dim=0)
            out.append(e)
        return torch.stack(out, dim=0)

    def forward(self, input_ids, attention_mask=None, point_clouds=None, cache=None, fps_start=None,
                training=False):
        x = self.embed_tokens(input_ids)                                           # pointllm.py:107
        if (input_ids.shape[1] != 1 or training) and point_clouds is not None:     # pointllm.py:112
            feats = self.point_proj(self.point_backbone(point_clouds, fps_start))  # :124,:129
            x = self.splice(input_ids, x, feats)
        B, Lq, _ = x.shape
        past = 0 if (cache is None or cache[0][0] is None) else cache[0][0].shape[2]
        pos = torch.arange(past, past + Lq, device=x.device)                                        # HF :394-397
        cos, sin = rope_tables(self.c.hidden_size // self.c.num_heads, self.c.rope_theta, pos)
        Lk = past + Lq
        causal = torch.arange(Lk, device=x.device)[None, :] <= pos[:, None]                         # [Lq,Lk]
        mask = causal[None, None]
        if attention_mask is not None:
            mask = mask & attention_mask.bool()[:, None, None, :Lk]
        for i, layer in enumerate(self.layers):
            x = layer(x, cos, sin, mask, None if cache is None else cache[i])
        return self.norm(x)


class TrajVLMOracle(nn.Module):
    """PointLLMLlamaForCausalLM (pointllm.py:181-253) with TrajPointLLMForCausalLM's
    freezing semantics (model_arch.py:33-51,110-124) applied by ``configure_trainable``."""

    def __init__(self, c: LlmCfg, p: PointCfg):
        super().__init__()
        self.c, self.p = c, p
        self.model = _Model(c, p)
        self.lm_head = nn.Linear(c.hidden_size, c.vocab_size, bias=False)

    def configure_trainable(self, unfreeze_pc_encoder=False, unfreeze_language_model=False):
        for q in self.model.point_backbone.parameters():
            q.requires_grad = bool(unfreeze_pc_encoder)
        for q in self.model.layers.parameters():
            q.requires_grad = bool(unfreeze_language_model)
        for q in self.model.embed_tokens.parameters():
            q.requires_grad = True
        return self

    def new_cache(self):
        return [[None, None] for _ in range(self.c.num_layers)]

    def forward(self, input_ids, attention_mask=None, point_clouds=None, cache=None, fps_start=None,
                training=False):
        h = self.model(input_ids, attention_mask, point_clouds, cache, fps_start, training)
        return self.lm_head(h)                                                     # pointllm.py:228


def train_loss(logits: torch.Tensor, tokens: torch.Tensor, prompt_len: int, pad_id: int) -> torch.Tensor:
    """Loss slice -- train.py:174-181."""
    lg = logits[:, prompt_len - 1:-1, :]
    tg = tokens[:, prompt_len:]
    return F.cross_entropy(lg.reshape(-1, lg.shape[-1]), tg.flatten(), ignore_index=pad_id)


@torch.no_grad()
def greedy_decode(model: TrajVLMOracle, prompts, prompt_masks, point_clouds, max_new_tokens: int,
                  fps_start=None, suppress_eos: bool = False, return_scores: bool = False):
    """Manual KV-cache greedy loop over ``forward`` (the decode oracle).

    Semantics = HF GenerationMixin greedy (generation/utils.py:2762-2806) as reached through
    model_arch.py:77-108 / pointllm.py:255-275: full-prompt step runs the point encoder, later
    steps feed one token with the cache; finished rows emit ``pad``; mask grows by ones.
    """
    c = model.c
    cache = model.new_cache()
    ids, mask = prompts.clone(), prompt_masks.clone().bool()
    unfinished = torch.ones(ids.shape[0], dtype=torch.long)
    scores = []
    cur = ids
    for step in range(max_new_tokens):
        logits = model(cur, mask, point_clouds if step == 0 else None, cache, fps_start)[:, -1].float()
        if suppress_eos:
            logits[:, c.eos_token_id] = float("-inf")
        if return_scores:
            scores.append(logits.clone())
        nxt = torch.argmax(logits, dim=-1)
        nxt = nxt * unfinished + c.pad_token_id * (1 - unfinished)
        ids = torch.cat((ids, nxt[:, None]), dim=1)
        mask = torch.cat((mask, torch.ones(ids.shape[0], 1, dtype=torch.bool)), dim=1)
        unfinished = unfinished & (nxt != c.eos_token_id).long()
        cur = nxt[:, None]
        if unfinished.max() == 0:
            break
    return (ids, scores) if return_scores else ids
