"""Parameter tree with the reference's ``state_dict`` names, and the packed operand buffers the kernels read.

Reference names (probed from the real module): ``model.embed_tokens.weight``,
``model.layers.{i}.self_attn.{q,k,v,o}_proj.weight``, ``model.layers.{i}.mlp.{gate,up,down}_proj.weight``,
``model.layers.{i}.{input,post_attention}_layernorm.weight``, ``model.norm.weight``, ``model.point_backbone.*``
(pointbert/point_encoder.py:101-160, dvae.py:189-205), ``model.point_proj.{0,2,4}.*`` (pointllm.py:67-81), ``lm_head.weight``.

Storage: LLM trunk / embeddings / head / projector live in bf16 (what DeepSpeed's bf16 engine makes of them,
train.py:97-98); q|k|v and gate|up parameters are *views* into packed [3d,d] / [2f,d] buffers so the fused projections
need no second copy and ``load_state_dict`` fills the packed buffers directly.  The 22 M-parameter point encoder is kept
in fp32 and folded/packed (BatchNorm folded into the 1x1 convs, K padded to 64) by ``pack_point_encoder``.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, List, Tuple

import torch
import torch.nn as nn

from .config import TrajVLMConfig


def _set_nested(root: nn.Module, dotted: str, tensor: torch.Tensor, buffer: bool = False, requires_grad: bool = False):
    parts = dotted.split(".")
    m = root
    for p in parts[:-1]:
        if not hasattr(m, p):
            m.add_module(p, nn.Module())
        m = getattr(m, p)
    if buffer:
        m.register_buffer(parts[-1], tensor)
    else:
        m.register_parameter(parts[-1], nn.Parameter(tensor, requires_grad=requires_grad))


def point_encoder_shapes(cfg: TrajVLMConfig) -> "OrderedDict[str, Tuple[Tuple[int, ...], str]]":
    """name -> (shape, kind) for ``model.point_backbone`` ; kind in {param, buffer, buffer_long}."""
    p = cfg.point
    T, E = p.trans_dim, p.encoder_dims
    o: "OrderedDict[str, Tuple[Tuple[int, ...], str]]" = OrderedDict()
    pre = "model.point_backbone."
    o[pre + "cls_token"] = ((1, 1, T), "param")
    o[pre + "cls_pos"] = ((1, 1, T), "param")

    def conv(name, cout, cin):
        o[pre + name + ".weight"] = ((cout, cin, 1), "param")
        o[pre + name + ".bias"] = ((cout,), "param")

    def bn(name, c):
        o[pre + name + ".weight"] = ((c,), "param")
        o[pre + name + ".bias"] = ((c,), "param")
        o[pre + name + ".running_mean"] = ((c,), "buffer")
        o[pre + name + ".running_var"] = ((c,), "buffer")
        o[pre + name + ".num_batches_tracked"] = ((), "buffer_long")

    def lin(name, cout, cin, bias=True):
        o[pre + name + ".weight"] = ((cout, cin), "param")
        if bias:
            o[pre + name + ".bias"] = ((cout,), "param")

    conv("encoder.first_conv.0", 128, p.point_dims); bn("encoder.first_conv.1", 128); conv("encoder.first_conv.3", 256, 128)
    conv("encoder.second_conv.0", 512, 512); bn("encoder.second_conv.1", 512); conv("encoder.second_conv.3", E, 512)
    lin("reduce_dim", T, E)
    lin("pos_embed.0", 128, 3); lin("pos_embed.2", T, 128)
    for j in range(p.depth):
        b = f"blocks.blocks.{j}."
        lin(b + "norm1", T, 0, bias=True); o[pre + b + "norm1.weight"] = ((T,), "param")
        lin(b + "norm2", T, 0, bias=True); o[pre + b + "norm2.weight"] = ((T,), "param")
        lin(b + "mlp.fc1", 4 * T, T); lin(b + "mlp.fc2", T, 4 * T)
        lin(b + "attn.qkv", 3 * T, T, bias=False); lin(b + "attn.proj", T, T)
    o[pre + "norm.weight"] = ((T,), "param"); o[pre + "norm.bias"] = ((T,), "param")
    return o


def projector_dims(cfg: TrajVLMConfig) -> List[Tuple[int, int]]:
    p = cfg.point
    last = p.trans_dim * (2 if p.use_max_pool else 1)
    dims = []
    for h in p.projection_hidden_dim:
        dims.append((h, last)); last = h
    dims.append((cfg.hidden_size, last))
    return dims


class ParamTree(nn.Module):
    """All parameters of the trajectory VLM under the reference's names (no compute in here)."""

    def __init__(self, cfg: TrajVLMConfig, device, llm_dtype=torch.bfloat16):
        super().__init__()
        self.cfg_ = cfg
        d, f, V = cfg.hidden_size, cfg.intermediate_size, cfg.vocab_size
        kw = dict(device=device, dtype=llm_dtype)
        _set_nested(self, "model.embed_tokens.weight", torch.zeros(V, d, **kw))
        self.wqkv: List[torch.Tensor] = []
        self.wgu: List[torch.Tensor] = []
        for i in range(cfg.num_hidden_layers):
            pre = f"model.layers.{i}."
            qkv = torch.zeros(3 * d, d, **kw)
            gu = torch.zeros(2 * f, d, **kw)
            self.wqkv.append(qkv); self.wgu.append(gu)
            _set_nested(self, pre + "self_attn.q_proj.weight", qkv[:d])
            _set_nested(self, pre + "self_attn.k_proj.weight", qkv[d:2 * d])
            _set_nested(self, pre + "self_attn.v_proj.weight", qkv[2 * d:])
            _set_nested(self, pre + "self_attn.o_proj.weight", torch.zeros(d, d, **kw))
            _set_nested(self, pre + "mlp.gate_proj.weight", gu[:f])
            _set_nested(self, pre + "mlp.up_proj.weight", gu[f:])
            _set_nested(self, pre + "mlp.down_proj.weight", torch.zeros(d, f, **kw))
            _set_nested(self, pre + "input_layernorm.weight", torch.ones(d, **kw))
            _set_nested(self, pre + "post_attention_layernorm.weight", torch.ones(d, **kw))
        _set_nested(self, "model.norm.weight", torch.ones(d, **kw))
        for name, (shape, kind) in point_encoder_shapes(cfg).items():
            if kind == "param":
                _set_nested(self, name, torch.zeros(shape, device=device, dtype=torch.float32))
            elif kind == "buffer":
                _set_nested(self, name, torch.ones(shape, device=device, dtype=torch.float32) if name.endswith("running_var")
                            else torch.zeros(shape, device=device, dtype=torch.float32), buffer=True)
            else:
                _set_nested(self, name, torch.zeros(shape, device=device, dtype=torch.long), buffer=True)
        dims = projector_dims(cfg)
        if len(dims) == 1:                                   # pointllm.py:80-81 single nn.Linear
            _set_nested(self, "model.point_proj.weight", torch.zeros(dims[0], **kw))
            _set_nested(self, "model.point_proj.bias", torch.zeros(dims[0][0], **kw))
        else:                                                # pointllm.py:67-79 nn.Sequential(Linear, GELU, ...)
            for li, (co, ci) in enumerate(dims):
                _set_nested(self, f"model.point_proj.{2 * li}.weight", torch.zeros(co, ci, **kw))
                _set_nested(self, f"model.point_proj.{2 * li}.bias", torch.zeros(co, **kw))
        _set_nested(self, "lm_head.weight", torch.zeros(V, d, **kw))

    def named(self) -> Dict[str, torch.Tensor]:
        out = dict(self.named_parameters())
        out.update(dict(self.named_buffers()))
        return out

    def projector_params(self) -> List[Tuple[str, str]]:
        dims = projector_dims(self.cfg_)
        if len(dims) == 1:
            return [("model.point_proj.weight", "model.point_proj.bias")]
        return [(f"model.point_proj.{2 * i}.weight", f"model.point_proj.{2 * i}.bias") for i in range(len(dims))]
