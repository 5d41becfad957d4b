"""Import the REAL reference implementation from /root/reference (dev container only).

TEST INFRASTRUCTURE.  The reference needs a few pure-Python deps that are not installed
(easydict, timm's DropPath, matplotlib, termcolor); they are stubbed here.  ``transformers``
must be imported before the stubs (its lazy ``is_timm_available`` probe reads module specs).
``/root/reference`` does not exist on the GPU box: only ``oracle/make_golden.py`` and tests that
skip when the tree is absent may call this.
"""
from __future__ import annotations

import importlib.machinery as _im
import os
import sys
import types

REF_ROOT = os.environ.get("EGOSCALER_REF", "/root/reference")


def available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "egoscaler", "models", "pointllm"))


def _stub(name, **attrs):
    if name in sys.modules:
        return
    m = types.ModuleType(name)
    m.__spec__ = _im.ModuleSpec(name, None)
    m.__dict__.update(attrs)
    sys.modules[name] = m


_loaded = {}


def load():
    """Returns (PointLLMConfig, PointLLMLlamaForCausalLM, ref_utils_module)."""
    if _loaded:
        return _loaded["cfg"], _loaded["model"], _loaded["utils"]
    if not available():
        raise RuntimeError(f"reference tree not found at {REF_ROOT}")
    import transformers  # noqa: F401  (must precede the stubs)
    import torch
    import torch.nn as nn

    class EasyDict(dict):
        def __setitem__(s, k, v):
            dict.__setitem__(s, k, EasyDict(v) if isinstance(v, dict) and not isinstance(v, EasyDict) else v)

        def __init__(s, d=None):
            dict.__init__(s)
            for k, v in (d or {}).items():
                s.__setitem__(k, v)
        __setattr__ = __setitem__

        def __getattr__(s, k):
            try:
                return s[k]
            except KeyError:
                raise AttributeError(k)

    class DropPath(nn.Module):  # timm 0.4.12 semantics; identity in eval
        def __init__(s, p=0.):
            super().__init__()
            s.drop_prob = p

        def forward(s, x):
            if s.drop_prob == 0. or not s.training:
                return x
            keep = 1 - s.drop_prob
            r = (keep + torch.rand((x.shape[0],) + (1,) * (x.ndim - 1), dtype=x.dtype, device=x.device)).floor_()
            return x.div(keep) * r

    _stub('easydict', EasyDict=EasyDict)
    _stub('timm'); _stub('timm.models'); _stub('timm.models.layers', DropPath=DropPath)
    _stub('matplotlib'); _stub('matplotlib.pyplot'); _stub('mpl_toolkits'); _stub('mpl_toolkits.mplot3d', Axes3D=object)
    _stub('termcolor', colored=lambda s, *a, **k: s)
    sys.path[:0] = [os.path.join(REF_ROOT, 'egoscaler', 'models', 'pointllm'), REF_ROOT]
    from pointllm.model.pointllm import PointLLMConfig, PointLLMLlamaForCausalLM
    # utils/utils.py:10 reads a non-existent attribute: alias it.
    from egoscaler.configs.camera import CameraConfig
    if not hasattr(CameraConfig.devices.aria, "focal_length"):
        CameraConfig.devices.aria.focal_length = CameraConfig.devices.aria.focal_len
    import importlib.util
    spec = importlib.util.spec_from_file_location(
        "_ref_traj_utils", os.path.join(REF_ROOT, "egoscaler", "models", "pointllm", "utils", "utils.py"))
    ref_utils = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_utils)
    _loaded.update(cfg=PointLLMConfig, model=PointLLMLlamaForCausalLM, utils=ref_utils)
    return PointLLMConfig, PointLLMLlamaForCausalLM, ref_utils


def build_reference_model(llm, point_cfg_name="PointTransformer_8192point_2layer"):
    """Construct the reference PointLLMLlamaForCausalLM for an ``oracle.traj_vlm.LlmCfg``."""
    PointLLMConfig, PointLLMLlamaForCausalLM, _ = load()
    cfg = PointLLMConfig(hidden_size=llm.hidden_size, intermediate_size=llm.intermediate_size,
                         num_hidden_layers=llm.num_layers, num_attention_heads=llm.num_heads,
                         num_key_value_heads=llm.num_heads, vocab_size=llm.vocab_size,
                         rms_norm_eps=llm.rms_eps, pad_token_id=llm.pad_token_id, bos_token_id=llm.bos_token_id,
                         eos_token_id=llm.eos_token_id, tie_word_embeddings=False)
    cfg.point_backbone = "PointBERT"
    cfg.point_backbone_config_name = point_cfg_name
    cfg.use_color = True
    cfg.mm_use_point_start_end = llm.mm_use_point_start_end
    cfg.DEFAULT_POINT_PATCH_TOKEN = "<point_patch>"
    cfg.DEFAULT_POINT_START_TOKEN = "<point_start>"
    cfg.DEFAULT_POINT_END_TOKEN = "<point_end>"
    m = PointLLMLlamaForCausalLM(cfg)
    pbc = m.get_model().point_backbone_config
    pbc['point_patch_token'] = llm.point_patch_token
    pbc['point_start_token'] = llm.point_start_token
    pbc['point_end_token'] = llm.point_end_token
    return m
