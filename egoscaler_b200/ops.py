"""Tensor-level wrappers over the C ABI (``include/egoscaler_b200.h``).

Each function takes CUDA ``torch.Tensor``s, checks layout, and calls the matching ``esb_*`` entry point with raw
device pointers on torch's current stream.  Nothing here computes on the CPU and nothing falls back to ATen:
a missing library or a non-CUDA tensor raises.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import lib as _l

ACT_NONE, ACT_GELU, ACT_RELU = 0, 1, 2
SCALE_NONE, SCALE_RT2, SCALE_SIMPLE, SCALE_DENORM = 0, 1, 2, 3


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _chk(t: torch.Tensor, dtype, name: str):
    if not t.is_cuda:
        raise _l.EsbError(f"{name}: expected a CUDA tensor (there is no CPU path)")
    if t.dtype != dtype:
        raise _l.EsbError(f"{name}: expected {dtype}, got {t.dtype}")


# ------------------------------------------------------------------------------------------------ GEMM
def gemm(a: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None,
         residual: Optional[torch.Tensor] = None, act: int = ACT_NONE, out: Optional[torch.Tensor] = None,
         out_dtype: torch.dtype = torch.bfloat16, accumulate: bool = False, alpha: float = 1.0,
         force_bn: int = 0) -> torch.Tensor:
    """``act(alpha * a @ w.T + bias) + residual`` with a [M,K] bf16 (row stride any multiple of 8), w [N,K] bf16."""
    _chk(a, torch.bfloat16, "gemm.a")
    _chk(w, torch.bfloat16, "gemm.w")
    assert a.dim() == 2 and w.dim() == 2 and a.shape[1] == w.shape[1], (a.shape, w.shape)
    assert a.stride(1) == 1 and w.stride(1) == 1
    M, K = a.shape
    N = w.shape[0]
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
    assert out.shape == (M, N) and out.stride(1) == 1
    if bias is not None:
        _chk(bias, torch.float32, "gemm.bias")
        assert bias.numel() == N and bias.is_contiguous()
    ldr = 0
    if residual is not None:
        _chk(residual, torch.bfloat16, "gemm.residual")
        assert residual.shape == (M, N) and residual.stride(1) == 1
        ldr = residual.stride(0)
    _l.call("esb_gemm_bf16", a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), out.data_ptr(), out.stride(0),
            M, N, K, _p(bias), _p(residual), ldr, act, 1 if out.dtype == torch.float32 else 0,
            1 if accumulate else 0, float(alpha), force_bn, _stream())
    return out


# ------------------------------------------------------------------------------------------------ trajectory tokens
def traj_quantize(x: torch.Tensor, num_bins: int = 256, id_offset: int = 0) -> torch.Tensor:
    """utils/utils.py:13-16 ``discretize_action`` (+ ``id_offset``); x float32 or float64, any shape."""
    if not x.is_cuda:
        raise _l.EsbError("traj_quantize: expected a CUDA tensor")
    assert x.dtype in (torch.float32, torch.float64)
    x = x.contiguous()
    out = torch.empty(x.shape, device=x.device, dtype=torch.int64)
    _l.call("esb_traj_quantize", x.data_ptr(), 1 if x.dtype == torch.float64 else 0, x.numel(), num_bins, id_offset,
            out.data_ptr(), _stream())
    return out


def traj_dequantize(ids: torch.Tensor, num_bins: int = 256, id_offset: int = 0, scale: int = SCALE_NONE,
                    d_max: float = 0.0, d_min: float = 0.0) -> torch.Tensor:
    """utils/utils.py:18-21 + :96 (+ scaler); ids [..., 6] int64 -> float32 [..., 6]."""
    _chk(ids, torch.int64, "traj_dequantize.ids")
    assert ids.shape[-1] == 6
    ids = ids.contiguous()
    out = torch.empty(ids.shape, device=ids.device, dtype=torch.float32)
    _l.call("esb_traj_dequantize", ids.data_ptr(), ids.numel() // 6, num_bins, id_offset, scale, float(d_max),
            float(d_min), out.data_ptr(), _stream())
    return out


def traj_scale_(poses: torch.Tensor, scale: int, d_max: float = 0.0, d_min: float = 0.0) -> torch.Tensor:
    _chk(poses, torch.float32, "traj_scale.poses")
    assert poses.shape[-1] == 6 and poses.is_contiguous()
    _l.call("esb_traj_scale", poses.data_ptr(), poses.numel() // 6, scale, float(d_max), float(d_min), _stream())
    return poses


def traj_ids_to_poses(ids: torch.Tensor, max_steps: int, num_bins: int = 256, base_id: int = 32006,
                      tsep_id: int = 32004, eos_id: int = 2, scale: int = SCALE_NONE, d_max: float = 0.0,
                      d_min: float = 0.0, pad_repeat: bool = True):
    """Generated ids [B,T] -> (poses [B,max_steps,6] float32, count [B] int32); utils/utils.py:47-104 on ids."""
    _chk(ids, torch.int64, "traj_ids_to_poses.ids")
    ids = ids.contiguous()
    B, T = ids.shape
    out = torch.zeros(B, max_steps, 6, device=ids.device, dtype=torch.float32)
    cnt = torch.zeros(B, device=ids.device, dtype=torch.int32)
    _l.call("esb_traj_ids_to_poses", ids.data_ptr(), B, T, num_bins, base_id, tsep_id, eos_id, scale, float(d_max),
            float(d_min), max_steps, 1 if pad_repeat else 0, out.data_ptr(), cnt.data_ptr(), _stream())
    return out, cnt
