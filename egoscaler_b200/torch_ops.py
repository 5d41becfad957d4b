"""PyTorch custom ops (``torch.library``) over the C ABI: ``torch.ops.esb.*``.

north_star: "Python host code calls CUDA through PyTorch custom ops over a thin C-ABI layer".  ``ops.py`` is the thin layer
(one tensor wrapper per ``esb_*`` entry point); this module registers the differentiable building blocks of the path as
dispatcher-visible operators with fake (meta) kernels and autograd formulas whose backward ops are themselves ``esb`` kernels,
so they compose with ``torch.autograd`` / ``torch.library.opcheck`` and can be called from reference-side code:

    torch.ops.esb.linear(x, w)                      nn.Linear without bias (HF LlamaAttention / LlamaMLP / lm_head), [rows, in] bf16
    torch.ops.esb.rmsnorm(x, w, eps)                LlamaRMSNorm (modeling_llama.py:53-67)
    torch.ops.esb.swiglu(gu)                        act_fn(gate) * up on the packed gate|up projection (modeling_llama.py:171-184)
    torch.ops.esb.attention(qkv, kmask, causal)     sdpa on a packed [B, L, 3, H, hd] projection (sdpa_attention.py:40-98)
    torch.ops.esb.cross_entropy(logits, targets, ignore_index)      train.py:174-181 (mean over non-ignored rows)
    torch.ops.esb.traj_quantize(x, num_bins, id_offset) / traj_dequantize(ids, num_bins, id_offset)     utils/utils.py:13-21

The module-level fast paths (``Trainer``, ``TrajVLMFunction``) call the same kernels directly through ``ops.py`` - these
operators are the op-granular surface of the boundary, not a second implementation.  CUDA only: no CPU kernels are registered.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import ops

_LIB = "esb"


def _define():
    lib = torch.library

    # ------------------------------------------------------------------ linear
    @lib.custom_op(f"{_LIB}::linear", mutates_args=(), device_types="cuda")
    def linear(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        return ops.gemm(x, w)

    @linear.register_fake
    def _(x, w):
        return x.new_empty(x.shape[0], w.shape[0])

    @lib.custom_op(f"{_LIB}::linear_dgrad", mutates_args=(), device_types="cuda")
    def linear_dgrad(dy: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
        return ops.gemm_nn(dy.contiguous(), w)                               # dX = dY . W on the [out, in] parameter

    @linear_dgrad.register_fake
    def _(dy, w):
        return dy.new_empty(dy.shape[0], w.shape[1])

    @lib.custom_op(f"{_LIB}::linear_wgrad", mutates_args=(), device_types="cuda")
    def linear_wgrad(dy: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
        return ops.gemm_tn(dy.contiguous(), x, out_dtype=torch.bfloat16)     # dW = dY^T . X

    @linear_wgrad.register_fake
    def _(dy, x):
        return dy.new_empty(dy.shape[1], x.shape[1])

    def linear_setup(ctx, inputs, output):
        ctx.save_for_backward(*inputs)

    def linear_bwd(ctx, dy):
        x, w = ctx.saved_tensors
        dx = torch.ops.esb.linear_dgrad(dy, w) if ctx.needs_input_grad[0] else None
        dw = torch.ops.esb.linear_wgrad(dy, x) if ctx.needs_input_grad[1] else None
        return dx, dw

    linear.register_autograd(linear_bwd, setup_context=linear_setup)

    # ------------------------------------------------------------------ rmsnorm
    @lib.custom_op(f"{_LIB}::rmsnorm_fwd", mutates_args=(), device_types="cuda")
    def rmsnorm_fwd(x: torch.Tensor, w: torch.Tensor, eps: float) -> Tuple[torch.Tensor, torch.Tensor]:
        y, rstd = ops.rmsnorm(x, w, eps, want_rstd=True)
        return y, rstd

    @rmsnorm_fwd.register_fake
    def _(x, w, eps):
        return torch.empty_like(x), x.new_empty(x.shape[0], dtype=torch.float32)

    @lib.custom_op(f"{_LIB}::rmsnorm_bwd", mutates_args=(), device_types="cuda")
    def rmsnorm_bwd(dy: torch.Tensor, x: torch.Tensor, w: torch.Tensor, rstd: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        dw = torch.zeros(w.shape, device=w.device, dtype=torch.float32)
        dx = ops.rmsnorm_bwd(dy.contiguous(), x.contiguous(), w, rstd, dw=dw)
        return dx, dw

    @rmsnorm_bwd.register_fake
    def _(dy, x, w, rstd):
        return torch.empty_like(x), w.new_empty(w.shape, dtype=torch.float32)

    def rms_setup(ctx, inputs, output):
        x, w, eps = inputs
        ctx.save_for_backward(x, w, output[1])

    def rms_bwd(ctx, dy, _drstd):
        x, w, rstd = ctx.saved_tensors
        dx, dw = torch.ops.esb.rmsnorm_bwd(dy, x, w, rstd)
        return dx, dw.to(w.dtype), None

    rmsnorm_fwd.register_autograd(rms_bwd, setup_context=rms_setup)

    # ------------------------------------------------------------------ swiglu
    @lib.custom_op(f"{_LIB}::swiglu", mutates_args=(), device_types="cuda")
    def swiglu(gu: torch.Tensor) -> torch.Tensor:
        return ops.silu_mul(gu)

    @swiglu.register_fake
    def _(gu):
        return gu.new_empty(gu.shape[0], gu.shape[1] // 2)

    @lib.custom_op(f"{_LIB}::swiglu_bwd", mutates_args=(), device_types="cuda")
    def swiglu_bwd(gu: torch.Tensor, dact: torch.Tensor) -> torch.Tensor:
        return ops.silu_mul_bwd(gu, dact.contiguous())

    @swiglu_bwd.register_fake
    def _(gu, dact):
        return torch.empty_like(gu)

    swiglu.register_autograd(lambda ctx, d: torch.ops.esb.swiglu_bwd(ctx.saved_tensors[0], d),
                             setup_context=lambda ctx, inputs, output: ctx.save_for_backward(inputs[0]))

    # ------------------------------------------------------------------ attention (packed q|k|v)
    @lib.custom_op(f"{_LIB}::attention_fwd", mutates_args=(), device_types="cuda")
    def attention_fwd(qkv: torch.Tensor, kmask: Optional[torch.Tensor], causal: bool) -> Tuple[torch.Tensor, torch.Tensor]:
        o, lse = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=causal, kmask=kmask, want_lse=True)
        return o, lse

    @attention_fwd.register_fake
    def _(qkv, kmask, causal):
        B, L, _, H, hd = qkv.shape
        return qkv.new_empty(B, L, H * hd), qkv.new_empty(B, H, L, dtype=torch.float32)

    @lib.custom_op(f"{_LIB}::attention_bwd", mutates_args=(), device_types="cuda")
    def attention_bwd(qkv: torch.Tensor, o: torch.Tensor, lse: torch.Tensor, do: torch.Tensor, kmask: Optional[torch.Tensor],
                      causal: bool) -> torch.Tensor:
        return ops.attention_bwd(qkv, o, lse, do.contiguous(), causal=causal, kmask=kmask)

    @attention_bwd.register_fake
    def _(qkv, o, lse, do, kmask, causal):
        return torch.empty_like(qkv)

    def attn_setup(ctx, inputs, output):
        qkv, kmask, causal = inputs
        ctx.save_for_backward(qkv, output[0], output[1], *([] if kmask is None else [kmask]))
        ctx.causal, ctx.has_mask = causal, kmask is not None

    def attn_bwd(ctx, do, _dlse):
        t = ctx.saved_tensors
        return torch.ops.esb.attention_bwd(t[0], t[1], t[2], do, t[3] if ctx.has_mask else None, ctx.causal), None, None

    attention_fwd.register_autograd(attn_bwd, setup_context=attn_setup)

    # ------------------------------------------------------------------ cross entropy (loss + its gradient in one pass)
    @lib.custom_op(f"{_LIB}::cross_entropy_fwd", mutates_args=(), device_types="cuda")
    def cross_entropy_fwd(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int) -> Tuple[torch.Tensor, torch.Tensor]:
        V = logits.shape[1]
        ld = (V + 7) // 8 * 8
        buf = torch.empty(logits.shape[0], ld, device=logits.device, dtype=torch.bfloat16)
        buf[:, :V].copy_(logits)
        loss = ops.cross_entropy(buf[:, :V], V, targets.contiguous(), ignore_index, write_grad=True)       # buf <- d loss / d logits
        return loss.reshape(()), buf[:, :V]

    @cross_entropy_fwd.register_fake
    def _(logits, targets, ignore_index):
        return logits.new_empty((), dtype=torch.float32), torch.empty_like(logits)

    def ce_setup(ctx, inputs, output):
        ctx.save_for_backward(output[1])

    def ce_bwd(ctx, dloss, _dgrad):
        (g,) = ctx.saved_tensors
        return (g.float() * dloss).to(g.dtype), None, None

    cross_entropy_fwd.register_autograd(ce_bwd, setup_context=ce_setup)

    # ------------------------------------------------------------------ trajectory tokens (integer: no gradient)
    @lib.custom_op(f"{_LIB}::traj_quantize", mutates_args=(), device_types="cuda")
    def traj_quantize(x: torch.Tensor, num_bins: int, id_offset: int) -> torch.Tensor:
        return ops.traj_quantize(x, num_bins, id_offset)

    @traj_quantize.register_fake
    def _(x, num_bins, id_offset):
        return x.new_empty(x.shape, dtype=torch.int64)

    @lib.custom_op(f"{_LIB}::traj_dequantize", mutates_args=(), device_types="cuda")
    def traj_dequantize(ids: torch.Tensor, num_bins: int, id_offset: int) -> torch.Tensor:
        return ops.traj_dequantize(ids, num_bins, id_offset)

    @traj_dequantize.register_fake
    def _(ids, num_bins, id_offset):
        return ids.new_empty(ids.shape, dtype=torch.float32)


_define()


# ---------------------------------------------------------------------------------------------------- convenience wrappers
def linear(x: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    return torch.ops.esb.linear(x, w)


def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float = 1e-6) -> torch.Tensor:
    return torch.ops.esb.rmsnorm_fwd(x, w, eps)[0]


def swiglu(gu: torch.Tensor) -> torch.Tensor:
    return torch.ops.esb.swiglu(gu)


def attention(qkv: torch.Tensor, kmask: Optional[torch.Tensor] = None, causal: bool = True) -> torch.Tensor:
    return torch.ops.esb.attention_fwd(qkv, kmask, causal)[0]


def cross_entropy(logits: torch.Tensor, targets: torch.Tensor, ignore_index: int = 0) -> torch.Tensor:
    return torch.ops.esb.cross_entropy_fwd(logits, targets, ignore_index)[0]
