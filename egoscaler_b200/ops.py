"""Tensor-level wrappers over the C ABI (``include/egoscaler_b200.h``).

Each function takes CUDA ``torch.Tensor``s, checks layout, and calls the matching ``esb_*`` entry point with raw
device pointers on torch's current stream.  Nothing here computes on the CPU and nothing falls back to ATen:
a missing library or a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import lib as _l

ACT_NONE, ACT_GELU, ACT_RELU = 0, 1, 2
gemm_timer = None      # bench.py installs an object with .before(M, N, K) / .after() to time the GEMM launches
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
SKINNY_MAX_M = 64
pdl_chain_enabled = False  # small kernels of the decode step on the programmatic launch chain too (esb_set_pdl_chain): measured 5.51 vs 5.29-5.45 ms/token off (tools/decode_variants.py), so off; the early launch_dependents triggers in those kernels stay (they let the PDL GEMMs prefetch)
# decode step with RMSNorm folded around the skinny GEMMs (llm._decode_step_folded: 7 kernels per layer instead of 9).  Measured
# round-robin against the plain step (tools/decode_variants.py, B=32, 100 tokens): 5.88 vs 5.90 ms/token - the norm kernels
# were already hidden behind the PDL weight prefetch, so it stays off and the plain step remains the parity-pinned path.
fold_norm_enabled = False
pdl_enabled = True        # programmatic dependent launch of the decode-step GEMMs (tools/decode_variants.py toggles it)
skinny_enabled = True
_SKINNY_WS_BYTES = 96 << 20        # counters (16 KiB, zero) + fp32 split-K partials; lm_head at 9 splits needs ~37 MB
_skinny_ws = {}


def _skinny_workspace(device) -> torch.Tensor:
    """Per-device workspace of esb_gemm_skinny_bf16 (zeroed once; the kernel leaves its counters at zero)."""
    key = (device.type, device.index if device.index is not None else torch.cuda.current_device())
    ws = _skinny_ws.get(key)
    if ws is None:
        ws = torch.zeros(_SKINNY_WS_BYTES, device=device, dtype=torch.uint8)
        _skinny_ws[key] = ws
    return ws


class TiledWeight:
    """A [N,K] bf16 weight re-laid out once into [ceil(N/128)][ceil(K/64)][128][64] tiles (esb_retile_weight_bf16): the
    layout the decode-step GEMM streams at HBM speed (one contiguous 16 KiB read per pipeline stage)."""
    __slots__ = ("data", "N", "K", "gateup")

    def __init__(self, data: torch.Tensor, N: int, K: int, gateup: bool = False):
        self.data, self.N, self.K, self.gateup = data, N, K, gateup


def retile_weight(w: torch.Tensor, col_scale: Optional[torch.Tensor] = None) -> TiledWeight:
    """``col_scale`` [K] bf16: the weight of the RMSNorm in front of this Linear, folded into its columns."""
    _chk(w, torch.bfloat16, "retile_weight.w")
    assert w.dim() == 2 and w.stride(1) == 1
    N, K = w.shape
    n = int(_l.lib().raw("esb_retiled_weight_elems")(N, K))
    out = torch.empty(n, device=w.device, dtype=torch.bfloat16)
    if col_scale is None:
        _l.call("esb_retile_weight_bf16", w.data_ptr(), w.stride(0), N, K, out.data_ptr(), _stream())
    else:
        _chk(col_scale, torch.bfloat16, "retile_weight.col_scale")
        assert col_scale.numel() == K and col_scale.is_contiguous()
        _l.call("esb_retile_weight_scaled_bf16", w.data_ptr(), w.stride(0), N, K, col_scale.data_ptr(), out.data_ptr(), _stream())
    return TiledWeight(out, N, K)


def retile_weight_gateup(w: torch.Tensor, col_scale: Optional[torch.Tensor] = None) -> TiledWeight:
    """The packed [2f, K] gate|up weight as tiles of 64 gate rows + the 64 matching up rows (esb_retile_weight_gateup_bf16):
    the layout whose accumulator tile carries both factors of SwiGLU (``gemm_chain`` with ``swiglu=True``)."""
    _chk(w, torch.bfloat16, "retile_weight_gateup.w")
    assert w.dim() == 2 and w.stride(1) == 1 and w.shape[0] % 2 == 0
    N, K = w.shape
    n = int(_l.lib().raw("esb_retiled_gateup_elems")(N // 2, K))
    out = torch.empty(n, device=w.device, dtype=torch.bfloat16)
    if col_scale is not None:
        _chk(col_scale, torch.bfloat16, "retile_weight_gateup.col_scale")
        assert col_scale.numel() == K and col_scale.is_contiguous()
    _l.call("esb_retile_weight_gateup_bf16", w.data_ptr(), w.stride(0), N // 2, K, _p(col_scale), out.data_ptr(), _stream())
    return TiledWeight(out, N, K, gateup=True)


class _ChainPhaseC(ctypes.Structure):          # esb_chain_phase (include/egoscaler_b200.h)
    _fields_ = [("X", ctypes.c_void_p), ("ldx", ctypes.c_int64), ("W_tiles", ctypes.c_void_p), ("C", ctypes.c_void_p),
                ("ldc", ctypes.c_int64), ("residual", ctypes.c_void_p), ("ldr", ctypes.c_int64), ("scale_src", ctypes.c_void_p),
                ("ssq_out", ctypes.c_void_p), ("N", ctypes.c_int32), ("K", ctypes.c_int32), ("scale_parts", ctypes.c_int32),
                ("swiglu", ctypes.c_int32), ("eps", ctypes.c_float), ("reserved", ctypes.c_int32)]


class ChainPhase:
    """One Linear of ``gemm_chain``: ``out = rstd * (x @ w.T) + residual`` (``swiglu``: ``out = silu(gate) * up``); arguments as
    in :func:`gemm_skinny`, ``w`` a :class:`TiledWeight` (``retile_weight_gateup`` for a SwiGLU phase)."""
    __slots__ = ("x", "w", "out", "residual", "row_scale", "eps", "ssq_out", "swiglu")

    def __init__(self, x, w, out, residual=None, row_scale=None, eps: float = 0.0, ssq_out=None, swiglu: bool = False):
        self.x, self.w, self.out, self.residual, self.row_scale, self.eps, self.ssq_out, self.swiglu = \
            x, w, out, residual, row_scale, eps, ssq_out, swiglu


_chain_ws = {}


def gemm_chain(phases, weights_const: bool = True) -> None:
    """Up to four dependent skinny Linears in ONE persistent launch (esb_gemm_chain_bf16): phase i+1 may read what phase i
    wrote.  The decode step runs o-proj -> gate|up + SwiGLU -> down-proj -> next q|k|v through it (llm._decode_step_chain)."""
    assert 1 <= len(phases) <= 4
    M = phases[0].x.shape[0]
    arr = (_ChainPhaseC * len(phases))()
    max_n = 0
    for c, ph in zip(arr, phases):
        _chk(ph.x, torch.bfloat16, "gemm_chain.x")
        _chk(ph.out, torch.bfloat16, "gemm_chain.out")
        assert isinstance(ph.w, TiledWeight) and ph.w.K == ph.x.shape[1] and ph.x.shape[0] == M and ph.x.stride(1) == 1
        assert ph.w.gateup == bool(ph.swiglu) and ph.out.stride(1) == 1
        assert tuple(ph.out.shape) == (M, ph.w.N // 2 if ph.swiglu else ph.w.N)
        sc, parts = (None, 0) if ph.row_scale is None else ph.row_scale
        if sc is not None:
            _chk(sc, torch.float32, "gemm_chain.row_scale")
            assert sc.is_contiguous() and sc.numel() >= (M if parts == 0 else parts * 64)
        if ph.ssq_out is not None:
            _chk(ph.ssq_out, torch.float32, "gemm_chain.ssq_out")
            assert ph.ssq_out.is_contiguous() and ph.ssq_out.numel() >= (ph.w.N + 127) // 128 * 64
        if ph.residual is not None:
            _chk(ph.residual, torch.bfloat16, "gemm_chain.residual")
            assert ph.residual.shape == ph.out.shape and ph.residual.stride(1) == 1
        c.X, c.ldx, c.W_tiles = ph.x.data_ptr(), ph.x.stride(0), ph.w.data.data_ptr()
        c.C, c.ldc = ph.out.data_ptr(), ph.out.stride(0)
        c.residual, c.ldr = _p(ph.residual), 0 if ph.residual is None else ph.residual.stride(0)
        c.scale_src, c.ssq_out = _p(sc), _p(ph.ssq_out)
        c.N, c.K, c.scale_parts, c.swiglu, c.eps, c.reserved = ph.w.N, ph.w.K, int(parts), int(bool(ph.swiglu)), float(ph.eps), 0
        max_n = max(max_n, ph.w.N)
    dev = phases[0].x.device
    key = (dev.type, dev.index if dev.index is not None else torch.cuda.current_device())
    need = int(_l.lib().raw("esb_gemm_chain_workspace_bytes")(64, max(max_n, 65536)))
    ws = _chain_ws.get(key)
    if ws is None or ws.numel() < need:
        ws = torch.zeros(need, device=dev, dtype=torch.uint8)
        _chain_ws[key] = ws
    _l.call("esb_gemm_chain_bf16", ctypes.addressof(arr), len(phases), M, ws.data_ptr(), ws.numel(),
            2 if (weights_const and pdl_enabled) else 0, _stream())


def gemm_skinny(a: torch.Tensor, w, residual: Optional[torch.Tensor] = None,
                out: Optional[torch.Tensor] = None, force_splits: int = 0, weights_const: bool = False,
                row_scale=None, eps: float = 0.0, ssq_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a @ w.T + residual`` for M <= 64 rows through the split-K weight-streaming kernel; ``w`` is a [N,K] bf16
    tensor or a :class:`TiledWeight`.  ``weights_const``: promise that ``w`` is not written by the kernel launched just
    before on this stream -> programmatic dependent launch with early weight prefetch (the decode step).
    ``row_scale`` = (fp32 tensor, parts): parts == 0 -> the tensor is rstd [M]; parts > 0 -> [parts, 64] partial sums of squares
    of ``a``'s rows (the ``ssq_out`` of the GEMM that produced ``a``), rstd = rsqrt(sum / K + eps).  ``ssq_out`` [ceil(N/128), 64]
    fp32 receives this GEMM's per-tile sums of squares of the bf16 output rows."""
    _chk(a, torch.bfloat16, "gemm_skinny.a")
    M, K = a.shape
    if isinstance(w, TiledWeight):
        assert w.K == K
        N, wp, ldw, tiled = w.N, w.data.data_ptr(), 0, 1
    else:
        _chk(w, torch.bfloat16, "gemm_skinny.w")
        N, wp, ldw, tiled = w.shape[0], w.data_ptr(), w.stride(0), 0
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=torch.bfloat16)
    ws = _skinny_workspace(a.device)
    if gemm_timer is not None:
        gemm_timer.before(M, N, K)
    flags = tiled | (2 if (weights_const and pdl_enabled) else 0)
    if row_scale is None and ssq_out is None:
        _l.call("esb_gemm_skinny_bf16", a.data_ptr(), a.stride(0), wp, ldw, out.data_ptr(), out.stride(0),
                M, N, K, _p(residual), 0 if residual is None else residual.stride(0), ws.data_ptr(), ws.numel(), force_splits,
                flags, _stream())
    else:
        sc, parts = (None, 0) if row_scale is None else row_scale
        if sc is not None:
            _chk(sc, torch.float32, "gemm_skinny.row_scale")
            assert sc.is_contiguous() and sc.numel() >= (M if parts == 0 else parts * 64)
        if ssq_out is not None:
            _chk(ssq_out, torch.float32, "gemm_skinny.ssq_out")
            assert ssq_out.is_contiguous() and ssq_out.numel() >= (N + 127) // 128 * 64
        _l.call("esb_gemm_skinny_normed_bf16", a.data_ptr(), a.stride(0), wp, ldw, out.data_ptr(), out.stride(0),
                M, N, K, _p(residual), 0 if residual is None else residual.stride(0), ws.data_ptr(), ws.numel(), force_splits,
                flags, _p(sc), int(parts), float(eps), _p(ssq_out), _stream())
    if gemm_timer is not None:
        gemm_timer.after()
    return out


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
    if (M <= SKINNY_MAX_M and skinny_enabled and bias is None and act == ACT_NONE and out.dtype == torch.bfloat16
            and not accumulate and alpha == 1.0 and force_bn == 0 and K >= 256):
        # cached single-token decode step: weight-streaming split-K kernel (gemm_skinny.cu)
        ws = _skinny_workspace(a.device)
        if gemm_timer is not None:
            gemm_timer.before(M, N, K)
        _l.call("esb_gemm_skinny_bf16", a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), out.data_ptr(), out.stride(0),
                M, N, K, _p(residual), ldr, ws.data_ptr(), ws.numel(), 0, 0, _stream())
        if gemm_timer is not None:
            gemm_timer.after()
        return out
    if gemm_timer is not None:
        gemm_timer.before(M, N, K)
    _l.call("esb_gemm_bf16", a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), out.data_ptr(), out.stride(0),
            M, N, K, _p(bias), _p(residual), ldr, act, 1 if out.dtype == torch.float32 else 0,
            1 if accumulate else 0, float(alpha), force_bn, _stream())
    if gemm_timer is not None:
        gemm_timer.after()
    return out


# Fused GEMM epilogues of the training / prefill step (llm.py switches on them).  Each one is bit-identical to the pair of
# launches it replaces (tests/test_gemm_gpu.py), so they can be toggled for A/B timing without touching parity.
def _env_flag(name: str, default: bool) -> bool:
    import os
    v = os.environ.get(name)
    return default if v is None else v not in ("0", "false", "False", "")


# decode step with the Linears between two attention calls as one persistent launch (llm._decode_step_chain, csrc/gemm_chain.cu);
# norms folded as above, SwiGLU in the gate|up epilogue: 3 kernels per layer instead of 9.  OFF (ESB_DECODE_CHAIN=1 enables it):
# the four Linears of a layer take 109 us in the chain against 108.5 us as four launches (tools/chain_bench.py) and the token step
# 5.95 against 5.49-5.61 ms (tools/decode_variants.py) - an in-kernel hand-off between dependent phases costs what a kernel
# boundary with programmatic launch costs: ~6 dependent L2 round trips of ~1.3 us (tools/chain_trace.py).
decode_chain_enabled = _env_flag("ESB_DECODE_CHAIN", False)
fuse_swiglu = _env_flag("ESB_FUSE_SWIGLU", True)        # gate|up GEMM + silu*mul            (esb_gemm_swiglu_bf16)
# d(act) GEMM + dSwiGLU (esb_gemm_dswiglu_bf16): bit-identical, but OFF - the epilogue has to read gate|up with one row per thread
# (TMEM lane = row), 32 lines per load instruction; measured 620 / 1003 / 1243 us per layer in three variants against 439-534 us
# for GEMM + silu_mul_bwd (profiles/r02_fused_bench_*.txt).  Needs the gate|up tile staged through shared memory to pay.
fuse_dswiglu = _env_flag("ESB_FUSE_DSWIGLU", False)
fuse_rope = _env_flag("ESB_FUSE_ROPE", True)            # q|k|v GEMM + rotary embedding       (esb_gemm_rope_bf16; head_dim 128 only)
top_layer_rows_enabled = _env_flag("ESB_TOP_LAYER_ROWS", True)   # training: top decoder layer on the loss rows only (llm._top_layer_rows)
fuse_rope_bwd = _env_flag("ESB_FUSE_ROPE_BWD", True)    # attention backward + transposed rotary embedding on dq / dk (esb_attn_bwd_rope)


def set_gemm_2cta(on: bool) -> bool:
    """Large GEMMs through the CTA-pair (cta_group::2, 256 x 256 tile) kernel; returns the previous setting."""
    return bool(_l.lib().raw("esb_set_gemm_2cta")(1 if on else 0))


def set_gemm_sm_reserve(k: int) -> int:
    """Leave ``k`` SMs out of the persistent GEMM grids (for NCCL while a gradient all-reduce is in flight); returns the previous k."""
    return int(_l.lib().raw("esb_set_gemm_sm_reserve")(int(k)))


def gemm_swiglu(a: torch.Tensor, wgu: torch.Tensor, want_gu: bool = True):
    """a [M,K], wgu [2f,K] = [gate ; up] -> (act [M,f], gu [M,2f] or None)."""
    _chk(a, torch.bfloat16, "gemm_swiglu.a"); _chk(wgu, torch.bfloat16, "gemm_swiglu.wgu")
    assert a.dim() == 2 and wgu.dim() == 2 and a.shape[1] == wgu.shape[1] and a.stride(1) == 1 and wgu.stride(1) == 1
    M, K = a.shape
    f = wgu.shape[0] // 2
    act = torch.empty(M, f, device=a.device, dtype=torch.bfloat16)
    gu = torch.empty(M, 2 * f, device=a.device, dtype=torch.bfloat16) if want_gu else None
    if gemm_timer is not None:
        gemm_timer.before(M, 2 * f, K)
    _l.call("esb_gemm_swiglu_bf16", a.data_ptr(), a.stride(0), wgu.data_ptr(), wgu.stride(0), act.data_ptr(), act.stride(0),
            _p(gu), 0 if gu is None else gu.stride(0), M, f, K, _stream())
    if gemm_timer is not None:
        gemm_timer.after()
    return act, gu


def gemm_dswiglu(dout: torch.Tensor, wd: torch.Tensor, gu: torch.Tensor, w_is_down_proj: bool = False) -> torch.Tensor:
    """dout [M,d], gu [M,2f] -> dgu [M,2f] = dSwiGLU(gu, dout . down_proj).  ``wd`` = down_proj^T [f,d], or with
    ``w_is_down_proj`` the parameter itself [d,f] (no contraction-major copy needed)."""
    _chk(dout, torch.bfloat16, "gemm_dswiglu.dout"); _chk(wd, torch.bfloat16, "gemm_dswiglu.wd"); _chk(gu, torch.bfloat16, "gemm_dswiglu.gu")
    M, K = dout.shape
    f = wd.shape[1] if w_is_down_proj else wd.shape[0]
    assert wd.shape == ((K, f) if w_is_down_proj else (f, K)) and gu.shape == (M, 2 * f)
    assert gu.stride(1) == 1 and dout.stride(1) == 1 and wd.stride(1) == 1
    dgu = torch.empty(M, 2 * f, device=dout.device, dtype=torch.bfloat16)
    if gemm_timer is not None:
        gemm_timer.before(M, f, K)
    _l.call("esb_gemm_dswiglu_bf16", dout.data_ptr(), dout.stride(0), wd.data_ptr(), wd.stride(0), 1 if w_is_down_proj else 0,
            gu.data_ptr(), gu.stride(0), dgu.data_ptr(), dgu.stride(0), M, f, K, _stream())
    if gemm_timer is not None:
        gemm_timer.after()
    return dgu


def gemm_rope(a: torch.Tensor, w: torch.Tensor, cos_tab, sin_tab, L: int, pos0: int, rope_cols: int, hd: int) -> torch.Tensor:
    """a [M,K] . w [N,K]^T with rotary embedding applied to the 128-wide heads below column ``rope_cols``."""
    _chk(a, torch.bfloat16, "gemm_rope.a"); _chk(w, torch.bfloat16, "gemm_rope.w")
    M, K = a.shape
    N = w.shape[0]
    assert w.shape[1] == K and a.stride(1) == 1 and w.stride(1) == 1 and cos_tab.shape[1] == 64 and cos_tab.is_contiguous()
    out = torch.empty(M, N, device=a.device, dtype=torch.bfloat16)
    if gemm_timer is not None:
        gemm_timer.before(M, N, K)
    _l.call("esb_gemm_rope_bf16", a.data_ptr(), a.stride(0), w.data_ptr(), w.stride(0), out.data_ptr(), out.stride(0), M, N, K,
            cos_tab.data_ptr(), sin_tab.data_ptr(), int(L), int(pos0), int(rope_cols), int(hd), _stream())
    if gemm_timer is not None:
        gemm_timer.after()
    return out


gemm_tn_enabled = True    # weight gradients dY^T . X through the TN tcgen05 GEMM (no transposed activation copies)
gemm_nn_enabled = True    # data gradients dY . W on the [out, in] weight itself (no contraction-major weight copies)


def _gemm_t(a, trans_a, b, trans_b, M, N, Kc, out, out_dtype, accumulate, alpha):
    if out is None:
        out = torch.empty(M, N, device=a.device, dtype=out_dtype)
    assert out.shape == (M, N) and out.stride(1) == 1
    if gemm_timer is not None:
        gemm_timer.before(M, N, Kc)
    _l.call("esb_gemm_t_bf16", a.data_ptr(), a.stride(0), 1 if trans_a else 0, b.data_ptr(), b.stride(0), 1 if trans_b else 0,
            out.data_ptr(), out.stride(0), M, N, Kc, 1 if out.dtype == torch.float32 else 0, 1 if accumulate else 0, float(alpha), _stream())
    if gemm_timer is not None:
        gemm_timer.after()
    return out


def gemm_tn(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.float32,
            accumulate: bool = False, alpha: float = 1.0) -> torch.Tensor:
    """``alpha * a.T @ b`` for row-major a [T, M], b [T, N] (bf16; row strides multiples of 8): dW = dY^T . X."""
    _chk(a, torch.bfloat16, "gemm_tn.a"); _chk(b, torch.bfloat16, "gemm_tn.b")
    assert a.dim() == 2 and b.dim() == 2 and a.shape[0] == b.shape[0] and a.stride(1) == 1 and b.stride(1) == 1
    if not gemm_tn_enabled:
        return gemm(transpose(a), transpose(b), out=out, out_dtype=out_dtype, accumulate=accumulate, alpha=alpha)
    return _gemm_t(a, True, b, True, a.shape[1], b.shape[1], a.shape[0], out, out_dtype, accumulate, alpha)


def gemm_nn(a: torch.Tensor, w: torch.Tensor, out: Optional[torch.Tensor] = None, out_dtype: torch.dtype = torch.bfloat16,
            wT: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``a @ w`` for a [M, Kc], w [Kc, N] row-major (a weight in its [out, in] state_dict layout): dX = dY . W.
    ``wT`` = an existing [N, Kc] copy of w to use when the NN kernel is switched off."""
    _chk(a, torch.bfloat16, "gemm_nn.a"); _chk(w, torch.bfloat16, "gemm_nn.w")
    assert a.dim() == 2 and w.dim() == 2 and a.shape[1] == w.shape[0] and a.stride(1) == 1 and w.stride(1) == 1
    if not gemm_nn_enabled:
        return gemm(a, wT if wT is not None else transpose(w), out=out, out_dtype=out_dtype)
    return _gemm_t(a, False, w, True, a.shape[0], w.shape[1], a.shape[1], out, out_dtype, False, 1.0)


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


# ------------------------------------------------------------------------------------------------ norms
def rmsnorm(x: torch.Tensor, w: torch.Tensor, eps: float, want_rstd: bool = False):
    """HF LlamaRMSNorm (modeling_llama.py:53-67). x [rows, d] bf16 (row stride multiple of 8), w [d] bf16."""
    _chk(x, torch.bfloat16, "rmsnorm.x"); _chk(w, torch.bfloat16, "rmsnorm.w")
    assert x.dim() == 2 and x.stride(1) == 1
    rows, d = x.shape
    y = torch.empty(rows, d, device=x.device, dtype=torch.bfloat16)
    rstd = torch.empty(rows, device=x.device, dtype=torch.float32) if want_rstd else None
    _l.call("esb_rmsnorm_fwd", x.data_ptr(), x.stride(0), w.data_ptr(), y.data_ptr(), d, _p(rstd), rows, d, float(eps), _stream())
    return (y, rstd) if want_rstd else y


def rmsnorm_bwd(dy, x, w, rstd, dres=None, dw: Optional[torch.Tensor] = None):
    for t, n in ((dy, "dy"), (x, "x"), (w, "w")):
        _chk(t, torch.bfloat16, "rmsnorm_bwd." + n)
    assert dy.is_contiguous() and x.is_contiguous()
    rows, d = x.shape
    dx = torch.empty_like(x)
    _l.call("esb_rmsnorm_bwd", dy.data_ptr(), x.data_ptr(), w.data_ptr(), rstd.data_ptr(), _p(dres), dx.data_ptr(), _p(dw),
            rows, d, _stream())
    return dx


def layernorm(x, gamma, beta, eps: float = 1e-5, add=None, sum_out=None):
    """nn.LayerNorm over the last dim of x [rows, d] bf16; optional fused ``s = x + add`` (sum written to sum_out)."""
    _chk(x, torch.bfloat16, "layernorm.x"); _chk(gamma, torch.float32, "layernorm.gamma"); _chk(beta, torch.float32, "layernorm.beta")
    assert x.is_contiguous()
    rows, d = x.shape
    y = torch.empty_like(x)
    period = 0
    if add is not None:
        _chk(add, torch.bfloat16, "layernorm.add")
        assert add.is_contiguous() and add.shape[-1] == d
        period = add.numel() // d
    _l.call("esb_layernorm_fwd", x.data_ptr(), _p(add), period, gamma.data_ptr(), beta.data_ptr(), _p(sum_out), y.data_ptr(),
            rows, d, float(eps), _stream())
    return y


# ------------------------------------------------------------------------------------------------ element-wise
def rope_(buf: torch.Tensor, L: int, n_heads: int, hd: int, cos_tab, sin_tab, pos0: int = 0, inverse: bool = False,
          kcache=None, vcache=None, kv_head0: int = 0, kv_heads: int = 0, v_col0: int = 0, cache_len: int = 0, pos_ptr=None):
    _chk(buf, torch.bfloat16, "rope.buf")
    assert buf.dim() == 2 and buf.stride(1) == 1
    _l.call("esb_rope", buf.data_ptr(), buf.stride(0), buf.shape[0], L, n_heads, hd, cos_tab.data_ptr(), sin_tab.data_ptr(),
            pos0, 1 if inverse else 0, _p(kcache), _p(vcache), kv_head0, kv_heads, v_col0, cache_len, _p(pos_ptr), _stream())
    return buf


def silu_mul(gu: torch.Tensor) -> torch.Tensor:
    _chk(gu, torch.bfloat16, "silu_mul.gu")
    assert gu.is_contiguous()
    rows, f2 = gu.shape
    act = torch.empty(rows, f2 // 2, device=gu.device, dtype=torch.bfloat16)
    _l.call("esb_silu_mul_fwd", gu.data_ptr(), act.data_ptr(), rows, f2 // 2, _stream())
    return act


def silu_mul_bwd(gu: torch.Tensor, dact: torch.Tensor) -> torch.Tensor:
    assert gu.is_contiguous() and dact.is_contiguous()
    dgu = torch.empty_like(gu)
    _l.call("esb_silu_mul_bwd", gu.data_ptr(), dact.data_ptr(), dgu.data_ptr(), gu.shape[0], gu.shape[1] // 2, _stream())
    return dgu


def gelu(z: torch.Tensor) -> torch.Tensor:
    _chk(z, torch.bfloat16, "gelu.z")
    assert z.is_contiguous()
    y = torch.empty_like(z)
    _l.call("esb_gelu", z.data_ptr(), None, y.data_ptr(), z.numel(), 0, _stream())
    return y


def gelu_bwd(z: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    assert z.is_contiguous() and dy.is_contiguous()
    dz = torch.empty_like(z)
    _l.call("esb_gelu", z.data_ptr(), dy.data_ptr(), dz.data_ptr(), z.numel(), 1, _stream())
    return dz


# ------------------------------------------------------------------------------------------------ gather / layout
def point_layout(ids: torch.Tensor, patch_id: int, start_id: int, end_id: int, n_point: int) -> torch.Tensor:
    _chk(ids, torch.int64, "point_layout.ids")
    ids = ids.contiguous()
    B, L = ids.shape
    info = torch.empty(B, 8, device=ids.device, dtype=torch.int32)
    _l.call("esb_point_layout", ids.data_ptr(), B, L, patch_id, start_id, end_id, n_point, info.data_ptr(), _stream())
    return info


def embed_splice(ids, embed, feats, splice_pos, n_point: int) -> torch.Tensor:
    _chk(ids, torch.int64, "embed_splice.ids"); _chk(embed, torch.bfloat16, "embed_splice.embed")
    ids = ids.contiguous()
    B, L = ids.shape
    d = embed.shape[1]
    x = torch.empty(B * L, d, device=ids.device, dtype=torch.bfloat16)
    _l.call("esb_embed_splice_fwd", ids.data_ptr(), embed.data_ptr(), _p(feats), _p(splice_pos), n_point, x.data_ptr(), B, L, d, embed.shape[0], _stream())
    return x


def embed_splice_bwd(ids, dx, splice_pos, n_point: int, dembed, dfeats, padding_idx: int, vocab: Optional[int] = None):
    ids = ids.contiguous()
    B, L = ids.shape
    d = dx.shape[-1]
    assert dx.is_contiguous()
    _l.call("esb_embed_splice_bwd", ids.data_ptr(), dx.data_ptr(), _p(splice_pos), n_point, _p(dembed), _p(dfeats), padding_idx,
            B, L, d, int(vocab if vocab is not None else (dembed.shape[0] if dembed is not None else 1 << 62)), _stream())


def embed_rows_pack(ids, dx, splice_pos, n_point: int, rmax: int, pad_id: int):
    """Non-point rows of dx [B*L, d] -> (rows [B*rmax, d] bf16, ids [B*rmax] int64; unused slots carry pad_id)."""
    ids = ids.contiguous()
    B, L = ids.shape
    d = dx.shape[-1]
    assert dx.is_contiguous()
    rows = torch.empty(B * rmax, d, device=dx.device, dtype=torch.bfloat16)
    rid = torch.empty(B * rmax, device=dx.device, dtype=torch.int64)
    _l.call("esb_embed_rows_pack", ids.data_ptr(), dx.data_ptr(), _p(splice_pos), n_point, rmax, rows.data_ptr(), rid.data_ptr(),
            pad_id, B, L, d, _stream())
    return rows, rid


def embed_scatter_rows(ids: torch.Tensor, rows: torch.Tensor, dembed: torch.Tensor, pad_id: int):
    """dembed[id] += sum of rows with that id, in increasing row order (deterministic; ids [R] int64, rows [R, d] bf16)."""
    _chk(ids, torch.int64, "embed_scatter_rows.ids"); _chk(rows, torch.bfloat16, "embed_scatter_rows.rows"); _chk(dembed, torch.float32, "embed_scatter_rows.dembed")
    assert ids.is_contiguous() and rows.is_contiguous() and dembed.is_contiguous() and rows.shape[0] == ids.numel()
    _l.call("esb_embed_scatter_rows", ids.data_ptr(), rows.data_ptr(), dembed.data_ptr(), pad_id, dembed.shape[0], ids.numel(),
            rows.shape[1], _stream())


def embed_scatter_sorted(ids: torch.Tensor, rows: torch.Tensor, dembed: torch.Tensor, pad_id: int):
    """``embed_scatter_rows`` (same sums, same order) through a stable sort of the ids instead of the O(R^2) scan; any R."""
    _chk(ids, torch.int64, "embed_scatter_sorted.ids"); _chk(rows, torch.bfloat16, "embed_scatter_sorted.rows"); _chk(dembed, torch.float32, "embed_scatter_sorted.dembed")
    assert ids.is_contiguous() and rows.is_contiguous() and dembed.is_contiguous() and rows.shape[0] == ids.numel()
    sorted_ids, order = torch.sort(ids, stable=True)             # index bookkeeping only; the sums run in esb_embed_scatter_sorted
    _l.call("esb_embed_scatter_sorted", sorted_ids.data_ptr(), order.data_ptr(), rows.data_ptr(), dembed.data_ptr(), pad_id,
            dembed.shape[0], ids.numel(), rows.shape[1], _stream())


def copy_rows(src: torch.Tensor, dst: torch.Tensor):
    """dst[b, r, :] = src[b, r, :] for 3-D views whose last dim is contiguous (any batch/row strides)."""
    assert src.dim() == 3 and dst.shape == src.shape and src.stride(2) == 1 and dst.stride(2) == 1 and src.dtype == dst.dtype
    es = src.element_size()
    Bn, R, Cn = src.shape
    _l.call("esb_copy_rows", src.data_ptr(), src.stride(0) * es, src.stride(1) * es, dst.data_ptr(), dst.stride(0) * es,
            dst.stride(1) * es, Bn, R, Cn * es, _stream())
    return dst


def transpose(x: torch.Tensor, out: Optional[torch.Tensor] = None, pad_to: int = 8) -> torch.Tensor:
    """[R, C] bf16 -> [C, R] (leading dimension padded to a multiple of ``pad_to``; returns the [C, R] view)."""
    _chk(x, torch.bfloat16, "transpose.x")
    assert x.dim() == 2 and x.stride(1) == 1
    R, C = x.shape
    if out is None:
        ld = (R + pad_to - 1) // pad_to * pad_to
        out = torch.zeros(C, ld, device=x.device, dtype=torch.bfloat16)[:, :R] if ld != R else \
            torch.empty(C, R, device=x.device, dtype=torch.bfloat16)
    _l.call("esb_transpose_bf16", x.data_ptr(), x.stride(0), out.data_ptr(), out.stride(0), R, C, _stream())
    return out


def colsum(x: torch.Tensor, out: torch.Tensor):
    _chk(x, torch.bfloat16, "colsum.x"); _chk(out, torch.float32, "colsum.out")
    M, N = x.shape
    _l.call("esb_colsum_bf16", x.data_ptr(), x.stride(0), M, N, out.data_ptr(), _stream())
    return out


def group_maxpool(f: torch.Tensor, n: int, concat: bool) -> torch.Tensor:
    _chk(f, torch.bfloat16, "group_maxpool.f")
    assert f.is_contiguous()
    rows, C = f.shape
    groups = rows // n
    out = torch.empty((rows, 2 * C) if concat else (groups, C), device=f.device, dtype=torch.bfloat16)
    _l.call("esb_group_maxpool", f.data_ptr(), out.data_ptr(), groups, n, C, 1 if concat else 0, _stream())
    return out


# ------------------------------------------------------------------------------------------------ loss / sampling / optimizer
def cross_entropy(logits: torch.Tensor, V: int, targets: torch.Tensor, ignore_index: int, write_grad: bool):
    """Mean CE over non-ignored rows (train.py:177-181). logits [rows, ld>=V] bf16 are overwritten with dlogits when
    ``write_grad``.  Returns a 1-element fp32 tensor."""
    _chk(logits, torch.bfloat16, "cross_entropy.logits"); _chk(targets, torch.int64, "cross_entropy.targets")
    assert logits.dim() == 2 and logits.stride(1) == 1 and targets.is_contiguous()
    rows = logits.shape[0]
    row_loss = torch.empty(rows, device=logits.device, dtype=torch.float32)
    scratch = torch.empty(1, device=logits.device, dtype=torch.float32)
    loss = torch.empty(1, device=logits.device, dtype=torch.float32)
    _l.call("esb_cross_entropy", logits.data_ptr(), logits.stride(0), rows, V, targets.data_ptr(), ignore_index,
            row_loss.data_ptr(), scratch.data_ptr(), loss.data_ptr(), 1 if write_grad else 0, _stream())
    return loss


def argmax_rows(logits: torch.Tensor, V: int, scores: Optional[torch.Tensor] = None, suppress_col: int = -1,
                out: Optional[torch.Tensor] = None, step_ptr=None, step_off: int = 0, scores_step_stride: int = 0) -> torch.Tensor:
    _chk(logits, torch.bfloat16, "argmax_rows.logits")
    rows = logits.shape[0]
    if out is None:
        out = torch.empty(rows, device=logits.device, dtype=torch.int64)
    _l.call("esb_argmax_rows", logits.data_ptr(), logits.stride(0), rows, V, out.data_ptr(), _p(scores),
            0 if scores is None else scores.stride(-2), suppress_col, _p(step_ptr), step_off, scores_step_stride, _stream())
    return out


def sample_rows(logits: torch.Tensor, V: int, u: torch.Tensor, temperature: float = 1.0, top_k: int = 50, top_p: float = 0.95,
                penalty: float = 1.0, hist: Optional[torch.Tensor] = None, hist_len: int = 0, pos_ptr=None,
                scores: Optional[torch.Tensor] = None, suppress_col: int = -1, out: Optional[torch.Tensor] = None,
                step_ptr=None, step_off: int = 0, scores_step_stride: int = 0, u_step_stride: int = 0,
                n_kept: Optional[torch.Tensor] = None) -> torch.Tensor:
    """HF sampling chain (repetition penalty -> temperature -> top-k -> top-p -> draw) on bf16 logits [rows, >=V];
    ``u`` fp32 uniforms ([rows], or [steps, rows] with ``u_step_stride`` and a step counter)."""
    _chk(logits, torch.bfloat16, "sample_rows.logits")
    _chk(u, torch.float32, "sample_rows.u")
    rows = logits.shape[0]
    if out is None:
        out = torch.empty(rows, device=logits.device, dtype=torch.int64)
    if hist is not None:
        _chk(hist, torch.int64, "sample_rows.hist")
    _l.call("esb_sample_rows", logits.data_ptr(), logits.stride(0), rows, V, float(temperature), int(top_k), float(top_p),
            float(penalty), _p(hist), 0 if hist is None else hist.stride(0), int(hist_len), _p(pos_ptr), u.data_ptr(),
            int(u_step_stride), _p(step_ptr), int(step_off), _p(scores), 0 if scores is None else scores.stride(-2),
            int(scores_step_stride), int(suppress_col), out.data_ptr(), _p(n_kept), _stream())
    return out


def greedy_update(nxt, tokens_out, unfinished, sequences, cur_len: int, mask, eos: int, pad: int, pos_ptr=None):
    B = nxt.shape[0]
    _l.call("esb_greedy_update", nxt.data_ptr(), tokens_out.data_ptr(), unfinished.data_ptr(), sequences.data_ptr(),
            sequences.stride(0), cur_len, _p(mask), 0 if mask is None else mask.stride(0), eos, pad, B, _p(pos_ptr), _stream())


def adamw_(p, g, m, v, p_bf16, lr, beta1, beta2, eps, wd, step: int, grad_scale: float = 1.0):
    for t, n in ((p, "p"), (g, "g"), (m, "m"), (v, "v")):
        _chk(t, torch.float32, "adamw." + n)
        assert t.is_contiguous()
    _l.call("esb_adamw", p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _p(p_bf16), p.numel(), float(lr), float(beta1),
            float(beta2), float(eps), float(wd), int(step), float(grad_scale), _stream())


def cast_bf16(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    _chk(x, torch.float32, "cast_bf16.x")
    x = x.contiguous()
    if out is None:
        out = torch.empty(x.shape, device=x.device, dtype=torch.bfloat16)
    _l.call("esb_cast_f32_bf16", x.data_ptr(), out.data_ptr(), x.numel(), _stream())
    return out


# ------------------------------------------------------------------------------------------------ attention
def _bls(t: torch.Tensor):
    """(batch stride, row stride) of a [B, L, H, hd] view whose (H, hd) block is contiguous."""
    assert t.dim() == 4 and t.stride(3) == 1 and t.stride(2) == t.shape[3]
    return t.stride(0), t.stride(1)


def attention(q, k, v, causal: bool, kmask: Optional[torch.Tensor] = None, q_pos0: int = 0, scale: Optional[float] = None,
              want_lse: bool = False, out: Optional[torch.Tensor] = None):
    """q [B,Lq,H,hd], k/v [B,Lk,H,hd] bf16 views (strided OK) -> o [B,Lq,H*hd] (+ lse [B,H,Lq])."""
    for t, n in ((q, "q"), (k, "k"), (v, "v")):
        _chk(t, torch.bfloat16, "attention." + n)
    B, Lq, H, hd = q.shape
    Lk = k.shape[1]
    if out is None:
        out = torch.empty(B, Lq, H * hd, device=q.device, dtype=torch.bfloat16)
    lse = torch.empty(B, H, Lq, device=q.device, dtype=torch.float32) if want_lse else None
    if kmask is not None:
        assert kmask.dtype == torch.uint8 and kmask.stride(1) == 1 and kmask.shape[1] >= Lk
    qs, ks, vs = _bls(q), _bls(k), _bls(v)
    _l.call("esb_attn_fwd", q.data_ptr(), qs[0], qs[1], k.data_ptr(), ks[0], ks[1], v.data_ptr(), vs[0], vs[1],
            out.data_ptr(), out.stride(0), out.stride(1), _p(lse), _p(kmask), 0 if kmask is None else kmask.stride(0),
            B, H, Lq, Lk, hd, 1 if causal else 0, q_pos0, float(scale if scale is not None else hd ** -0.5), _stream())
    return (out, lse) if want_lse else out


def attention_decode(q, kcache, vcache, H: int, hd: int, kmask=None, len_ptr=None, len_static: int = 0,
                     scale: Optional[float] = None, out: Optional[torch.Tensor] = None):
    """q [B, >=H*hd] (row stride any), caches [B, Lmax, H*hd] -> o [B, H*hd]."""
    B = q.shape[0]
    if out is None:
        out = torch.empty(B, H * hd, device=q.device, dtype=torch.bfloat16)
    _l.call("esb_attn_decode", q.data_ptr(), q.stride(0), kcache.data_ptr(), vcache.data_ptr(), kcache.stride(0), kcache.stride(1),
            out.data_ptr(), out.stride(0), _p(kmask), 0 if kmask is None else kmask.stride(0), _p(len_ptr), len_static, B, H, hd,
            float(scale if scale is not None else hd ** -0.5), _stream())
    return out


# ------------------------------------------------------------------------------------------------ point cloud
def fps(pts: torch.Tensor, G: int, start: torch.Tensor, pad_ld: int = 64):
    """-> (centers [B,G,3] fp32, center_idx [B,G] int32, centers_pad [B*G, pad_ld] bf16)."""
    _chk(pts, torch.float32, "fps.pts"); _chk(start, torch.int64, "fps.start")
    pts = pts.contiguous()
    B, N, C = pts.shape
    centers = torch.empty(B, G, 3, device=pts.device, dtype=torch.float32)
    idx = torch.empty(B, G, device=pts.device, dtype=torch.int32)
    cpad = torch.empty(B * G, pad_ld, device=pts.device, dtype=torch.bfloat16)
    _l.call("esb_fps", pts.data_ptr(), B, N, C, G, start.data_ptr(), centers.data_ptr(), idx.data_ptr(), cpad.data_ptr(),
            pad_ld, _stream())
    return centers, idx, cpad


def knn_group(pts: torch.Tensor, centers: torch.Tensor, K: int, out_ld: int = 64, want_idx: bool = False):
    pts = pts.contiguous()
    B, N, C = pts.shape
    G = centers.shape[1]
    nb = torch.empty(B * G * K, out_ld, device=pts.device, dtype=torch.bfloat16)
    idx = torch.empty(B, G, K, device=pts.device, dtype=torch.int32) if want_idx else None
    _l.call("esb_knn_group", pts.data_ptr(), centers.data_ptr(), B, N, C, G, K, out_ld, nb.data_ptr(), _p(idx), _stream())
    return (nb, idx) if want_idx else nb


def attention_bwd(qkv: torch.Tensor, o: torch.Tensor, lse: torch.Tensor, dout: torch.Tensor, causal: bool,
                  kmask: Optional[torch.Tensor] = None, scale: Optional[float] = None, rope=None, q_rows=None) -> torch.Tensor:
    """Self-attention backward on a packed projection: qkv [B,L,3,H,hd], o/dout [B,L,H*hd], lse [B,H,L]
    -> dqkv [B,L,3,H,hd] (gradients w.r.t. the post-RoPE q,k and v)."""
    _chk(qkv, torch.bfloat16, "attention_bwd.qkv"); _chk(dout, torch.bfloat16, "attention_bwd.dout")
    B, L, three, H, hd = qkv.shape
    assert three == 3 and qkv.is_contiguous() and o.stride(2) == 1 and dout.stride(2) == 1
    # q_rows = (r0, r1): only these query rows of every sample took part in the forward (o / lse / dout have r1 - r0 rows); dq of the
    # other rows is zero, dk / dv cover all L keys
    r0, r1 = (0, L) if q_rows is None else (int(q_rows[0]), int(q_rows[1]))
    Lq = r1 - r0
    assert o.shape[1] == Lq and dout.shape[1] == Lq and lse.shape[2] == Lq
    dqkv = torch.empty_like(qkv) if q_rows is None else torch.zeros_like(qkv)
    dvec = torch.empty(2 * B * H * ((Lq + 63) // 64 * 64), device=qkv.device, dtype=torch.float32)   # (lse, D) workspace
    bs, rs = qkv.stride(0), qkv.stride(1)
    es = qkv.element_size()
    d = H * hd
    qoff = r0 * rs * es
    args = (qkv.data_ptr() + qoff, bs, rs, qkv.data_ptr() + d * es, bs, rs, qkv.data_ptr() + 2 * d * es, bs, rs,
            o.data_ptr(), o.stride(0), o.stride(1), dout.data_ptr(), dout.stride(0), dout.stride(1), lse.data_ptr(),
            dvec.data_ptr(), dqkv.data_ptr() + qoff, bs, rs, dqkv.data_ptr() + d * es, bs, rs, dqkv.data_ptr() + 2 * d * es, bs, rs,
            _p(kmask), 0 if kmask is None else kmask.stride(0), B, H, Lq, L, hd, 1 if causal else 0, r0,
            float(scale if scale is not None else hd ** -0.5))
    if rope is None:
        _l.call("esb_attn_bwd", *args, _stream())
    else:                   # (cos, sin) tables: dq / dk come back w.r.t. the pre-rotation projections (fused inverse RoPE)
        cos_tab, sin_tab = rope
        assert hd == 128 and cos_tab.shape[1] == 64 and cos_tab.shape[0] >= L and cos_tab.is_contiguous() and sin_tab.is_contiguous()
        _l.call("esb_attn_bwd_rope", *args, cos_tab.data_ptr(), sin_tab.data_ptr(), _stream())
    return dqkv


# ------------------------------------------------------------------------------------------------ trainable point encoder
def layernorm_bwd(dy, x, gamma, dgamma=None, dbeta=None, dres=None, eps: float = 1e-5):
    """nn.LayerNorm backward; ``x`` is the LayerNorm input.  Returns dx (+ dres); dgamma / dbeta fp32 are accumulated."""
    _chk(dy, torch.bfloat16, "layernorm_bwd.dy"); _chk(x, torch.bfloat16, "layernorm_bwd.x"); _chk(gamma, torch.float32, "layernorm_bwd.gamma")
    assert dy.is_contiguous() and x.is_contiguous()
    rows, d = x.shape
    dx = torch.empty_like(x)
    _l.call("esb_layernorm_bwd", dy.data_ptr(), x.data_ptr(), gamma.data_ptr(), _p(dres), dx.data_ptr(), _p(dgamma), _p(dbeta),
            rows, d, float(eps), _stream())
    return dx


def colsum2(a, b, out0, out1):
    """out0[c] += sum_r a[r,c]; out1[c] += sum_r a[r,c] * b[r,c] (b None: a*a)."""
    _chk(a, torch.bfloat16, "colsum2.a")
    M, N = a.shape
    _l.call("esb_colsum2_bf16", a.data_ptr(), a.stride(0), _p(b), 0 if b is None else b.stride(0), M, N, out0.data_ptr(), _p(out1),
            _stream())


def bn_coeffs(mode: int, sum0, sum1, R: int, gamma, beta, mean, rstd, A, B, C, running_mean=None, running_var=None,
              dgamma=None, dbeta=None, eps: float = 1e-5, momentum: float = 0.1):
    _l.call("esb_bn_coeffs", mode, sum0.data_ptr(), sum1.data_ptr(), R, gamma.numel(), gamma.data_ptr(), _p(beta), float(eps),
            float(momentum), mean.data_ptr(), rstd.data_ptr(), _p(running_mean), _p(running_var), A.data_ptr(), _p(B), C.data_ptr(),
            _p(dgamma), _p(dbeta), _stream())


def affine_cols(x, A, C, y=None, B=None, relu: bool = False, out=None):
    """out[r,c] = act(A[c]*x[r,c] + B[c]*y[r,c] + C[c])."""
    _chk(x, torch.bfloat16, "affine_cols.x")
    assert x.is_contiguous() and (y is None or y.is_contiguous())
    rows, Cn = x.shape
    if out is None:
        out = torch.empty_like(x)
    _l.call("esb_affine_cols", x.data_ptr(), _p(y), A.data_ptr(), _p(B), C.data_ptr(), out.data_ptr(), rows, Cn, 1 if relu else 0,
            _stream())
    return out


def relu_mask(dy, y):
    _chk(dy, torch.bfloat16, "relu_mask.dy"); _chk(y, torch.bfloat16, "relu_mask.y")
    assert dy.is_contiguous() and y.is_contiguous() and dy.shape == y.shape
    out = torch.empty_like(dy)
    _l.call("esb_relu_mask", dy.data_ptr(), y.data_ptr(), out.data_ptr(), dy.numel(), _stream())
    return out


def group_maxpool_bwd(f, n: int, dpool=None, dcat=None):
    """Backward of ops.group_maxpool: f [G*n, C]; dpool [G, C] (plain) or dcat [G*n, 2C] (concat form) -> df [G*n, C]."""
    _chk(f, torch.bfloat16, "group_maxpool_bwd.f")
    assert f.is_contiguous()
    R, C = f.shape
    df = torch.empty_like(f)
    for t in (dpool, dcat):
        if t is not None:
            _chk(t, torch.bfloat16, "group_maxpool_bwd.d")
            assert t.is_contiguous()
    _l.call("esb_group_maxpool_bwd", f.data_ptr(), _p(dpool), _p(dcat), df.data_ptr(), R // n, n, C, _stream())
    return df


def scale_rows_add(y, scale, rows_per_batch: int, x=None):
    """DropPath: x + scale[row // rows_per_batch] * y (x None: scale * y)."""
    _chk(y, torch.bfloat16, "scale_rows_add.y"); _chk(scale, torch.float32, "scale_rows_add.scale")
    assert y.is_contiguous() and (x is None or x.is_contiguous())
    rows, d = y.shape
    out = torch.empty_like(y)
    _l.call("esb_scale_rows_add", _p(x), y.data_ptr(), scale.data_ptr(), out.data_ptr(), rows, d, rows_per_batch, _stream())
    return out


# ------------------------------------------------------------------------------------------------ metrics
def traj_metrics(gen: torch.Tensor, gt: torch.Tensor, gen_len: Optional[torch.Tensor] = None, gt_len: Optional[torch.Tensor] = None,
                 cols=(0, 6), rot_col: int = 3) -> torch.Tensor:
    """models/utils/metrics.py per sample: gen [B,Tg,D], gt [B,Tt,D] fp32 -> [B,4] fp64 (ADE, FDE, IDE, angular distance)."""
    _chk(gen, torch.float32, "traj_metrics.gen"); _chk(gt, torch.float32, "traj_metrics.gt")
    gen, gt = gen.contiguous(), gt.contiguous()
    B, Tg, D = gen.shape
    Tt = gt.shape[1]
    for t in (gen_len, gt_len):
        if t is not None:
            _chk(t, torch.int32, "traj_metrics.len")
    out = torch.empty(B, 4, device=gen.device, dtype=torch.float64)
    _l.call("esb_traj_metrics", gen.data_ptr(), gt.data_ptr(), _p(gen_len), _p(gt_len), B, Tg, Tt, D, int(cols[0]), int(cols[1]),
            int(rot_col), out.data_ptr(), _stream())
    return out


def traj_poses_to_ids(poses: torch.Tensor, lens: Optional[torch.Tensor] = None, num_bins: int = 256, base_id: int = 32006,
                      ts_id: int = 32003, tsep_id: int = 32004, te_id: int = 32005, eos_id: int = 2, pad_id: int = 0,
                      ranges: Optional[torch.Tensor] = None, out_len_total: Optional[int] = None):
    """Normalised (or, with ``ranges`` fp64 [6] = min/max of x,y,z, physical) poses [B,T,6] -> (ids [B,Lout] int64, len [B])."""
    _chk(poses, torch.float32, "traj_poses_to_ids.poses")
    poses = poses.contiguous()
    B, T, six = poses.shape
    assert six == 6
    Lout = out_len_total if out_len_total is not None else 3 + 7 * T
    out = torch.empty(B, Lout, device=poses.device, dtype=torch.int64)
    olen = torch.empty(B, device=poses.device, dtype=torch.int32)
    if lens is not None:
        _chk(lens, torch.int32, "traj_poses_to_ids.lens")
    if ranges is not None:
        _chk(ranges, torch.float64, "traj_poses_to_ids.ranges")
    _l.call("esb_traj_poses_to_ids", poses.data_ptr(), _p(lens), B, T, num_bins, base_id, ts_id, tsep_id, te_id, eos_id, pad_id,
            0 if ranges is None else 1, _p(ranges), Lout, out.data_ptr(), olen.data_ptr(), _stream())
    return out, olen


def traj_preprocess(traj: torch.Tensor, lens: Optional[torch.Tensor], num_steps: int, smooth: bool = False):
    """traj_utils.py smoothing_traj (optional) then preprocess_traj, batched: traj [B,T,D] fp32, lens [B] int32 or None
    -> (out [B,num_steps,D] fp32, padding mask [B,num_steps] int32)."""
    _chk(traj, torch.float32, "traj_preprocess.traj")
    traj = traj.contiguous()
    B, T, D = traj.shape
    if lens is not None:
        _chk(lens, torch.int32, "traj_preprocess.lens")
    out = torch.empty(B, num_steps, D, device=traj.device, dtype=torch.float32)
    mask = torch.empty(B, num_steps, device=traj.device, dtype=torch.int32)
    _l.call("esb_traj_preprocess", traj.data_ptr(), _p(lens), B, T, D, int(num_steps), 1 if smooth else 0, out.data_ptr(),
            mask.data_ptr(), _stream())
    return out, mask


def rgbd_to_points(rgbd: torch.Tensor, boxes: Optional[torch.Tensor], principal_point: float, focal_x: float, focal_y: float,
                   d_thres: Optional[float] = None, cap: Optional[int] = None):
    """pcm_tools.py:68-96 get_points_colors on device: rgbd [H,W,4] fp32, boxes [nb,4] int32 {xmin,ymin,xmax,ymax} or None
    -> (pts [cap,6] fp32 = xyz ++ rgb/255 in row-major pixel order, count int64 [1] device)."""
    _chk(rgbd, torch.float32, "rgbd_to_points.rgbd")
    rgbd = rgbd.contiguous()
    H, W, four = rgbd.shape
    assert four == 4
    nb = 0
    if boxes is not None and boxes.numel():
        _chk(boxes, torch.int32, "rgbd_to_points.boxes")
        boxes = boxes.contiguous()
        assert boxes.shape[1] == 4
        nb = boxes.shape[0]
    cap = H * W if cap is None else int(cap)
    out = torch.empty(cap, 6, device=rgbd.device, dtype=torch.float32)
    count = torch.empty(1, device=rgbd.device, dtype=torch.int64)
    wsb = int(_l.lib().raw("esb_rgbd_workspace_bytes")(H, W))
    ws = torch.empty(wsb, device=rgbd.device, dtype=torch.uint8)
    _l.call("esb_rgbd_to_points", rgbd.data_ptr(), H, W, _p(boxes) if nb else None, nb, float(principal_point), float(focal_x),
            float(focal_y), 0 if d_thres is None else 1, 0.0 if d_thres is None else float(d_thres), out.data_ptr(), cap,
            count.data_ptr(), ws.data_ptr(), wsb, _stream())
    return out, count


def pc_sample_norm(points: torch.Tensor, idx: Optional[torch.Tensor] = None, n: Optional[int] = None,
                   num_samples: Optional[int] = None) -> torch.Tensor:
    """pointllm/data/utils.py:146-157 pc_norm of the rows ``idx`` [B,S] (None: the first ``num_samples`` rows) of points
    [B,N,6] fp32 (only the first ``n`` rows of each cloud are valid) -> [B,S,6] fp32, ready for the point encoder."""
    _chk(points, torch.float32, "pc_sample_norm.points")
    if points.dim() == 2:
        points = points.unsqueeze(0)
    points = points.contiguous()
    B, N, six = points.shape
    assert six == 6
    n = N if n is None else int(n)
    if idx is not None:
        _chk(idx, torch.int64, "pc_sample_norm.idx")
        idx = idx.contiguous().view(B, -1)
        S = idx.shape[1]
    else:
        S = n if num_samples is None else int(num_samples)
    out = torch.empty(B, S, 6, device=points.device, dtype=torch.float32)
    _l.call("esb_pc_sample_norm", points.data_ptr(), N * 6, n, _p(idx), B, S, out.data_ptr(), _stream())
    return out


class pdl_chain:
    """Context manager around the single-token decode step: launches inside it use programmatic stream serialization
    (esb_set_pdl_chain); a no-op when ``pdl_chain_enabled`` is False."""

    def __enter__(self):
        if pdl_chain_enabled:
            _l.lib().raw("esb_set_pdl_chain")(1)
        return self

    def __exit__(self, *exc):
        _l.lib().raw("esb_set_pdl_chain")(0)
        return False
