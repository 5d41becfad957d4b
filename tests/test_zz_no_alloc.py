"""GPU: esb_gemm_skinny_static_bf16 - the no-allocation variant of the decode-step GEMM (split-K workspace in a file-scope device
array of the library) - against esb_gemm_skinny_bf16 with a caller workspace, through the C ABI, on two calls in a row.
Collected last on purpose (file name)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import lib


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _free_bytes():
    free, _ = torch.cuda.mem_get_info()
    return free


def _case(M, N, K, seed, tiled=False, residual=True):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
    res = torch.randn(M, N, device="cuda", generator=g).bfloat16() if residual else None
    wt = None
    if tiled:
        n = int(lib.lib().raw("esb_retiled_weight_elems")(N, K))
        wt = torch.empty(n, device="cuda", dtype=torch.bfloat16)
        lib.call("esb_retile_weight_bf16", w.data_ptr(), w.stride(0), N, K, wt.data_ptr(), _stream())
    return x, w, wt, res


def _existing(x, w, wt, res, splits):
    M, K = x.shape
    N = w.shape[0]
    out = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    nbytes = int(lib.lib().raw("esb_gemm_skinny_workspace_bytes")(M, N, K))
    ws = torch.zeros(nbytes, device="cuda", dtype=torch.uint8)
    wp, ldw, flag = (wt.data_ptr(), 0, 1) if wt is not None else (w.data_ptr(), w.stride(0), 0)
    lib.call("esb_gemm_skinny_bf16", x.data_ptr(), x.stride(0), wp, ldw, out.data_ptr(), out.stride(0), M, N, K,
             None if res is None else res.data_ptr(), 0 if res is None else res.stride(0), ws.data_ptr(), ws.numel(), splits, flag,
             _stream())
    return out


def _static(x, w, wt, res, splits, out):
    M, K = x.shape
    N = w.shape[0]
    wp, ldw, flag = (wt.data_ptr(), 0, 1) if wt is not None else (w.data_ptr(), w.stride(0), 0)
    lib.call("esb_gemm_skinny_static_bf16", x.data_ptr(), x.stride(0), wp, ldw, out.data_ptr(), out.stride(0), M, N, K,
             None if res is None else res.data_ptr(), 0 if res is None else res.stride(0), splits, flag, _stream())


def test_static_scratch_gemm_matches_workspace_gemm_on_two_calls_in_a_row():
    assert int(lib.lib().raw("esb_gemm_skinny_static_scratch_bytes")()) == (8 << 20) + 16384
    # two different problems back to back: the second must not see anything of the first in the shared scratch
    cases = [(_case(32, 4096, 4096, 1, tiled=True), 0), (_case(7, 1000, 11008, 2, tiled=False, residual=False), 4),
             (_case(64, 12288, 4096, 3, tiled=True), 0)]
    want = [_existing(*c, splits) for c, splits in cases]
    outs = [torch.full_like(wv, float("nan")) for wv in want]
    torch.cuda.synchronize()
    _static(*cases[0][0], cases[0][1], outs[0])            # first call on this device resolves the scratch address
    torch.cuda.synchronize()
    outs[0].fill_(float("nan"))
    free0, alloc0 = _free_bytes(), torch.cuda.memory_allocated()
    for (c, splits), o in zip(cases, outs):                # back to back on one stream, no synchronisation in between
        _static(*c, splits, o)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == alloc0 and _free_bytes() == free0, "the entry point must not allocate device memory"
    for (c, splits), o, wv in zip(cases, outs, want):
        x, w, _, res = c
        ref = x.float() @ w.float().t() + (0 if res is None else res.float())
        torch.testing.assert_close(o.float(), ref, atol=2e-2, rtol=1.6e-2)      # tolerance of tests/test_gemm_gpu.py
        torch.testing.assert_close(o.float(), wv.float(), atol=2e-2, rtol=1.6e-2)
    # same split factor -> same summation order -> same bits as the existing entry point
    assert torch.equal(outs[1], want[1])


def test_static_scratch_gemm_rejects_sizes_beyond_its_limits():
    x, w, _, _ = _case(4, 128, 64, 5)
    out = torch.empty(4, 128, device="cuda", dtype=torch.bfloat16)
    for M, N, K in ((65, 128, 64), (4, 16385, 64), (4, 128, 16392)):
        with pytest.raises(lib.EsbError, match="static scratch limits"):
            lib.call("esb_gemm_skinny_static_bf16", x.data_ptr(), x.stride(0), w.data_ptr(), w.stride(0), out.data_ptr(), out.stride(0),
                     M, N, K, None, 0, 0, 0, _stream())
