"""GPU: the persistent chain of decode-step Linears (esb_gemm_chain_bf16, csrc/gemm_chain.cu) against the one-launch-per-Linear
kernels it replaces (esb_gemm_skinny*_bf16 + silu_mul) and a plain fp32 torch reference of the same contractions."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops


def _rand(shape, g, scale=1.0):
    return (torch.randn(*shape, device="cuda", generator=g) * scale).bfloat16()


def _rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-20))


def test_grid_is_sized_from_resident_ctas():
    from egoscaler_b200 import lib
    n = [int(lib.lib().raw("esb_gemm_chain_ctas_per_sm")(m)) for m in (32, 64)]
    print("resident CTAs per SM (M<=32, M<=64):", n, lib.lib().last_error())
    assert min(n) >= 1


@pytest.mark.parametrize("M,N,K", [(5, 200, 72), (32, 4096, 4096), (32, 12288, 4096), (40, 1000, 512), (64, 4096, 11008),
                                   (1, 128, 64), (32, 32262, 4096), (17, 256, 688)])
def test_single_phase_matches_skinny_and_fp32(M, N, K):
    g = torch.Generator(device="cuda").manual_seed(M * 131 + N * 7 + K)
    x, w, res = _rand((M, K), g), _rand((N, K), g, K ** -0.5), _rand((M, N), g)
    tw = ops.retile_weight(w)
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm_chain([ops.ChainPhase(x, tw, out, residual=res)])
    ref = x.float() @ w.float().t() + res.float()
    torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=1.6e-2)
    assert _rel(out, ops.gemm_skinny(x, tw, residual=res)) < 2e-3
    out2 = torch.empty_like(out)
    ops.gemm_chain([ops.ChainPhase(x, tw, out2, residual=res)])
    assert torch.equal(out, out2)                                   # fixed summation order
    assert int(ops._chain_ws[("cuda", torch.cuda.current_device())][:16384].max()) == 0     # counters left at zero


@pytest.mark.parametrize("M,f,K", [(32, 11008, 4096), (7, 104, 72), (48, 688, 256), (32, 64, 64)])
def test_swiglu_phase_matches_gemm_plus_silu_mul(M, f, K):
    g = torch.Generator(device="cuda").manual_seed(M + f + K)
    x, w = _rand((M, K), g), _rand((2 * f, K), g, K ** -0.5)
    ln = (1.0 + 0.1 * torch.randn(K, device="cuda", generator=g)).bfloat16()
    _, rstd = ops.rmsnorm(x, ln, 1e-6, want_rstd=True)
    a = torch.full((M, f), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm_chain([ops.ChainPhase(x, ops.retile_weight_gateup(w, ln), a, row_scale=(rstd, 0), eps=1e-6, swiglu=True)])
    gu = ops.gemm_skinny(x, ops.retile_weight(w, ln), row_scale=(rstd, 0), eps=1e-6)
    want = ops.silu_mul(gu)
    assert _rel(a, want) < 3e-3
    hn = ops.rmsnorm(x, ln, 1e-6).float()
    y = hn @ w.float().t()
    ref = torch.nn.functional.silu(y[:, :f]) * y[:, f:]
    assert _rel(a, ref) < 1.5e-2


@pytest.mark.parametrize("M,d,f,n_next", [(32, 4096, 11008, 12288), (9, 256, 688, 768), (33, 384, 1000, 1152)])
def test_four_phase_chain_matches_separate_launches(M, d, f, n_next):
    """o-proj + residual -> gate|up + SwiGLU -> down-proj + residual -> next q|k|v, norms folded: against the same four Linears
    as separate esb_gemm_skinny_normed_bf16 launches + silu_mul (llm._decode_step_folded)."""
    g = torch.Generator(device="cuda").manual_seed(d + f + M)
    eps = 1e-6
    o, x = _rand((M, d), g), _rand((M, d), g)
    wo, wgu = _rand((d, d), g, d ** -0.5), _rand((2 * f, d), g, d ** -0.5)
    wd, wq = _rand((d, f), g, f ** -0.5), _rand((n_next, d), g, d ** -0.5)
    ln2 = (1.0 + 0.1 * torch.randn(d, device="cuda", generator=g)).bfloat16()
    ln1 = (1.0 + 0.1 * torch.randn(d, device="cuda", generator=g)).bfloat16()
    nt = (d + 127) // 128
    to, td = ops.retile_weight(wo), ops.retile_weight(wd)
    tq = ops.retile_weight(wq, ln1)
    # ---- separate launches
    ssq = torch.zeros(2, nt, 64, device="cuda")
    h0 = ops.gemm_skinny(o, to, residual=x, ssq_out=ssq[0])
    gu0 = ops.gemm_skinny(h0, ops.retile_weight(wgu, ln2), row_scale=(ssq[0], nt), eps=eps)
    a0 = ops.silu_mul(gu0)
    x0 = ops.gemm_skinny(a0, td, residual=h0, ssq_out=ssq[1])
    q0 = ops.gemm_skinny(x0, tq, row_scale=(ssq[1], nt), eps=eps)
    # ---- one launch
    h = torch.full((M, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    a = torch.full((M, f), float("nan"), device="cuda", dtype=torch.bfloat16)
    xn = torch.full((M, d), float("nan"), device="cuda", dtype=torch.bfloat16)
    q = torch.full((M, n_next), float("nan"), device="cuda", dtype=torch.bfloat16)
    ssq1 = torch.full((2, nt, 64), float("nan"), device="cuda")
    phases = [ops.ChainPhase(o, to, h, residual=x, ssq_out=ssq1[0]),
              ops.ChainPhase(h, ops.retile_weight_gateup(wgu, ln2), a, row_scale=(ssq1[0], nt), eps=eps, swiglu=True),
              ops.ChainPhase(a, td, xn, residual=h, ssq_out=ssq1[1]),
              ops.ChainPhase(xn, tq, q, row_scale=(ssq1[1], nt), eps=eps)]
    ops.gemm_chain(phases)
    for name, got, want in (("h", h, h0), ("a", a, a0), ("x", xn, x0), ("qkv", q, q0)):
        assert torch.isfinite(got.float()).all(), name
        assert _rel(got, want) < 4e-3, (name, _rel(got, want))
    torch.testing.assert_close(ssq1[0, :, :M].sum(0), h.float().pow(2).sum(1), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(ssq1[1, :, :M].sum(0), xn.float().pow(2).sum(1), rtol=1e-4, atol=1e-4)
    # replay: same bits, hand-off counter and tile counters back at zero
    keep = [t.clone() for t in (h, a, xn, q)]
    for _ in range(3):
        ops.gemm_chain(phases)
    for t, k in zip((h, a, xn, q), keep):
        assert torch.equal(t, k)
    assert int(ops._chain_ws[("cuda", torch.cuda.current_device())][:16384].max()) == 0
