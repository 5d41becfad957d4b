"""GPU: tcgen05 GEMM vs a plain fp32 torch reference of the same contraction (floating-point kernel)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops


def _ref(a, w, bias=None, res=None, act=0, alpha=1.0):
    y = alpha * (a.float() @ w.float().t())
    if bias is not None:
        y = y + bias
    if act == 1:
        y = torch.nn.functional.gelu(y)
    elif act == 2:
        y = torch.relu(y)
    if res is not None:
        y = y + res.float()
    return y


SHAPES = [(128, 256, 64), (128, 256, 256), (256, 512, 4096), (1392, 768, 256), (1026, 1152, 384), (300, 688, 256),
          (77, 40, 688), (5568, 4096, 4096), (1216, 32262, 256), (32, 12288, 4096), (513, 384, 1536)]


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("bn", [0, 64, 128, 256])
def test_gemm_matches_fp32_reference(M, N, K, bn):
    if bn != 0 and M * N > 4e7:
        pytest.skip("large shape only with auto tile")
    g = torch.Generator(device="cuda").manual_seed(M * 31 + N * 7 + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    out = ops.gemm(a, w, force_bn=bn)
    ref = _ref(a, w)
    # bf16 output rounding: 2^-8 relative, fp32 accumulation error is far below that
    torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=1.6e-2)
    assert (out.float() - ref).abs().max() / ref.abs().max() < 8e-3


def test_gemm_epilogues_and_fp32_out():
    g = torch.Generator(device="cuda").manual_seed(5)
    M, N, K = 513, 1000, 384
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(N, device="cuda", generator=g)
    res = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    for act in (0, 1, 2):
        out = ops.gemm(a, w, bias=bias, residual=res, act=act, out_dtype=torch.float32, alpha=0.5)
        torch.testing.assert_close(out, _ref(a, w, bias, res, act, 0.5), atol=2e-3, rtol=1e-3)
    out2 = ops.gemm(a, w, out=out.clone(), accumulate=True)
    torch.testing.assert_close(out2, out + _ref(a, w), atol=2e-3, rtol=1e-3)
    # unaligned ldc (vocab-like N) and strided A
    a2 = torch.randn(M, 2 * K, device="cuda", generator=g).bfloat16()[:, :K]
    w2 = w[:999]
    out3 = ops.gemm(a2, w2, bias=bias[:999].contiguous())
    torch.testing.assert_close(out3.float(), _ref(a2, w2, bias[:999]), atol=3e-2, rtol=1.6e-2)


SKINNY = [(32, 4096, 4096), (1, 12288, 4096), (32, 32262, 4096), (64, 22016, 4096), (17, 4096, 11008), (33, 1000, 320),
          (8, 256, 688), (32, 688, 256)]


@pytest.mark.parametrize("M,N,K", SKINNY)
@pytest.mark.parametrize("splits", [0, 1, 3])
def test_gemm_skinny_matches_fp32_reference(M, N, K, splits):
    """Decode-step weight-streaming GEMM (split-K, in-kernel deterministic reduction) vs fp32 torch, with residual."""
    g = torch.Generator(device="cuda").manual_seed(M * 31 + N * 7 + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    res = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    out = ops.gemm_skinny(a, w, residual=res, force_splits=splits)
    ref = _ref(a, w, res=res)
    torch.testing.assert_close(out.float(), ref, atol=3e-2, rtol=1.6e-2)
    out_nores = ops.gemm_skinny(a, w, force_splits=splits)
    torch.testing.assert_close(out_nores.float(), _ref(a, w), atol=2e-2, rtol=1.6e-2)
    # pre-tiled weight layout (the decode path's resident copy): identical arithmetic, identical bits
    wt = ops.retile_weight(w)
    assert torch.equal(ops.gemm_skinny(a, wt, residual=res, force_splits=splits), out)
    # deterministic: the reduction order over K slices is fixed, whichever CTA arrives last
    for _ in range(3):
        assert torch.equal(ops.gemm_skinny(a, w, residual=res, force_splits=splits), out)
    # counters are left at zero (graph-replay safe)
    assert int(ops._skinny_workspace(a.device)[:16384].max()) == 0


def test_gemm_dispatches_skinny_for_decode_rows():
    g = torch.Generator(device="cuda").manual_seed(3)
    a = torch.randn(32, 4096, device="cuda", generator=g).bfloat16()
    w = (torch.randn(4096, 4096, device="cuda", generator=g) / 64).bfloat16()
    ops.skinny_enabled = True
    y1 = ops.gemm(a, w)
    ops.skinny_enabled = False
    try:
        y0 = ops.gemm(a, w)
    finally:
        ops.skinny_enabled = True
    torch.testing.assert_close(y1.float(), y0.float(), atol=2e-2, rtol=1.6e-2)


@pytest.mark.parametrize("M,N,K,splits", [(32, 4096, 4096, 0), (32, 1000, 512, 1), (50, 640, 1024, 4), (7, 4096, 11008, 0),
                                          (32, 12288, 4096, 0)])
def test_skinny_folded_rmsnorm(M, N, K, splits):
    """RMSNorm folded around the decode-step GEMM (esb_gemm_skinny_normed_bf16 + esb_retile_weight_scaled_bf16):
    (a) rstd * (x . (W*g)^T) == Linear(LlamaRMSNorm(x)) within bf16 rounding; (b) ssq_out of a producing GEMM (with residual) is
    the exact per-row sum of squares of its bf16 output, deterministic; (c) a consumer fed with those partials == the
    consumer fed with rstd from the stand-alone rmsnorm kernel."""
    g = torch.Generator().manual_seed(M * 7 + N + K)
    eps = 1e-6
    x = (torch.randn(M, K, generator=g) * 3.0).bfloat16().cuda()
    gam = (1.0 + 0.2 * torch.randn(K, generator=g)).bfloat16().cuda()
    w = (torch.randn(N, K, generator=g) * 0.03).bfloat16().cuda()
    xf = x.float()
    rstd_ref = torch.rsqrt((xf * xf).mean(-1, keepdim=True) + eps)
    ref = ((xf * rstd_ref).bfloat16().float() * gam.float()).bfloat16().float() @ w.float().T      # HF order of roundings
    tw = ops.retile_weight(w, gam)
    _, rstd = ops.rmsnorm(x, gam, eps, want_rstd=True)
    got = ops.gemm_skinny(x, tw, row_scale=(rstd, 0), eps=eps, force_splits=splits)
    err = float((got.float() - ref).norm() / ref.norm())
    assert err < 8e-3, err
    # (b) producer: out = a @ w2^T + res, statistics of out
    a = (torch.randn(M, N, generator=g)).bfloat16().cuda()
    w2 = (torch.randn(K, N, generator=g) * 0.05).bfloat16().cuda()
    res = torch.randn(M, K, generator=g).bfloat16().cuda()
    nt = (K + 127) // 128
    ssq = torch.full((nt, 64), -1.0, device="cuda")
    out = ops.gemm_skinny(a, w2, residual=res, ssq_out=ssq, force_splits=splits)
    plain = ops.gemm_skinny(a, w2, residual=res, force_splits=splits)
    assert torch.equal(out, plain)
    of = out.float()
    pad = torch.zeros(M, nt * 128, device="cuda")
    pad[:, :K] = of * of
    tile_ref = pad.view(M, nt, 128).sum(-1).T                                  # [nt, M]
    torch.testing.assert_close(ssq[:, :M], tile_ref, rtol=2e-6, atol=1e-6)
    ssq2 = torch.empty_like(ssq)
    ops.gemm_skinny(a, w2, residual=res, ssq_out=ssq2, force_splits=splits)
    assert torch.equal(ssq[:, :M], ssq2[:, :M])
    # (c) consumer from the partial sums == consumer from the rmsnorm kernel's rstd (same rsqrtf, sums differ by an ulp)
    w3 = (torch.randn(256, K, generator=g) * 0.03).bfloat16().cuda()
    tw3 = ops.retile_weight(w3, gam)
    c1 = ops.gemm_skinny(out, tw3, row_scale=(ssq, nt), eps=eps)
    _, r2 = ops.rmsnorm(out, gam, eps, want_rstd=True)
    c2 = ops.gemm_skinny(out, tw3, row_scale=(r2, 0), eps=eps)
    assert float((c1.float() - c2.float()).abs().max()) <= 2e-2 * float(c2.float().abs().max())
    assert float((c1.float() - c2.float()).norm() / c2.float().norm()) < 2e-3


@pytest.mark.parametrize("M,N,K", [(1000, 1024, 32768), (4000, 1000, 32768), (5568, 4096, 11008), (5568, 4096, 22016),
                                   (5568, 4096, 12288)])
def test_gemm_grouped_tile_walk(M, N, K):
    """Shapes whose operand panels exceed the L2 budget, so launch_gemm walks the tiles in several M-groups (first shape) or
    N-groups (the others - the down-proj forward and the two large data-gradient GEMMs of the 7B trunk among them): every
    output tile must still be produced exactly once."""
    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) / K ** 0.5).bfloat16()
    out = torch.full((M, N), float("nan"), device="cuda", dtype=torch.bfloat16)
    ops.gemm(a, w, out=out)
    ref = _ref(a, w)
    assert bool(torch.isfinite(out).all())
    torch.testing.assert_close(out.float(), ref, atol=2e-2, rtol=1.6e-2)


# ------------------------------------------------------------------ fused epilogues == the two launches they replace, bit for bit
def _rnd(*shape, seed, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).bfloat16()


@pytest.mark.parametrize("M,f,K", [(1392, 688, 256), (300, 128, 256), (77, 40, 64), (5568, 11008, 4096)])
def test_gemm_swiglu_fused_equals_gemm_then_silu_mul(M, f, K):
    """HF LlamaMLP gate|up projection + act_fn(gate) * up (modeling_llama.py:171-184); ragged f (688 = tiny config) and the
    full-size launch of the benchmark."""
    a = _rnd(M, K, seed=1)
    wgu = _rnd(2 * f, K, seed=2, scale=K ** -0.5)
    gu_ref = ops.gemm(a, wgu)
    act_ref = ops.silu_mul(gu_ref)
    act, gu = ops.gemm_swiglu(a, wgu, want_gu=True)
    assert torch.equal(gu, gu_ref)
    assert torch.equal(act, act_ref)
    act2, none = ops.gemm_swiglu(a, wgu, want_gu=False)
    assert none is None and torch.equal(act2, act_ref)
    ref = torch.nn.functional.silu(a.float() @ wgu[:f].float().t()) * (a.float() @ wgu[f:].float().t())
    assert (act.float() - ref).abs().max() <= 2e-2 * ref.abs().max()


@pytest.mark.parametrize("M,f,K", [(1392, 688, 256), (77, 40, 64), (5568, 11008, 4096)])
def test_gemm_dswiglu_fused_equals_gemm_then_silu_mul_bwd(M, f, K):
    dout = _rnd(M, K, seed=3)
    wdT = _rnd(f, K, seed=4, scale=K ** -0.5)
    gu = _rnd(M, 2 * f, seed=5)
    ref = ops.silu_mul_bwd(gu, ops.gemm(dout, wdT))
    got = ops.gemm_dswiglu(dout, wdT, gu)
    assert torch.equal(got, ref)


@pytest.mark.parametrize("B,L,H,K,pos0", [(2, 696, 2, 256, 0), (3, 77, 2, 256, 5), (8, 696, 32, 4096, 0)])
def test_gemm_rope_fused_equals_gemm_then_rope(B, L, H, K, pos0):
    """q|k|v projection + apply_rotary_pos_emb (modeling_llama.py:138-168) at head_dim 128: tiny-hd128 config and the
    full-size launch (N = 12288)."""
    from egoscaler_b200.llm import rope_tables
    hd, d = 128, H * 128
    cos, sin = rope_tables(hd, 10000.0, 1024, "cuda")
    a = _rnd(B * L, K, seed=6)
    w = _rnd(3 * d, K, seed=7, scale=K ** -0.5)
    ref = ops.gemm(a, w)
    ops.rope_(ref, L, 2 * H, hd, cos, sin, pos0)
    got = ops.gemm_rope(a, w, cos, sin, L, pos0, 2 * d, hd)
    assert torch.equal(got, ref)
    assert torch.equal(got[:, 2 * d:], ops.gemm(a, w)[:, 2 * d:])             # v is not rotated


def test_trunk_fused_and_unfused_paths_agree_bitwise():
    """llm.LlamaTrunk forward + backward with every fused epilogue on vs off (ops.fuse_* switches): identical hidden states
    and identical input gradients - the switches only change how many launches a layer takes."""
    from egoscaler_b200 import synth
    from egoscaler_b200.config import TrajVLMConfig
    from egoscaler_b200.model import TrajPointLLMForCausalLM
    cfg = TrajVLMConfig.tiny_hd128()
    m = TrajPointLLMForCausalLM(cfg, "cuda")
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    B, L, d = 2, 200, cfg.hidden_size
    x = _rnd(B * L, d, seed=8)
    dout = _rnd(B * L, d, seed=9)
    kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    kmask[0, 90:97] = 0
    res = {}
    old = (ops.fuse_swiglu, ops.fuse_dswiglu, ops.fuse_rope, ops.fuse_rope_bwd)
    try:
        for on in (False, True):
            ops.fuse_swiglu = ops.fuse_dswiglu = ops.fuse_rope = ops.fuse_rope_bwd = on
            saved = []
            h = m.trunk.forward(x.clone(), B, L, kmask, saved=saved)
            dx = m.trunk.backward(dout.clone(), B, L, kmask, saved)
            h_inf = m.trunk.forward(x.clone(), B, L, kmask)                     # inference form (gu not kept)
            res[on] = (h, dx, h_inf)
    finally:
        ops.fuse_swiglu, ops.fuse_dswiglu, ops.fuse_rope, ops.fuse_rope_bwd = old
    for a_, b_ in zip(res[False], res[True]):
        assert torch.equal(a_, b_)


@pytest.mark.parametrize("B,L,H", [(2, 200, 2), (1, 77, 3), (8, 696, 32)])
def test_attention_bwd_fused_inverse_rope_equals_two_launches(B, L, H):
    """Attention backward with the transposed rotary embedding fused into the dq / dk epilogues (esb_attn_bwd_rope) ==
    esb_attn_bwd followed by esb_rope(inverse), bit for bit, incl. the benchmark's launch shape."""
    from egoscaler_b200.llm import rope_tables
    hd, d = 128, H * 128
    cos, sin = rope_tables(hd, 10000.0, 1024, "cuda")
    qkv = _rnd(B, L, 3, H, hd, seed=11)
    do = _rnd(B, L, d, seed=12)
    kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    kmask[0, L // 2:L // 2 + 6] = 0
    o, lse = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True, kmask=kmask, want_lse=True)
    ref = ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask)
    ops.rope_(ref.view(B * L, 3 * d), L, 2 * H, hd, cos, sin, 0, inverse=True)
    got = ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask, rope=(cos, sin))
    assert torch.equal(got, ref)


@pytest.mark.parametrize("T,M,N", [(1100, 1100, 640), (1392, 256, 768), (1216, 32262, 256), (300, 688, 256), (77, 40, 24), (5568, 4096, 11008), (1026, 384, 1152)])
def test_gemm_tn_weight_gradient_form(T, M, N):
    """dW[M,N] = A^T . B with A [T,M], B [T,N] row-major (MN-major tcgen05 operands) vs fp32 torch and vs the transpose + NT path;
    fp32 accumulate-into-C, strided A (the logits buffer has a padded leading dimension)."""
    a_buf = _rnd(T, (M + 7) // 8 * 8 + 8, seed=21)                           # padded leading dimension (multiple of 8)
    a = a_buf[:, :M]
    b = _rnd(T, N, seed=22)
    ref = a.float().t() @ b.float()
    got = ops.gemm_tn(a, b)
    assert got.dtype == torch.float32
    assert (got - ref).abs().max() <= 2e-3 * ref.abs().max() + 1e-3
    old = ops.gemm_tn_enabled
    ops.gemm_tn_enabled = False
    try:
        via_nt = ops.gemm_tn(a, b)
    finally:
        ops.gemm_tn_enabled = old
    assert (got - via_nt).abs().max() <= 1e-4 * ref.abs().max() + 1e-4         # same products, same fp32 accumulation order per k-block
    acc = torch.ones(M, N, device="cuda")
    ops.gemm_tn(a, b, out=acc, accumulate=True)
    assert (acc - 1.0 - ref).abs().max() <= 2e-3 * ref.abs().max() + 1e-3


@pytest.mark.parametrize("M,Kc,N", [(1100, 320, 640), (1392, 768, 256), (1216, 32262, 256), (300, 256, 688), (77, 24, 40), (5568, 11008, 4096), (5568, 4096, 4096)])
def test_gemm_nn_data_gradient_form(M, Kc, N):
    """dX[M,N] = A . W with A [M,Kc] and W [Kc,N] = a weight in its [out, in] state_dict layout (MN-major B operand) vs fp32 torch
    and vs the NT GEMM on a transposed copy (same products and k-block order: equal up to rare rounding flips)."""
    a_buf = _rnd(M, (Kc + 7) // 8 * 8 + 8, seed=23)
    a = a_buf[:, :Kc]
    w = _rnd(Kc, N, seed=24, scale=Kc ** -0.5)
    got = ops.gemm_nn(a, w)
    ref = a.float() @ w.float()
    assert (got.float() - ref).abs().max() <= 1e-2 * ref.abs().max()
    if Kc % 8 == 0:                                                          # the NT kernel needs a 16-byte-aligned row pitch
        nt = ops.gemm(a, w.t().contiguous())
        assert (got.float() - nt.float()).abs().max() <= 8e-3 * ref.abs().max()    # at most a bf16 rounding flip apart
        assert (got != nt).float().mean() < 1e-2


def test_gemm_dswiglu_on_the_down_proj_parameter_itself():
    M, f, K = 1392, 688, 256
    dout = _rnd(M, K, seed=3)
    wd = _rnd(K, f, seed=4, scale=K ** -0.5)                              # down_proj.weight [d, f]
    gu = _rnd(M, 2 * f, seed=5)
    ref = ops.silu_mul_bwd(gu, ops.gemm(dout, wd.t().contiguous()))
    got = ops.gemm_dswiglu(dout, wd, gu, w_is_down_proj=True)
    assert (got.float() - ref.float()).abs().max() <= 8e-3 * ref.float().abs().max() and (got != ref).float().mean() < 1e-2


def test_cta_pair_gemm_matches_single_cta_kernel():
    """tcgen05.mma.cta_group::2 kernel (clusters of two CTAs, 256 x 256 tile, esb_set_gemm_2cta) == the single-CTA kernel bit for
    bit on plain / residual / fp32 / SwiGLU / dSwiGLU / RoPE epilogues.  Runs in a subprocess with a timeout: a mistake in the
    cross-CTA mbarrier protocol would hang, and a hang must fail this test, not the box."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    try:
        r = subprocess.run([sys.executable, os.path.join(root, "tools", "gemm2_check.py")], capture_output=True, text=True, timeout=240)
    except subprocess.TimeoutExpired:
        pytest.fail("CTA-pair GEMM check timed out (hang in the two-CTA pipeline)")
    assert r.returncode == 0, r.stderr[-2000:]
    res = json.loads(r.stdout.strip().splitlines()[-1])
    assert res["ok"], {k: v for k, v in res["checks"].items() if not v}
    assert len(res["checks"]) >= 12
