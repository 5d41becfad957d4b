"""GPU: op-level numerics of the floating-point kernels against plain fp32 torch references of the same op
(RMSNorm / LayerNorm / RoPE / SwiGLU / GELU / attention fwd+bwd+decode / CE / AdamW / splice / pooling)."""
import math

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops
from egoscaler_b200.llm import rope_tables


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    return (torch.randn(*shape, device="cuda", generator=g) * scale).bfloat16()


def close(a, b, tol):
    a, b = a.float(), b.float()
    err = (a - b).abs().max().item()
    ref = b.abs().max().item()
    assert err <= tol * max(ref, 1e-6), (err, ref)


@pytest.mark.parametrize("rows,d", [(7, 256), (1392, 4096), (33, 688)])
def test_rmsnorm_fwd_bwd(rows, d):
    x, w, dy, dres = rnd(rows, d, seed=1), (1 + 0.1 * rnd(d, seed=2).float()).bfloat16(), rnd(rows, d, seed=3), rnd(rows, d, seed=4)
    y, rstd = ops.rmsnorm(x, w, 1e-6, want_rstd=True)
    xf = x.float().requires_grad_(True)
    wf = w.float().requires_grad_(True)
    ref = wf * (xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6))
    close(y, ref, 1e-2)
    ref.backward(dy.float())
    dw = torch.zeros(d, device="cuda")
    dx = ops.rmsnorm_bwd(dy, x, w, rstd, dres=dres, dw=dw)
    close(dx, xf.grad + dres.float(), 1.5e-2)
    close(dw, wf.grad, 1e-2)
    # strided input rows
    big = rnd(rows, 2 * d, seed=5)
    close(ops.rmsnorm(big[:, :d], w, 1e-6), ops.rmsnorm(big[:, :d].contiguous(), w, 1e-6), 0.0)


def test_rmsnorm_register_path_equals_general_path():
    """d = 16 x threads (the trunk's 4096): the register-resident kernels against the general two-pass kernels
    (ESB_RMSNORM_GENERIC=1 selects those), with and without residual / weight gradient, over several rows per CTA."""
    import os
    rows, d = 5568, 4096
    x, w, dy, dres = rnd(rows, d, seed=11), (1 + 0.1 * rnd(d, seed=12).float()).bfloat16(), rnd(rows, d, seed=13), rnd(rows, d, seed=14)
    res = {}
    for mode in ("0", "1"):
        os.environ["ESB_RMSNORM_GENERIC"] = mode
        try:
            y, rstd = ops.rmsnorm(x, w, 1e-6, want_rstd=True)
            dx_plain = ops.rmsnorm_bwd(dy, x, w, rstd)
            dw = torch.zeros(d, device="cuda")
            dx_full = ops.rmsnorm_bwd(dy, x, w, rstd, dres=dres, dw=dw)
        finally:
            os.environ.pop("ESB_RMSNORM_GENERIC", None)
        res[mode] = (y, rstd, dx_plain, dx_full, dw)
    # same formulas; the compiler contracts the sums of products into FMAs differently in the two code shapes, so fp32
    # statistics agree to round-off and a few bf16 outputs land on the neighbouring value
    torch.testing.assert_close(res["0"][1], res["1"][1], rtol=2e-6, atol=0)
    for a, b in zip((res["0"][0], res["0"][2], res["0"][3]), (res["1"][0], res["1"][2], res["1"][3])):
        diff = (a.float() - b.float()).abs()
        assert float((diff > 0).float().mean()) < 2e-2
        assert bool((diff <= 2.0 ** -7 * b.float().abs() + 1e-6).all())              # at most one bf16 ulp
    close(res["0"][4], res["1"][4], 1e-4)          # fp32 atomics over CTAs: order differs with the grid size


@pytest.mark.parametrize("rows,d", [(1026, 384), (50, 1152)])
def test_layernorm_fused_add(rows, d):
    x, add = rnd(rows, d, seed=1), rnd(rows, d, seed=2)
    g, b = 1 + 0.1 * torch.randn(d, device="cuda"), 0.1 * torch.randn(d, device="cuda")
    s_ref = (x.float() + add.float()).bfloat16()
    xs = x.clone()
    y = ops.layernorm(xs, g, b, add=add, sum_out=xs)
    assert torch.equal(xs, s_ref)
    close(y, F.layer_norm(s_ref.float(), (d,), g, b, 1e-5), 1e-2)
    close(ops.layernorm(x, g, b), F.layer_norm(x.float(), (d,), g, b, 1e-5), 1e-2)


@pytest.mark.parametrize("hd,H", [(64, 4), (128, 2)])
def test_rope_forward_inverse_and_cache(hd, H):
    B, L, d = 2, 37, H * hd
    cos, sin = rope_tables(hd, 10000.0, 128, "cuda")
    qkv = rnd(B * L, 3 * d, seed=3)
    ref = qkv.clone().float().view(B, L, 3, H, hd)
    c = torch.cat((cos, cos), -1)[None, 5:5 + L, None, :].float()
    s = torch.cat((sin, sin), -1)[None, 5:5 + L, None, :].float()
    rot = lambda t: torch.cat((-t[..., hd // 2:], t[..., :hd // 2]), -1)
    for j in (0, 1):
        ref[:, :, j] = ref[:, :, j] * c + rot(ref[:, :, j]) * s
    kc = torch.zeros(B, 64, d, device="cuda", dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    out = ops.rope_(qkv.clone(), L, 2 * H, hd, cos, sin, pos0=5, kcache=kc, vcache=vc, kv_head0=H, kv_heads=H, v_col0=2 * d,
                    cache_len=64)
    close(out.view(B, L, 3, H, hd), ref, 1.2e-2)
    assert torch.equal(kc[:, 5:5 + L].reshape(B, L, d), out.view(B, L, 3, d)[:, :, 1])
    assert torch.equal(vc[:, 5:5 + L].reshape(B, L, d), qkv.view(B, L, 3, d)[:, :, 2])
    back = ops.rope_(out.clone(), L, 2 * H, hd, cos, sin, pos0=5, inverse=True)
    close(back.view(B, L, 3, d)[:, :, :2], qkv.view(B, L, 3, d)[:, :, :2], 2e-2)


def test_swiglu_and_gelu():
    rows, f = 300, 688
    gu, da = rnd(rows, 2 * f, seed=1), rnd(rows, f, seed=2)
    guf = gu.float().requires_grad_(True)
    ref = F.silu(guf[:, :f]) * guf[:, f:]
    close(ops.silu_mul(gu), ref, 1e-2)
    ref.backward(da.float())
    close(ops.silu_mul_bwd(gu, da), guf.grad, 1.5e-2)
    z, dy = rnd(64, 1024, seed=3), rnd(64, 1024, seed=4)
    zf = z.float().requires_grad_(True)
    r = F.gelu(zf)
    close(ops.gelu(z), r, 1e-2)
    r.backward(dy.float())
    close(ops.gelu_bwd(z, dy), zf.grad, 1.5e-2)


def _attn_ref(q, k, v, causal, kmask, q_pos0=0):
    B, Lq, H, hd = q.shape
    Lk = k.shape[1]
    s = torch.einsum("bihe,bjhe->bhij", q.float(), k.float()) * hd ** -0.5
    m = torch.ones(B, 1, Lq, Lk, dtype=torch.bool, device=q.device)
    if causal:
        m = m & (torch.arange(Lk, device=q.device)[None, :] <= (torch.arange(Lq, device=q.device)[:, None] + q_pos0))[None, None]
    if kmask is not None:
        m = m & kmask[:, None, None, :Lk].bool()
    s = s.masked_fill(~m, float("-inf"))
    p = torch.softmax(s, -1)
    return torch.einsum("bhij,bjhe->bihe", p, v.float()).reshape(B, Lq, H * hd), torch.logsumexp(s, -1)


@pytest.mark.parametrize("hd,H,L,causal", [(64, 4, 696, True), (128, 2, 200, True), (64, 6, 513, False), (128, 3, 64, True), (64, 2, 77, True),
                                              (96, 12, 513, False), (96, 2, 130, True)])
def test_attention_fwd_bwd(hd, H, L, causal):
    B, d = 2, H * hd
    qkv = rnd(B, L, 3, H, hd, seed=7)
    kmask = None
    if causal:
        kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
        kmask[0, L // 2:L // 2 + 5] = 0                           # mid-sequence padding like collate_fn
        kmask[1, L - 9:] = 0
    o, lse = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=causal, kmask=kmask, want_lse=True)
    qf = qkv.float().requires_grad_(True)
    ref_o, ref_lse = _attn_ref(qf[:, :, 0], qf[:, :, 1], qf[:, :, 2], causal, kmask)
    close(o, ref_o, 1.5e-2)
    assert (lse - ref_lse).abs().max() < 2e-2
    do = rnd(B, L, d, seed=8)
    ref_o.backward(do.float())
    dqkv = ops.attention_bwd(qkv, o, lse, do, causal=causal, kmask=kmask)
    for j, name in enumerate("qkv"):
        close(dqkv[:, :, j], qf.grad[:, :, j], 2.5e-2)


def _collate_like_mask(B, L):
    """Key-padding masks shaped like collate_fn's (dataset.py:150-194): 515 point-block tokens, a description padded to 20
    (pads sit MID-sequence, before the separator), trajectory tokens, tail padding."""
    kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    g = torch.Generator().manual_seed(5)
    for b in range(B):
        dl = int(torch.randint(8, 21, (1,), generator=g))
        kmask[b, 515 + dl:535] = 0
        kmask[b, L - int(torch.randint(0, 24, (1,), generator=g)):] = 0 if b % 2 else 1
    return kmask


def test_attention_fwd_bwd_at_bench_shape():
    """The exact launch shape bench.py times (BASELINE configs[1]): B=8, 32 heads x 128, L=696, causal + collate-like key
    padding, against fp32 torch: forward, lse and all three gradients of the tcgen05 kernels."""
    B, H, hd, L = 8, 32, 128, 696
    d = H * hd
    qkv = rnd(B, L, 3, H, hd, seed=17)
    kmask = _collate_like_mask(B, L)
    o, lse = ops.attention(qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], causal=True, kmask=kmask, want_lse=True)
    qf = qkv.float().requires_grad_(True)
    ref_o, ref_lse = _attn_ref(qf[:, :, 0], qf[:, :, 1], qf[:, :, 2], True, kmask)
    close(o, ref_o, 1.5e-2)
    assert (lse - ref_lse).abs().max() < 2e-2
    do = rnd(B, L, d, seed=18)
    ref_o.backward(do.float())
    dqkv = ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask)
    for j in range(3):
        close(dqkv[:, :, j], qf.grad[:, :, j], 2.5e-2)
    # per-row check on the rows just after the padded gap (where a mask bug would show first)
    for b in (0, 3):
        r = slice(535, 540)
        close(o[b, r], ref_o[b, r], 1.5e-2)


def test_attention_prefill_at_bench_shape():
    """Decode prefill of BASELINE configs[3]: P=544 prompt queries against the static KV cache (row stride = max_len),
    32 heads x 128, then one cached token on top of it."""
    B, H, hd, P, Lmax = 4, 32, 128, 544, 696
    d = H * hd
    kc, vc = rnd(B, Lmax, d, seed=21), rnd(B, Lmax, d, seed=22)
    kmask = _collate_like_mask(B, Lmax)
    kmask[:, P:] = 1
    q = rnd(B, P, H, hd, seed=23)
    o = ops.attention(q, kc[:, :P].view(B, P, H, hd), vc[:, :P].view(B, P, H, hd), causal=True, kmask=kmask, q_pos0=0)
    ref, _ = _attn_ref(q, kc[:, :P].view(B, P, H, hd), vc[:, :P].view(B, P, H, hd), True, kmask)
    close(o, ref, 1.5e-2)
    q1 = rnd(B, 3 * d, seed=24)
    od = ops.attention_decode(q1, kc, vc, H, hd, kmask=kmask, len_static=P + 1)
    ref1, _ = _attn_ref(q1[:, :d].reshape(B, 1, H, hd), kc[:, :P + 1].view(B, P + 1, H, hd), vc[:, :P + 1].view(B, P + 1, H, hd),
                        False, kmask)
    close(od, ref1.view(B, d), 1.5e-2)


@pytest.mark.parametrize("hd,H", [(64, 4), (128, 32)])
def test_attention_prefill_with_cache_and_decode(hd, H):
    B, d, Lmax, P = 3, H * hd, 96, 70
    kc, vc = rnd(B, Lmax, d, seed=1), rnd(B, Lmax, d, seed=2)
    kmask = torch.ones(B, Lmax, device="cuda", dtype=torch.uint8)
    kmask[1, 10:14] = 0
    # chunked prefill: 6 new queries at positions P-6..P-1 against the cache
    q = rnd(B, 6, H, hd, seed=3)
    o = ops.attention(q, kc[:, :P].view(B, P, H, hd), vc[:, :P].view(B, P, H, hd), causal=True, kmask=kmask, q_pos0=P - 6)
    ref, _ = _attn_ref(q, kc[:, :P].view(B, P, H, hd), vc[:, :P].view(B, P, H, hd), True, kmask, P - 6)
    close(o, ref, 1.5e-2)
    # single-token decode
    q1 = rnd(B, 3 * d, seed=4)
    od = ops.attention_decode(q1, kc, vc, H, hd, kmask=kmask, len_static=P)
    ref1, _ = _attn_ref(q1[:, :d].reshape(B, 1, H, hd), kc[:, :P].view(B, P, H, hd), vc[:, :P].view(B, P, H, hd), False, kmask)
    close(od, ref1.view(B, d), 1.5e-2)
    ln = torch.tensor([P], device="cuda", dtype=torch.int32)
    assert torch.equal(ops.attention_decode(q1, kc, vc, H, hd, kmask=kmask, len_ptr=ln), od)
    ln2 = torch.tensor([P - 5], device="cuda", dtype=torch.int32)        # context = len_static + *len_ptr
    assert torch.equal(ops.attention_decode(q1, kc, vc, H, hd, kmask=kmask, len_ptr=ln2, len_static=5), od)


def test_cross_entropy_and_argmax():
    rows, V, ld = 304, 32262, 32264
    buf = rnd(rows, ld, seed=5, scale=2.0)
    logits = buf[:, :V]
    tg = torch.randint(3, V, (rows,), device="cuda")
    tg[::7] = 0
    lf = logits.float().requires_grad_(True)
    ref = F.cross_entropy(lf, tg, ignore_index=0)
    ref.backward()
    am = ops.argmax_rows(logits, V)
    assert torch.equal(am, logits.float().argmax(-1))
    sc = torch.empty(rows, V, device="cuda")
    am2 = ops.argmax_rows(logits, V, scores=sc, suppress_col=int(am[0]))
    c0 = int(am[0])
    keep = torch.ones(V, dtype=torch.bool, device="cuda"); keep[c0] = False
    assert am2[0] != am[0] and torch.isinf(sc[:, c0]).all() and torch.equal(sc[1][keep], logits[1].float()[keep])
    loss = ops.cross_entropy(logits, V, tg, 0, write_grad=True)
    assert abs(float(loss) - float(ref)) < 2e-3 * float(ref)
    close(logits, lf.grad, 1e-2)
    # tie -> first index
    t = torch.zeros(2, 64, device="cuda", dtype=torch.bfloat16)
    t[0, 5] = t[0, 9] = 3.0
    assert ops.argmax_rows(t, 64).tolist() == [5, 0]
    # rows that are not 16-byte aligned (odd leading dimension / offset view) take the scalar path; ragged tails; ties across vectors
    for V2, ld2, off in ((1001, 1003, 1), (32262, 32264, 0), (77, 80, 3), (8, 8, 0), (4099, 4104, 0)):
        b2 = rnd(9, ld2 + 8, seed=V2, scale=2.0)
        lg = b2[:, off:off + V2]
        lg[3, V2 - 1] = 50.0                                  # maximum in the ragged tail
        lg[4, 2] = lg[4, V2 // 2] = 60.0                      # tie: first index wins
        sc2 = torch.empty(9, V2, device="cuda")
        got = ops.argmax_rows(lg, V2, scores=sc2)
        assert torch.equal(got, lg.float().argmax(-1)), (V2, ld2, off)
        assert int(got[3]) == V2 - 1 and int(got[4]) == 2
        assert torch.equal(sc2, lg.float())


def test_adamw_matches_torch():
    n = 10007
    p0, g = torch.randn(n, device="cuda"), torch.randn(n, device="cuda")
    ref = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref], lr=2e-3)
    p, m, v = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    pb = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for step in range(1, 4):
        ref.grad = g * step
        opt.step()
        ops.adamw_(p, g * step * 2, m, v, pb, 2e-3, 0.9, 0.999, 1e-8, 0.01, step, grad_scale=0.5)
    torch.testing.assert_close(p, ref.detach(), atol=1e-6, rtol=1e-5)
    assert torch.equal(pb, p.bfloat16())
    # unaligned views take the scalar loop and must give bit-identical results to the 16-byte path
    big = [torch.zeros(n + 1, device="cuda") for _ in range(3)]
    pu, mu, vu = (t[1:] for t in big)
    pu.copy_(p0)
    pbu = torch.empty(n + 1, device="cuda", dtype=torch.bfloat16)[1:]
    pa, ma, va = p0.clone(), torch.zeros(n, device="cuda"), torch.zeros(n, device="cuda")
    for step in range(1, 4):
        ops.adamw_(pu, g * step * 2, mu, vu, pbu, 2e-3, 0.9, 0.999, 1e-8, 0.01, step, grad_scale=0.5)
        ops.adamw_(pa, g * step * 2, ma, va, pb, 2e-3, 0.9, 0.999, 1e-8, 0.01, step, grad_scale=0.5)
    assert torch.equal(pu, pa) and torch.equal(mu, ma) and torch.equal(vu, va) and torch.equal(pbu, pb)


def test_embed_splice_fwd_bwd_and_layout():
    B, L, d, V, Np = 2, 40, 64, 100, 9
    ids = torch.randint(3, 90, (B, L), device="cuda")
    ids[:, 2] = 97; ids[:, 3:3 + Np] = 96; ids[:, 3 + Np] = 98; ids[1, 30:] = 0
    info = ops.point_layout(ids, 96, 97, 98, Np).cpu()
    assert info[0, :7].tolist() == [1, 1, Np, 2, 1, 1, 3]
    emb, feats = rnd(V, d, seed=1), rnd(B, Np, d, seed=2)
    pos = torch.tensor([3, 3], device="cuda", dtype=torch.int32)
    x = ops.embed_splice(ids, emb, feats, pos, Np).view(B, L, d)
    ref = emb[ids]
    ref[:, 3:3 + Np] = feats
    assert torch.equal(x, ref)
    dx = rnd(B * L, d, seed=3)
    dE = torch.zeros(V, d, device="cuda")
    dF = torch.empty(B, Np, d, device="cuda", dtype=torch.bfloat16)
    ops.embed_splice_bwd(ids, dx, pos, Np, dE, dF, 0)
    assert torch.equal(dF, dx.view(B, L, d)[:, 3:3 + Np])
    refE = torch.zeros(V, d, device="cuda")
    keep = torch.ones(B, L, dtype=torch.bool, device="cuda"); keep[:, 3:3 + Np] = False; keep &= ids != 0
    refE.index_put_((ids[keep],), dx.view(B, L, d)[keep].float(), accumulate=True)
    torch.testing.assert_close(dE, refE, atol=1e-4, rtol=1e-4)
    assert float(dE[0].abs().max()) == 0.0


def test_pool_transpose_colsum_copy():
    f = rnd(8 * 32, 256, seed=1)
    mx = f.view(8, 32, 256).float().max(1).values
    assert torch.equal(ops.group_maxpool(f, 32, False).float(), mx)
    cat = ops.group_maxpool(f, 32, True)
    assert torch.equal(cat[:, 256:], f) and torch.equal(cat[:, :256].float(), mx.repeat_interleave(32, 0))
    x = rnd(77, 130, seed=2)
    assert torch.equal(ops.transpose(x), x.t())
    cs = torch.zeros(130, device="cuda")
    ops.colsum(x, cs)
    torch.testing.assert_close(cs, x.float().sum(0), atol=1e-3, rtol=1e-4)
    src = rnd(3, 20, 64, seed=3)
    dst = torch.zeros(3, 5, 64, device="cuda", dtype=torch.bfloat16)
    ops.copy_rows(src[:, 7:12, :], dst)
    assert torch.equal(dst, src[:, 7:12])


# ------------------------------------------------------------------ trainable point encoder kernels (encoder_train.cu)
def test_layernorm_bwd():
    rows, d = 1026, 384
    x, dy, dres = rnd(rows, d, seed=1), rnd(rows, d, seed=2), rnd(rows, d, seed=3)
    gamma = (torch.randn(d, device="cuda") * 0.3 + 1).float()
    beta = torch.randn(d, device="cuda")
    xf, gf, bf = x.float().requires_grad_(True), gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    y = F.layer_norm(xf, (d,), gf, bf, 1e-5)
    y.backward(dy.float())
    dg, db = torch.zeros(d, device="cuda"), torch.zeros(d, device="cuda")
    dx = ops.layernorm_bwd(dy, x, gamma, dgamma=dg, dbeta=db, dres=dres)
    close(dx, xf.grad + dres.float(), 1.5e-2)
    close(dg, gf.grad, 5e-3)
    close(db, bf.grad, 5e-3)


def test_colsum2_is_deterministic_and_exact():
    """Batch statistics of the train-mode BatchNorm: same bits on every call (fixed summation order over the row slices), equal
    to an fp64 sum within fp32 round-off; accumulates onto the outputs."""
    for R, C in ((131072, 384), (16384, 128), (300, 70), (7, 40)):
        z, w = rnd(R, C, seed=R % 97), rnd(R, C, seed=R % 89)
        outs = []
        for _ in range(3):
            s0, s1 = torch.ones(C, device="cuda"), torch.full((C,), 2.0, device="cuda")
            ops.colsum2(z, w, s0, s1)
            outs.append((s0, s1))
        assert all(torch.equal(outs[0][0], o[0]) and torch.equal(outs[0][1], o[1]) for o in outs[1:])
        torch.testing.assert_close(outs[0][0].double(), 1 + z.double().sum(0), rtol=1e-5, atol=1e-3)
        torch.testing.assert_close(outs[0][1].double(), 2 + (z.double() * w.double()).sum(0), rtol=1e-5, atol=1e-3)


def test_batchnorm_train_forward_backward():
    """colsum2 + bn_coeffs + affine_cols + relu_mask == nn.BatchNorm1d(train) + ReLU and their autograd."""
    R, C = 4096, 128
    z = rnd(R, C, seed=4, scale=2.0)
    dy = rnd(R, C, seed=5)
    bn = torch.nn.BatchNorm1d(C).cuda().train()
    with torch.no_grad():
        bn.weight.copy_(torch.randn(C) * 0.2 + 1); bn.bias.copy_(torch.randn(C) * 0.1)
    gamma, beta = bn.weight.detach().clone(), bn.bias.detach().clone()
    rm, rv = bn.running_mean.clone(), bn.running_var.clone()
    zf = z.float().requires_grad_(True)
    yref = torch.relu(bn(zf))
    yref.backward(dy.float())
    f = lambda: torch.zeros(C, device="cuda")
    s0, s1, mean, rstd, A, B, Cc = f(), f(), f(), f(), f(), f(), f()
    ops.colsum2(z, None, s0, s1)
    ops.bn_coeffs(0, s0, s1, R, gamma, beta, mean, rstd, A, None, Cc, running_mean=rm, running_var=rv)
    y = ops.affine_cols(z, A, Cc, relu=True)
    close(y, yref, 1e-2)
    close(rm, bn.running_mean, 1e-3)
    close(rv, bn.running_var, 1e-3)
    dym = ops.relu_mask(dy, y)
    t0, t1, dg, db = f(), f(), f(), f()
    ops.colsum2(dym, z, t0, t1)
    ops.bn_coeffs(1, t0, t1, R, gamma, None, mean, rstd, A, B, Cc, dgamma=dg, dbeta=db)
    dz = ops.affine_cols(dym, A, Cc, y=z, B=B)
    close(dz, zf.grad, 2e-2)
    close(dg, bn.weight.grad, 1e-2)
    close(db, bn.bias.grad, 1e-2)


def test_group_maxpool_bwd_and_droppath():
    G, n, C = 40, 32, 256
    f = rnd(G * n, C, seed=6)
    ff = f.float().view(G, n, C).requires_grad_(True)
    # plain: max over n
    dpool = rnd(G, C, seed=7)
    ff.max(1)[0].backward(dpool.float())
    df = ops.group_maxpool_bwd(f, n, dpool=dpool)
    assert torch.equal(df.float().view(G, n, C), ff.grad)
    # concat: cat(global.expand, f)
    ff.grad = None
    dcat = rnd(G * n, 2 * C, seed=8)
    cat = torch.cat([ff.max(1, keepdim=True)[0].expand(-1, n, -1), ff], -1)
    cat.backward(dcat.float().view(G, n, 2 * C))
    df2 = ops.group_maxpool_bwd(f, n, dcat=dcat)
    close(df2.view(G, n, C), ff.grad, 1e-2)
    # DropPath scale: x + s[b] * y
    B, L, d = 3, 513, 384
    x, y = rnd(B * L, d, seed=9), rnd(B * L, d, seed=10)
    sc = torch.tensor([0.0, 1.0 / 0.9, 1.0 / 0.9], device="cuda")
    out = ops.scale_rows_add(y, sc, L, x=x)
    ref = x.float().view(B, L, d) + sc[:, None, None] * y.float().view(B, L, d)
    close(out.view(B, L, d), ref, 1e-2)
    close(ops.scale_rows_add(y, sc, L).view(B, L, d), sc[:, None, None] * y.float().view(B, L, d), 1e-2)


def test_embed_rows_pack_and_deterministic_scatter():
    """The sparse embedding-gradient exchange (trainer._embed_exchange): pack == the non-point rows in order; the deterministic
    scatter == the atomicAdd scatter of esb_embed_splice_bwd up to fp32 summation order, is bit-reproducible, and skips pad /
    out-of-range ids."""
    B, L, d, Np, V = 3, 40, 64, 7, 500
    g = torch.Generator(device="cuda").manual_seed(3)
    ids = torch.randint(1, 30, (B, L), device="cuda", generator=g)             # few distinct ids: many duplicates
    ids[:, -5:] = 0                                                              # pad tail
    ids[1, 3] = V + 5                                                            # out of range: ignored
    dx = rnd(B * L, d, seed=4)
    pos = torch.tensor([2, 10, 5], device="cuda", dtype=torch.int32)
    rows, rid = ops.embed_rows_pack(ids, dx, pos, Np, L - Np, 0)
    for b in range(B):
        keep = [l for l in range(L) if not (int(pos[b]) <= l < int(pos[b]) + Np)]
        assert torch.equal(rows.view(B, L - Np, d)[b], dx.view(B, L, d)[b, keep])
        assert torch.equal(rid.view(B, L - Np)[b], ids[b, keep])
    ref = torch.zeros(V, d, device="cuda")
    ops.embed_splice_bwd(ids, dx, pos, Np, ref, torch.empty(B, Np, d, device="cuda", dtype=torch.bfloat16), 0, vocab=V)
    out1 = torch.zeros(V, d, device="cuda")
    ops.embed_scatter_rows(rid, rows, out1, 0)
    out2 = torch.zeros(V, d, device="cuda")
    ops.embed_scatter_rows(rid, rows, out2, 0)
    assert torch.equal(out1, out2)
    torch.testing.assert_close(out1, ref, atol=1e-4, rtol=1e-5)
    assert float(out1[0].abs().max()) == 0.0
    exp = torch.zeros(V, d, device="cuda", dtype=torch.float64)
    ok = (rid > 0) & (rid < V)
    exp.index_add_(0, rid[ok], rows[ok].double())
    torch.testing.assert_close(out1.double(), exp, atol=1e-4, rtol=1e-5)
    # mixed batch (one sample without a point block): rmax = L, unused slots carry pad
    pos2 = torch.tensor([2, -1, 5], device="cuda", dtype=torch.int32)
    rows2, rid2 = ops.embed_rows_pack(ids, dx, pos2, Np, L, 0)
    assert torch.equal(rows2.view(B, L, d)[1], dx.view(B, L, d)[1]) and torch.equal(rid2.view(B, L)[1], ids[1])
    assert bool((rid2.view(B, L)[0, L - Np:] == 0).all()) and float(rows2.view(B, L, d)[0, L - Np:].abs().max()) == 0.0
    # accumulate on top of existing content, and many rows (several bitmask words)
    R = 5000
    ridb = torch.randint(0, 64, (R,), device="cuda", generator=g)
    rowsb = rnd(R, d, seed=6)
    acc = torch.ones(V, d, device="cuda")
    ops.embed_scatter_rows(ridb, rowsb, acc, 0)
    expb = torch.ones(V, d, device="cuda", dtype=torch.float64)
    okb = ridb > 0
    expb.index_add_(0, ridb[okb], rowsb[okb].double())
    torch.testing.assert_close(acc.double(), expb, atol=2e-3, rtol=1e-5)
    # the sorted variant (what the trainer calls): same sums in the same order -> same bits, also beyond 65536 rows
    acc2 = torch.ones(V, d, device="cuda")
    ops.embed_scatter_sorted(ridb, rowsb, acc2, 0)
    assert torch.equal(acc2, acc)
    out3 = torch.zeros(V, d, device="cuda")
    ops.embed_scatter_sorted(rid, rows, out3, 0)
    assert torch.equal(out3, out1)
    R2 = 70000
    ridc = torch.randint(0, V + 20, (R2,), device="cuda", generator=g)
    rowsc = rnd(R2, d, seed=8)
    big = torch.zeros(V, d, device="cuda")
    ops.embed_scatter_sorted(ridc, rowsc, big, 0)
    expc = torch.zeros(V, d, device="cuda", dtype=torch.float64)
    okc = (ridc > 0) & (ridc < V)
    expc.index_add_(0, ridc[okc], rowsc[okc].double())
    torch.testing.assert_close(big.double(), expc, atol=5e-3, rtol=1e-5)


def test_attention_bwd_on_a_query_row_range():
    """ops.attention_bwd(q_rows=(r0, r1)): only rows r0..r1-1 of every sample were queries (the top decoder layer of a training
    step reads just the loss rows, llm._top_layer_rows) - dq of the other rows is zero, dk / dv cover all keys.  vs fp32 torch."""
    B, H, hd, L, r0, r1 = 2, 3, 128, 300, 211, 299
    d = H * hd
    qkv = rnd(B, L, 3, H, hd, seed=31)
    kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    kmask[0, 100:109] = 0
    q = qkv[:, r0:r1, 0]
    o, lse = ops.attention(q, qkv[:, :, 1], qkv[:, :, 2], causal=True, kmask=kmask, q_pos0=r0, want_lse=True)
    qf = qkv.float().requires_grad_(True)
    ref_o, ref_lse = _attn_ref(qf[:, r0:r1, 0], qf[:, :, 1], qf[:, :, 2], True, kmask, q_pos0=r0)
    close(o, ref_o, 1.5e-2)
    assert (lse - ref_lse).abs().max() < 2e-2
    do = rnd(B, r1 - r0, d, seed=32)
    ref_o.backward(do.float())
    dqkv = ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask, q_rows=(r0, r1))
    for j in range(3):
        close(dqkv[:, :, j], qf.grad[:, :, j], 2.5e-2)
    assert float(dqkv[:, :r0, 0].abs().max()) == 0.0 and float(dqkv[:, r1:, 0].abs().max()) == 0.0
    # with the fused inverse rotary embedding == unfused + rope kernel
    from egoscaler_b200.llm import rope_tables
    cos, sin = rope_tables(hd, 10000.0, 512, "cuda")
    ref = ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask, q_rows=(r0, r1))
    ops.rope_(ref.view(B * L, 3 * d), L, 2 * H, hd, cos, sin, 0, inverse=True)
    got = ops.attention_bwd(qkv, o, lse, do, causal=True, kmask=kmask, q_rows=(r0, r1), rope=(cos, sin))
    assert torch.equal(got, ref)
