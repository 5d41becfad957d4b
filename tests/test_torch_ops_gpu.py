"""GPU: the ``torch.library`` custom ops (egoscaler_b200/torch_ops.py, ``torch.ops.esb.*``) - schema / fake-tensor / autograd
registration through ``torch.library.opcheck`` and values + gradients against plain PyTorch fp32 references of the same ops
(HF LlamaRMSNorm / LlamaMLP activation / sdpa attention / nn.Linear / F.cross_entropy; np.digitize for the tokenizer)."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from egoscaler_b200 import torch_ops  # noqa: F401  (registers torch.ops.esb.*)
from oracle import traj_tokenizer as tt


def rnd(*shape, seed, scale=1.0, grad=False):
    g = torch.Generator(device="cuda").manual_seed(seed)
    t = (torch.randn(*shape, device="cuda", generator=g) * scale).bfloat16()
    return t.requires_grad_(grad)


def close(a, b, tol):
    a, b = a.float(), b.float()
    assert (a - b).abs().max().item() <= tol * max(b.abs().max().item(), 1e-6), ((a - b).abs().max().item(), b.abs().max().item())


def test_opcheck_schema_fake_and_autograd_registration():
    x, w = rnd(96, 64, seed=1, grad=True), rnd(40, 64, seed=2, scale=0.1, grad=True)
    tests = ("test_schema", "test_faketensor", "test_autograd_registration")
    torch.library.opcheck(torch.ops.esb.linear.default, (x, w), test_utils=tests)
    torch.library.opcheck(torch.ops.esb.swiglu.default, (rnd(33, 128, seed=3, grad=True),), test_utils=tests)
    torch.library.opcheck(torch.ops.esb.rmsnorm_fwd.default, (x, rnd(64, seed=4, grad=True), 1e-6), test_utils=tests)
    torch.library.opcheck(torch.ops.esb.traj_quantize.default, (torch.rand(100, 6, device="cuda") * 2 - 1, 256, 32006), test_utils=tests[:2])


def test_linear_rmsnorm_swiglu_chain_gradients_match_torch():
    """A LLaMA MLP block written with the custom ops vs the same block in fp32 torch: y = down(silu(gate(n)) * up(n)), n = rmsnorm(x)."""
    T, d, f = 300, 256, 688
    x = rnd(T, d, seed=1, grad=True)
    g = rnd(d, seed=2, scale=0.1)
    g = (g.float() + 1.0).bfloat16().requires_grad_(True)
    wgu = rnd(2 * f, d, seed=3, scale=d ** -0.5, grad=True)
    wd = rnd(d, f, seed=4, scale=f ** -0.5, grad=True)
    dy = rnd(T, d, seed=5)
    n = torch_ops.rmsnorm(x, g, 1e-6)
    y = torch_ops.linear(torch_ops.swiglu(torch_ops.linear(n, wgu)), wd)
    y.backward(dy)
    xf, gf, wguf, wdf = (t.detach().float().requires_grad_(True) for t in (x, g, wgu, wd))
    nf = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + 1e-6) * gf
    guf = nf @ wguf.t()
    yf = (F.silu(guf[:, :f]) * guf[:, f:]) @ wdf.t()
    yf.backward(dy.float())
    close(y, yf, 2e-2)
    for a, b in ((x, xf), (wgu, wguf), (wd, wdf), (g, gf)):
        assert a.grad is not None and a.grad.dtype == a.dtype
        cos = F.cosine_similarity(a.grad.float().flatten(), b.grad.flatten(), dim=0).item()
        assert cos > 0.999, cos


def test_attention_and_cross_entropy_ops_match_torch():
    B, L, H, hd, V = 2, 200, 2, 128, 1000
    qkv = rnd(B, L, 3, H, hd, seed=7, grad=True)
    kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    kmask[1, 90:95] = 0
    o = torch_ops.attention(qkv, kmask, True)
    do = rnd(B, L, H * hd, seed=8)
    o.backward(do)
    qf = qkv.detach().float().requires_grad_(True)
    q, k, v = (qf[:, :, j].transpose(1, 2) for j in range(3))
    m = torch.tril(torch.ones(L, L, dtype=torch.bool, device="cuda"))[None, None] & kmask.bool()[:, None, None, :]
    of = F.scaled_dot_product_attention(q, k, v, attn_mask=m).transpose(1, 2).reshape(B, L, H * hd)
    of.backward(do.float())
    close(o, of, 1.5e-2)
    close(qkv.grad, qf.grad, 2.5e-2)
    logits = rnd(64, V, seed=9, scale=2.0, grad=True)
    tg = torch.randint(1, V, (64,), device="cuda")
    tg[::5] = 0
    loss = torch_ops.cross_entropy(logits, tg, 0)
    (loss * 3.0).backward()
    lf = logits.detach().float().requires_grad_(True)
    ref = F.cross_entropy(lf, tg, ignore_index=0)
    (ref * 3.0).backward()
    assert abs(loss.item() - ref.item()) < 2e-3 * ref.item()
    close(logits.grad, lf.grad, 1.5e-2)


def test_traj_token_ops_bit_exact():
    x = torch.cat([torch.rand(4096, device="cuda") * 2.1 - 1.05, torch.linspace(-1, 1, 256, dtype=torch.float64).float().cuda(),
                   torch.tensor([float("nan"), float("inf"), float("-inf")], device="cuda")])
    x = x[: x.numel() // 6 * 6].view(-1, 6)
    ids = torch.ops.esb.traj_quantize(x, 256, 32006)
    assert np.array_equal(ids.cpu().numpy() - 32006, tt.discretize_action(x.cpu().numpy(), 256))
    ok = (ids - 32006 >= 0).all(dim=1)
    back = torch.ops.esb.traj_dequantize(ids[ok], 256, 32006)
    assert np.array_equal(back.cpu().numpy(), tt.token_to_action((ids[ok] - 32006).cpu().numpy(), 256).astype(np.float32))
