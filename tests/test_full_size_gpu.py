"""GPU: the hot path at BASELINE.json's FULL size (LLaMA-7B trunk, PointBERT v1.2, 513 point tokens, vocab 32262) checked
through size-independent properties - the reference itself cannot run here at this size, so parity at full size rests on:

  * cached decode == teacher-forced re-forward of the same sequence (KV cache, skinny split-K GEMMs and decode attention
    against the prefill kernels), CUDA-graph decode == eager decode bit for bit, seeded sampling reproducible;
  * rows of a decode batch are independent (batch of 4 == four batches of 1);
  * data-parallel linearity of the training step: token-weighted loss and every gradient of the batch {a, b} equal the
    combination of the single-sample runs - the identity the batch-sharded N-GPU path relies on;
  * AdamW steps on a fixed batch lower the loss.
Arithmetic parity against the real reference is pinned at the tiny size by the goldens (test_model_gpu / test_variants_gpu).
"""
import math
import os
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from egoscaler_b200 import synth
from egoscaler_b200.config import TrajVLMConfig
from egoscaler_b200.trainer import Trainer


def _rel(a: torch.Tensor, b: torch.Tensor) -> float:
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.fixture(scope="module")
def full():
    import bench
    cfg = TrajVLMConfig.full()
    m = bench.build_model(torch, cfg, torch.device("cuda", 0))
    return m


def _decode_inputs(cfg, B, seed):
    bt = synth.make_batch(cfg, B, seed=seed)
    g = torch.Generator().manual_seed(seed)
    start = torch.randint(0, bt["pcrgbs"].shape[1], (B,), generator=g)
    return bt, start


def test_full_size_decode_matches_reforward_and_graph(full):
    m, cfg = full, full.cfg
    B, T = 2, 10
    bt, start = _decode_inputs(cfg, B, 5)
    P = bt["prompts"].shape[1]
    kw = dict(input_ids=bt["prompts"], attention_mask=bt["prompt_masks"], point_clouds=bt["pcrgbs"], max_length=T, do_sample=False,
              fps_start=start, suppress_eos=True)
    a = m.generate(use_cuda_graph=True, **kw)
    b = m.generate(use_cuda_graph=False, **kw)
    assert a.sequences.shape == (B, P + T)
    assert torch.equal(a.sequences, b.sequences)
    for x, y in zip(a.scores, b.scores):
        assert torch.equal(x, y)
    # teacher-forced forward over prompt + generated tokens: logits at P-1+t must reproduce decode step t
    seq = a.sequences[:, :P + T - 1]
    mask = torch.cat([bt["prompt_masks"].cuda().to(torch.uint8), torch.ones(B, T - 1, device="cuda", dtype=torch.uint8)], 1)
    m.eval()
    full_logits = m(input_ids=seq, attention_mask=mask, point_clouds=bt["pcrgbs"].cuda(), fps_start=start.cuda()).logits.float()
    gen = a.sequences[:, P:]
    for t in range(T):
        ref = full_logits[:, P - 1 + t, :].clone()
        got = a.scores[t].float().clone()
        ref[:, cfg.eos_token_id] = 0.0
        got[:, cfg.eos_token_id] = 0.0                                     # suppressed column holds -inf in the scores
        assert _rel(got, ref) < 3e-2, (t, _rel(got, ref))
        # the emitted id is the teacher-forced arg-max wherever that arg-max is decided by more than the kernel noise
        noise = float((got - ref).abs().max())
        ref[:, cfg.eos_token_id] = float("-inf")
        top2 = ref.topk(2, dim=-1)
        for r in range(B):
            if float(top2.values[r, 0] - top2.values[r, 1]) > 4 * noise:
                assert int(gen[r, t]) == int(top2.indices[r, 0])


def test_full_size_training_step_matches_cpu_oracle(full):
    """The kernels the benchmark times (BN=256 grouped-walk GEMM, tcgen05 attention fwd/bwd at 32 heads x 128, L=696)
    against the CPU oracle port on the SAME full-size weights: B=1 sample of BASELINE configs[1] with mid-sequence
    padding (dataset.py:150-194).  Logits on the loss rows <= 3e-2 * max|logit|, loss 2e-2, gradients of lm_head /
    model.norm cosine >= 0.999, of point_proj.* / the touched embed_tokens rows (32 bf16 layers of backward) >= 0.997
    (train.py:165-184, model_arch.py:33-51).
    The oracle needs ~30 GB of host RAM and ~10-20 s."""
    import psutil
    from oracle.traj_vlm import LlmCfg, PointCfg, TrajVLMOracle, train_loss
    if psutil.virtual_memory().available < 45e9:
        pytest.skip("host RAM < 45 GB: the fp32 full-size oracle does not fit")
    m, cfg = full, full.cfg
    bt = synth.make_batch(cfg, 1, seed=21, random_desc_len=True)
    assert int((bt["tokens"][0, :540] == cfg.pad_token_id).sum()) > 0          # mid-sequence padding is present
    P = bt["prompts"].shape[1]
    L = bt["tokens"].shape[1]
    start = torch.tensor([4321])
    torch.set_num_threads(os.cpu_count() or 1)
    with torch.device("meta"):
        orc = TrajVLMOracle(LlmCfg.full(), PointCfg())
    orc = orc.to_empty(device="cpu")
    sd = m.state_dict()
    with torch.no_grad():
        for k, v in orc.state_dict().items():
            v.copy_(sd[k].to("cpu", v.dtype))
    del sd
    orc.configure_trainable()
    orc.train()
    orc.model.layers.eval()
    orc.model.point_backbone.eval()
    ref_logits = orc(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=start, training=True)
    ref_loss = train_loss(ref_logits, bt["tokens"], P, cfg.pad_token_id)
    ref_loss.backward()
    ref_rows = ref_logits.detach()[0, P - 1:L - 1].clone()
    del ref_logits
    # ---- CUDA path: logits of the loss rows, then loss + gradients
    m.eval()
    got_rows = m(input_ids=bt["tokens"], attention_mask=bt["attention_masks"], point_clouds=bt["pcrgbs"],
                 fps_start=start).logits[0, P - 1:L - 1].detach().float().cpu()
    scale = float(ref_rows.abs().max())
    err = float((got_rows - ref_rows).abs().max())
    assert err <= 3e-2 * scale, (err, scale)
    assert _rel(got_rows, ref_rows) < 4e-2                                  # bf16 through 32 layers vs fp32: measured 2.8e-2
    m.train()
    tr = Trainer(m, lr=2e-5)
    tr.zero_grad()
    loss = tr.forward_backward(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], P, start)
    print(f"full-size loss: cuda {float(loss):.4f}  oracle {float(ref_loss):.4f}  logits max-err {err:.3e} of {scale:.2f}")
    assert abs(float(loss) - float(ref_loss)) <= 2e-2 * float(ref_loss), (float(loss), float(ref_loss))
    refg = {k: p.grad for k, p in orc.named_parameters() if p.grad is not None}
    checked = 0
    for k in tr.names:
        g, r = tr.grad[k].float().cpu(), refg[k].float()
        if k == "model.embed_tokens.weight":
            rows = torch.nonzero(r.norm(dim=1) > 0).flatten()
            assert torch.equal(rows, torch.nonzero(g.norm(dim=1) > 0).flatten())
            g, r = g[rows], r[rows]
        cos = float((g.double().flatten() @ r.double().flatten()) / (g.double().norm() * r.double().norm()).clamp_min(1e-30))
        # lm_head / model.norm see one bf16 GEMM; embed_tokens / point_proj gradients cross the 32-layer bf16 backward
        # (measured cosine 0.9989 for embed_tokens on the first B200 run of this test)
        through_trunk = k.startswith("model.embed_tokens") or k.startswith("model.point_proj")
        assert cos >= (0.997 if through_trunk else 0.999), (k, cos)
        assert _rel(g, r) <= (8e-2 if through_trunk else 5e-2), (k, _rel(g, r))
        checked += 1
    assert checked >= 9
    del tr, orc
    m.eval()
    torch.cuda.empty_cache()


def test_full_size_decode_rows_independent(full):
    m, cfg = full, full.cfg
    B, T = 4, 6
    bt, start = _decode_inputs(cfg, B, 6)
    kw = dict(max_length=T, do_sample=False, suppress_eos=True)
    allr = m.generate(input_ids=bt["prompts"], attention_mask=bt["prompt_masks"], point_clouds=bt["pcrgbs"], fps_start=start, **kw)
    for r in range(B):
        one = m.generate(input_ids=bt["prompts"][r:r + 1], attention_mask=bt["prompt_masks"][r:r + 1],
                         point_clouds=bt["pcrgbs"][r:r + 1], fps_start=start[r:r + 1], **kw)
        # step 0 sees identical inputs; later steps only while the emitted ids still agree
        for t in range(T):
            x, y = allr.scores[t][r].float(), one.scores[t][0].float()
            x[cfg.eos_token_id] = 0.0
            y[cfg.eos_token_id] = 0.0
            assert _rel(x, y) < 2e-2, (r, t, _rel(x, y))
            if int(allr.sequences[r, -T + t]) != int(one.sequences[0, -T + t]):
                break


def test_full_size_sampling_seeded(full):
    m, cfg = full, full.cfg
    bt, start = _decode_inputs(cfg, 2, 7)
    kw = dict(input_ids=bt["prompts"], attention_mask=bt["prompt_masks"], point_clouds=bt["pcrgbs"], max_length=8, fps_start=start,
              suppress_eos=True)
    a = m.generate(seed=1, use_cuda_graph=True, **kw)                      # reference defaults: top_k=50, top_p=0.95
    b = m.generate(seed=1, use_cuda_graph=False, **kw)
    c = m.generate(seed=2, use_cuda_graph=True, **kw)
    assert torch.equal(a.sequences, b.sequences) and not torch.equal(a.sequences, c.sequences)
    P = bt["prompts"].shape[1]
    for t, sc in enumerate(a.scores):
        kept = torch.isfinite(sc).sum(-1)
        assert bool((kept >= 1).all()) and bool((kept <= 58).all())
        assert bool(torch.isfinite(sc[torch.arange(2), a.sequences[:, P + t]]).all())


def test_full_size_train_step_linearity_and_descent(full):
    m, cfg = full, full.cfg
    m.train()
    tr = Trainer(m, lr=1e-4)
    bt = synth.make_batch(cfg, 2, seed=11, random_desc_len=True)
    P = bt["prompts"].shape[1]
    start = torch.tensor([17, 4242])

    def run(rows):
        tr.zero_grad()
        sel = {k: v[rows] for k, v in bt.items()}
        loss = tr.forward_backward(sel["tokens"], sel["attention_masks"], sel["pcrgbs"], P, start[rows])
        n = int((sel["tokens"][:, P:] != cfg.pad_token_id).sum())
        return float(loss), n, {k: g.float().clone() for k, g in tr.grad.items()}

    l_ab, n_ab, g_ab = run(slice(0, 2))
    l_a, n_a, g_a = run(slice(0, 1))
    l_b, n_b, g_b = run(slice(1, 2))
    assert n_a + n_b == n_ab and n_a > 0 and n_b > 0
    assert 9.0 < l_ab < 13.0                                               # ~ ln(32262) + sigma^2/2 at random init
    assert abs(l_ab * n_ab - (l_a * n_a + l_b * n_b)) < 2e-3 * l_ab * n_ab
    checked = 0
    for k in g_ab:
        comb = (g_a[k] * n_a + g_b[k] * n_b) / n_ab
        if float(comb.norm()) == 0.0:
            assert float(g_ab[k].norm()) == 0.0
            continue
        assert _rel(g_ab[k], comb) < 2e-2, (k, _rel(g_ab[k], comb))
        checked += 1
    assert checked >= 6                                                    # embed, head, norm, projector weights + biases
    dev_bt = {k: v.cuda() for k, v in bt.items()}
    losses = [float(tr.train_step(dev_bt, start.cuda())) for _ in range(4)]
    assert all(math.isfinite(x) for x in losses)
    assert losses[-1] < losses[0] - 1e-3, losses
    m.eval()
