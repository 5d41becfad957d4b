"""GPU: the sampling decode step (csrc/sampling.cu) vs the oracle restatement of the HF processor chain the reference's
generate() drives (model_arch.py:94-108 -> HF logits_process.py, oracle/sampling.py pinned against the real classes)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops, synth
from egoscaler_b200.config import TrajVLMConfig
from egoscaler_b200 import decode  # noqa: F401
from oracle.sampling import hf_filter, inverse_cdf_draw

CASES = [(32262, 50, 0.95, 1.0, 1.0), (32262, 50, 0.95, 0.7, 1.3), (1000, 5, 0.5, 1.0, 1.0), (4096, 200, 0.99, 1.5, 1.1),
         (32262, 1, 0.95, 1.0, 1.0), (32262, 1024, 0.9, 1.0, 1.0), (777, 50, 1.0, 1.0, 1.0)]


@pytest.mark.parametrize("V,k,p,temp,pen", CASES)
def test_filter_and_draw_match_oracle(V, k, p, temp, pen):
    g = torch.Generator().manual_seed(V + k)
    rows = 32
    vp = (V + 7) // 8 * 8
    logits = torch.zeros(rows, vp, dtype=torch.bfloat16)
    logits[:, :V] = (torch.randn(rows, V, generator=g) * 3).bfloat16()
    hist = torch.randint(0, V, (rows, 40), generator=g)
    u = torch.rand(rows, generator=g)
    dl = logits.cuda()
    scores = torch.empty(rows, V, device="cuda", dtype=torch.float32)
    nk = torch.empty(rows, device="cuda", dtype=torch.int32)
    out = ops.sample_rows(dl[:, :V], V, u.cuda(), temp, k, p, pen, hist=hist.cuda(), hist_len=37, scores=scores, n_kept=nk)
    ref = hf_filter(logits[:, :V].float(), hist[:, :37], temp, k, p, pen)
    got = scores.cpu()
    # identical survivor multiset per row; element-wise disagreements only among value-ties at the smallest kept value
    assert torch.equal(torch.sort(ref, dim=-1)[0], torch.sort(got, dim=-1)[0])
    same = (ref == got) | (torch.isinf(ref) & torch.isinf(got))
    pre = hf_filter(logits[:, :V].float(), hist[:, :37], temp, k, 1.0, pen)
    for b in range(rows):
        if not bool(same[b].all()):
            kept_min = ref[b][torch.isfinite(ref[b])].min()
            assert bool((pre[b][~same[b]] == kept_min).all())
    assert torch.equal(nk.cpu().long(), torch.isfinite(got).sum(-1))
    # the draw: same inverse-CDF walk on the kernel's own survivors (an ulp of expf may move a boundary: allow 2 rows)
    pick = inverse_cdf_draw(got, u)
    o = out.cpu()
    assert int((pick != o).sum()) <= 2
    assert bool(torch.isfinite(got[torch.arange(rows), o]).all())          # every pick is a survivor


def test_draw_distribution():
    """20000 draws from one logit row: empirical frequencies == softmax over the survivors."""
    g = torch.Generator().manual_seed(5)
    V, n = 32262, 20000
    row = (torch.randn(1, V, generator=g) * 2).bfloat16()
    logits = row.expand(n, V).contiguous().cuda()
    u = torch.rand(n, generator=g).cuda()
    out = ops.sample_rows(logits, V, u, 1.0, 50, 0.95, 1.0).cpu()
    proc = hf_filter(row.float(), None, 1.0, 50, 0.95, 1.0)[0]
    probs = torch.softmax(proc, -1)
    freq = torch.bincount(out, minlength=V).float() / n
    assert float((freq - probs).abs().max()) < 0.012
    assert float(freq[probs == 0].sum()) == 0.0


def test_suppressed_column_and_step_indexing():
    g = torch.Generator().manual_seed(9)
    V, rows, T = 512, 4, 3
    logits = (torch.randn(rows, V, generator=g)).bfloat16()
    logits[:, 2] = 50.0                                                       # would always win
    u = torch.rand(T, rows, generator=g).cuda()
    scores = torch.zeros(T, rows, V, device="cuda")
    step = torch.tensor([5], device="cuda", dtype=torch.int32)
    out = ops.sample_rows(logits.cuda(), V, u, 1.0, 50, 0.95, 1.0, scores=scores, suppress_col=2, step_ptr=step, step_off=-4,
                          scores_step_stride=rows * V, u_step_stride=rows).cpu()
    assert bool((out != 2).all())
    assert float(scores[0].abs().sum()) == 0.0 and float(scores[2].abs().sum()) == 0.0
    assert bool(torch.isinf(scores[1][:, 2]).all())
    lg = logits.float().clone()
    lg[:, 2] = float("-inf")
    ref = hf_filter(lg, None, 1.0, 50, 0.95, 1.0)
    assert torch.equal(torch.sort(ref, dim=-1)[0], torch.sort(scores[1].cpu(), dim=-1)[0])
    assert torch.equal(inverse_cdf_draw(scores[1].cpu(), u[1].cpu()), out)


def _tiny_model():
    from egoscaler_b200.model import TrajPointLLMForCausalLM
    cfg = TrajVLMConfig.tiny()
    m = TrajPointLLMForCausalLM(cfg, "cuda")
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    return m


def test_generate_sampling_mode():
    """generate() with the reference's defaults (do_sample=True, top_k=50, top_p=0.95): shapes, HF-style scores,
    seeded reproducibility, graph == eager, num_return_sequences, repetition penalty."""
    m = _tiny_model()
    bt = synth.make_batch(m.cfg, 2, seed=1)
    P = bt["prompts"].shape[1]
    kw = dict(input_ids=bt["prompts"], attention_mask=bt["prompt_masks"], point_clouds=bt["pcrgbs"], max_length=12,
              fps_start=torch.zeros(2, dtype=torch.long), suppress_eos=True, seed=123)
    a = m.generate(use_cuda_graph=False, **kw)                             # do_sample defaults to True like the reference
    b = m.generate(use_cuda_graph=True, **kw)
    assert a.sequences.shape == (2, P + 12) and len(a.scores) == 12
    assert torch.equal(a.sequences, b.sequences)
    for x, y in zip(a.scores, b.scores):
        assert torch.equal(x, y)
    gen = a.sequences[:, P:].cpu()
    for t, sc in enumerate(a.scores):
        sc = sc.cpu()
        kept = torch.isfinite(sc).sum(-1)
        assert bool((kept >= 1).all()) and bool((kept <= 50 + 8).all())
        assert bool(torch.isfinite(sc[torch.arange(2), gen[:, t]]).all())      # emitted ids are survivors
    c = m.generate(use_cuda_graph=True, **{**kw, "seed": 124})
    assert not torch.equal(a.sequences, c.sequences)
    # num_return_sequences: rows repeat_interleave'd like HF
    d = m.generate(num_return_sequences=3, **kw)
    assert d.sequences.shape == (6, P + 12)
    assert torch.equal(d.sequences[:, :P].cpu(), bt["prompts"].repeat_interleave(3, 0))
    # repetition penalty > 1 (exact arithmetic is covered by test_filter_and_draw_match_oracle): the loop runs with the
    # growing history and penalised ids that survive carry the penalised score
    e = m.generate(repetition_penalty=1.5, **kw)
    assert e.sequences.shape == (2, P + 12) and len(e.scores) == 12
