"""GPU: the head_dim-128 kernels the benchmark times (tcgen05 attention forward / backward, decode attention) against
fixtures generated from the REAL reference on a trunk with the full-size head width (2 heads x 128, oracle/make_golden.py
``hd128`` / ``trained``):

* ``traj_vlm_hd128_unfrozen_lm.npz`` - logits, loss and the gradient of EVERY LLM-layer tensor (--unfreeze_language_model,
  model_arch.py:40-45): the attention backward kernels feed all of them;
* ``traj_vlm_hd128_trained.npz`` - the reference over-fitted on the batch so that its greedy decode is a real
  ``<pA> ... <tsep> ... <te> </s>`` trajectory with a top-2 margin > 8 at every step: all emitted ids must be IDENTICAL,
  and the on-device ids -> poses (+ rt2 scaler) must equal the reference's str_to_float on the decoded string bit for bit.
"""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops, synth
from egoscaler_b200.config import TrajVLMConfig
from egoscaler_b200.model import TrajPointLLMForCausalLM
from egoscaler_b200 import decode  # noqa: F401
from egoscaler_b200.trainer import Trainer
from conftest import GOLDEN, apply_trained_delta


def _cmp_grad(name, g, gv, cos_min=0.999, rel_max=3e-2):
    ref = gv[f"gradval/{name}"]
    got = g.flatten().cpu()[torch.from_numpy(gv[f"gradidx/{name}"])].numpy()
    cos = float((got * ref).sum() / (np.linalg.norm(got) * np.linalg.norm(ref) + 1e-30))
    rel = float(np.linalg.norm(got - ref) / (np.linalg.norm(ref) + 1e-30))
    assert cos >= cos_min and rel <= rel_max, (name, cos, rel)


def test_hd128_forward_and_every_llm_gradient_vs_reference():
    gv = np.load(os.path.join(GOLDEN, "traj_vlm_hd128_unfrozen_lm.npz"))
    cfg = TrajVLMConfig.tiny_hd128()
    assert cfg.head_dim == 128
    m = TrajPointLLMForCausalLM(cfg, "cuda", unfreeze_language_model=True)
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    bt = synth.make_batch(cfg, 2, seed=1)
    start = torch.from_numpy(gv["fps_start"])
    m.eval()
    L = m(input_ids=bt["tokens"], attention_mask=bt["attention_masks"], point_clouds=bt["pcrgbs"], fps_start=start).logits.detach().float().cpu()
    sub = L[:, gv["logit_rows"]][:, :, gv["logit_cols"]].numpy()
    assert np.abs(sub - gv["logits_sub"]).max() <= 2e-2 * np.abs(gv["logits_sub"]).max()
    m.train()
    tr = Trainer(m, lr=1e-3)
    tr.zero_grad()
    loss = tr.forward_backward(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], int(gv["prompt_len"]), start)
    assert abs(float(loss) - float(gv["loss"])) <= 2e-2 * float(gv["loss"])
    d, f = cfg.hidden_size, cfg.intermediate_size
    n = 0
    for li in range(cfg.num_hidden_layers):
        pre = f"model.layers.{li}."
        parts = {
            pre + "self_attn.q_proj.weight": tr.wgrad["wqkv"][li][:d], pre + "self_attn.k_proj.weight": tr.wgrad["wqkv"][li][d:2 * d],
            pre + "self_attn.v_proj.weight": tr.wgrad["wqkv"][li][2 * d:], pre + "self_attn.o_proj.weight": tr.wgrad["wo"][li],
            pre + "mlp.gate_proj.weight": tr.wgrad["wgu"][li][:f], pre + "mlp.up_proj.weight": tr.wgrad["wgu"][li][f:],
            pre + "mlp.down_proj.weight": tr.wgrad["wdown"][li], pre + "input_layernorm.weight": tr.wgrad["ln1"][li],
            pre + "post_attention_layernorm.weight": tr.wgrad["ln2"][li]}
        for name, g in parts.items():
            _cmp_grad(name, g.contiguous(), gv, cos_min=0.998, rel_max=6e-2 if "layernorm" in name else 4e-2)
            n += 1
    for k in tr.names:
        _cmp_grad(k, tr.grad[k], gv)
        n += 1
    assert n >= 27


@pytest.fixture(scope="module")
def trained():
    g = np.load(os.path.join(GOLDEN, "traj_vlm_hd128_trained.npz"))
    cfg = TrajVLMConfig.tiny_hd128()
    m = TrajPointLLMForCausalLM(cfg, "cuda")
    sd = apply_trained_delta(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), g)
    m.load_state_dict(sd, strict=True)
    return m, g, synth.make_batch(cfg, 2, seed=1)


def test_trained_logits_and_loss_vs_reference(trained):
    m, g, bt = trained
    start = torch.from_numpy(g["fps_start"])
    m.eval()
    L = m(input_ids=bt["tokens"], attention_mask=bt["attention_masks"], point_clouds=bt["pcrgbs"], fps_start=start).logits.detach().float().cpu()
    sub = L[:, g["logit_rows"]][:, :, g["logit_cols"]].numpy()
    assert np.abs(sub - g["logits_sub"]).max() <= 2e-2 * np.abs(g["logits_sub"]).max()
    P = int(g["prompt_len"])
    assert np.array_equal(L[:, P - 1:-1].argmax(-1).numpy()[bt["tokens"][:, P:].numpy() != 0],
                          bt["tokens"][:, P:].numpy()[bt["tokens"][:, P:].numpy() != 0])       # teacher-forced arg-max = targets


@pytest.mark.parametrize("graph,chain", [(False, False), (True, False), (False, True), (True, True)])
def test_trained_greedy_decode_emits_the_reference_trajectory(trained, graph, chain):
    """``chain``: the token step with the Linears between two attention calls as one persistent launch (ops.gemm_chain,
    llm._decode_step_chain) instead of one launch per Linear - same reference trajectory either way."""
    m, g, bt = trained
    ref = g["greedy_ids"]                                              # [2, 135]: 19 x (6 bins + <tsep>) + <te> + </s>
    T = 152
    keep = ops.decode_chain_enabled
    ops.decode_chain_enabled = chain
    try:
        out = m.generate(input_ids=bt["prompts"], attention_mask=bt["prompt_masks"], point_clouds=bt["pcrgbs"], max_length=T,
                         do_sample=False, fps_start=torch.from_numpy(g["fps_start"]), use_cuda_graph=graph)
    finally:
        ops.decode_chain_enabled = keep
    P = bt["prompts"].shape[1]
    gen = out.sequences[:, P:].cpu().numpy()
    n = ref.shape[1]
    assert float(g["greedy_step_top2gap"].min()) > 1.0 and len(np.unique(ref)) > 100
    assert np.array_equal(gen[:, :n], ref), "greedy ids differ from the reference's although every top-2 margin is > 8"
    assert (gen[:, n:] == 0).all()                                     # finished rows emit pad (HF greedy search)
    sl = torch.stack(out.scores, 1).float().cpu()[:, :n][:, :, g["logit_cols"]].numpy()
    refl = g["greedy_step_logits_sub"][:, :n]
    assert np.abs(sl - refl).max() <= 3e-2 * np.abs(refl).max()
    # post-processing of train.py:240-246 on device: un-scaled poses and the rt2 scaler (utils/utils.py:23-34) bit-exact
    ids = out.sequences[:, P:].contiguous()
    raw, cnt = ops.traj_ids_to_poses(ids, max_steps=20, pad_repeat=False)
    d_max, d_min = g["maxmin"].tolist()
    rt2, cnt2 = ops.traj_ids_to_poses(ids, max_steps=20, pad_repeat=False, scale=ops.SCALE_RT2, d_max=d_max, d_min=d_min)
    assert cnt.tolist() == [20, 20] and cnt2.tolist() == [20, 20]
    for b in range(2):
        assert np.array_equal(raw[b].cpu().numpy(), g[f"greedy_poses_raw/{b}"])
        assert np.array_equal(rt2[b].cpu().numpy(), g[f"greedy_poses_rt2/{b}"])


def test_top_layer_on_loss_rows_only_equals_the_full_layer():
    """llm.LlamaTrunk with ``out_rows``: the top layer computed for the loss rows only must give the same hidden rows and the same
    input gradient as the full computation (frozen trunk and --unfreeze_language_model weight gradients)."""
    cfg = TrajVLMConfig.tiny_hd128()
    m = TrajPointLLMForCausalLM(cfg, "cuda", unfreeze_language_model=True)
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    B, L, d, r0, r1 = 2, 260, cfg.hidden_size, 180, 259
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(B * L, d, device="cuda", generator=g).bfloat16()
    dsub = torch.randn(B * (r1 - r0), d, device="cuda", generator=g).bfloat16()
    kmask = torch.ones(B, L, device="cuda", dtype=torch.uint8)
    kmask[1, 50:58] = 0
    mk = lambda: {k: [torch.zeros(getattr(lw, k).shape, device="cuda", dtype=torch.float32) for lw in m.trunk.layers]
                  for k in ("wqkv", "wo", "wgu", "wdown", "ln1", "ln2")}
    # full computation, gradient only on the row range
    saved = []
    h_full = m.trunk.forward(x.clone(), B, L, kmask, saved=saved)
    dfull = torch.zeros(B, L, d, device="cuda", dtype=torch.bfloat16)
    dfull[:, r0:r1] = dsub.view(B, r1 - r0, d)
    wg_full = mk()
    dx_full = m.trunk.backward(dfull.view(B * L, d), B, L, kmask, saved, wgrad=wg_full)
    # row-range computation
    saved = []
    h_sub = m.trunk.forward(x.clone(), B, L, kmask, saved=saved, out_rows=(r0, r1))
    assert h_sub.shape == (B * (r1 - r0), d) and saved[-1].rows == (r0, r1)
    wg_sub = mk()
    dx_sub = m.trunk.backward(dsub.clone(), B, L, kmask, saved, wgrad=wg_sub)
    rel = lambda a, b: float((a.double() - b.double()).norm() / b.double().norm().clamp_min(1e-30))
    assert rel(h_sub.view(B, r1 - r0, d), h_full.view(B, L, d)[:, r0:r1]) < 5e-3
    assert rel(dx_sub, dx_full) < 1e-2
    for k in wg_full:
        for a, b in zip(wg_sub[k], wg_full[k]):
            assert rel(a, b) < 1.5e-2, k
    # inference form
    h_inf = m.trunk.forward(x.clone(), B, L, kmask, out_rows=(r0, r1))
    assert rel(h_inf, h_sub) < 5e-3
