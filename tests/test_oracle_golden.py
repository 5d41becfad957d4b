"""CPU: pin the restatement in oracle/ against fixtures generated from the REAL reference
(oracle/make_golden.py; the reference ships no tests for this path)."""
import os

import numpy as np
import pytest
import torch

from oracle import traj_tokenizer as tt
from oracle.traj_vlm import LlmCfg, PointCfg, TrajVLMOracle, train_loss, greedy_decode
from egoscaler_b200 import synth
from egoscaler_b200.config import TrajVLMConfig
from conftest import GOLDEN, apply_trained_delta


@pytest.fixture(scope="module")
def gtok():
    return np.load(os.path.join(GOLDEN, "traj_tokenizer.npz"))


def test_discretize_matches_reference(gtok):
    assert np.array_equal(tt.discretize_action(gtok["x32"], 256), gtok["q32"])
    assert np.array_equal(tt.discretize_action(gtok["x64"], 256), gtok["q64"])
    assert np.array_equal(tt.discretize_action(gtok["x32"], 64), gtok["q32_64bins"])
    assert np.array_equal(tt.discretize_action(gtok["demo_norm"], 256), gtok["demo_q"])


def test_token_to_action_matches_reference(gtok):
    assert np.array_equal(tt.token_to_action(np.arange(256), 256), gtok["deq256"])
    assert np.array_equal(tt.token_to_action(np.arange(64), 64), gtok["deq64"])


def test_str_to_float_matches_reference(gtok):
    maxmin = gtok["s2f_maxmin"].tolist()
    for i, s in enumerate(gtok["s2f_strings"].tolist()):
        r = tt.str_to_float_rt2(s, maxmin)
        ref = gtok[f"s2f_{i}"]
        if ref.shape[0] == 0:
            assert r is None
        else:
            assert r.dtype == np.float32 and np.array_equal(r, ref)


def test_scalers_and_denorm_match_reference(gtok):
    maxmin = gtok["s2f_maxmin"].tolist()
    assert np.array_equal(tt.rt2_scaler(gtok["scaler_in"].copy(), maxmin), gtok["rt2_out"])
    assert np.array_equal(tt.simple_scaler(gtok["simple_in"].copy(), maxmin), gtok["simple_out"])
    assert np.array_equal(tt.denorm(gtok["denorm_in"], True), gtok["denorm_norm_out"])
    out = tt.denorm(gtok["denorm_in"], False, gtok["denorm_max_abs"], gtok["denorm_mean"], gtok["denorm_std"])
    assert np.array_equal(out, gtok["denorm_std_out"])


def test_ids_to_poses_equals_string_path(gtok):
    rng = np.random.default_rng(0)
    b = rng.integers(0, 256, (20, 6))
    ids = [tt.TS_TOKEN]
    for row in b:
        ids += (row + tt.TRAJ_TOKEN_BASE).tolist() + [tt.TSEP_TOKEN]
    ids += [tt.TE_TOKEN, 2, 0, 0]
    s = "<ts>" + "".join(" ".join(f"<p{v}>" for v in row) + "<tsep>" for row in b) + "<te>"
    a = tt.ids_to_poses(ids)
    r = tt.str_to_float_rt2(s, None, scale=False)
    assert np.array_equal(a, r)


# ------------------------------------------------------------------ model
@pytest.fixture(scope="module")
def gm():
    return np.load(os.path.join(GOLDEN, "traj_vlm_tiny.npz"))


@pytest.fixture(scope="module")
def oracle_model():
    torch.manual_seed(0)
    m = TrajVLMOracle(LlmCfg.tiny(), PointCfg())
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    m.configure_trainable()
    m.eval()
    return m


def test_oracle_state_dict_keys_match_reference(oracle_model, gm):
    keys = set(oracle_model.state_dict().keys())
    gk = {k.split("/", 1)[1] for k in gm.files if k.startswith("gradnorm/")}
    assert gk <= keys
    assert "model.layers.0.self_attn.q_proj.weight" in keys and "lm_head.weight" in keys
    assert "model.point_backbone.encoder.first_conv.1.running_mean" in keys


def test_oracle_forward_loss_grads_match_reference(oracle_model, gm):
    m = oracle_model
    bt = synth.make_batch(TrajVLMConfig.tiny(), 2, seed=1)
    assert np.array_equal(bt["tokens"].numpy(), gm["tokens"])
    P = int(gm["prompt_len"])
    start = torch.from_numpy(gm["fps_start"])
    logits = m(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=start, training=True)
    loss = train_loss(logits, bt["tokens"], P, 0)
    loss.backward()
    L = logits.detach()
    sub = L[:, gm["logit_rows"]][:, :, gm["logit_cols"]].numpy()
    np.testing.assert_allclose(sub, gm["logits_sub"], atol=2e-4, rtol=1e-4)
    np.testing.assert_allclose(torch.logsumexp(L, -1).numpy(), gm["logits_lse"], atol=2e-4, rtol=1e-5)
    assert (L.argmax(-1).numpy() == gm["logits_argmax"]).mean() > 0.999
    assert abs(loss.item() - float(gm["loss"])) < 1e-4
    n = 0
    for k, p in m.named_parameters():
        if f"gradnorm/{k}" not in gm.files:
            assert p.grad is None or not p.requires_grad or p.grad.abs().max() == 0, k
            continue
        g = p.grad.flatten()
        ref = gm[f"gradval/{k}"]
        got = g[torch.from_numpy(gm[f"gradidx/{k}"])].numpy()
        np.testing.assert_allclose(got, ref, atol=1e-6 + 1e-3 * np.abs(ref).max(), rtol=1e-3, err_msg=k)
        assert abs(g.norm().item() - float(gm[f"gradnorm/{k}"])) <= 1e-3 * float(gm[f"gradnorm/{k}"]) + 1e-9, k
        n += 1
    assert n >= 9      # embed, lm_head, norm, 6 projector tensors


def test_oracle_point_encoder_matches_reference(oracle_model, gm):
    bt = synth.make_batch(TrajVLMConfig.tiny(), 2, seed=1)
    with torch.no_grad():
        pf = oracle_model.model.point_backbone(bt["pcrgbs"], torch.from_numpy(gm["fps_start"]))
        pp = oracle_model.model.point_proj(pf)
    np.testing.assert_allclose(pf[:, ::16, ::8].numpy(), gm["point_feats_sub"], atol=1e-4, rtol=1e-4)
    np.testing.assert_allclose(pp[:, ::16, ::8].numpy(), gm["point_proj_sub"], atol=1e-4, rtol=1e-4)


def test_oracle_greedy_decode_matches_reference(oracle_model, gm):
    bt = synth.make_batch(TrajVLMConfig.tiny(), 2, seed=1)
    T = gm["greedy_ids"].shape[1]
    ids, scores = greedy_decode(oracle_model, bt["prompts"], bt["prompt_masks"], bt["pcrgbs"], T,
                                fps_start=torch.from_numpy(gm["fps_start"]), return_scores=True)
    P = bt["prompts"].shape[1]
    gen = ids[:, P:].numpy()
    gap = gm["greedy_step_top2gap"]
    agree = gen[:, :gap.shape[1]] == gm["greedy_ids"][:, :gap.shape[1]]
    # ids must agree wherever the reference's own top-2 margin is above fp32 noise
    assert agree[gap > 1e-3].all()
    sl = torch.stack(scores, 1)[:, :, gm["logit_cols"]].numpy()
    n = min(sl.shape[1], gm["greedy_step_logits_sub"].shape[1])
    np.testing.assert_allclose(sl[:, :n], gm["greedy_step_logits_sub"][:, :n], atol=5e-4, rtol=1e-4)


def test_teacher_forced_poses(gm):
    P = int(gm["prompt_len"])
    for b in range(2):
        p = tt.ids_to_poses(gm["tokens"][b, P:].tolist())
        # the prompt holds <ts> + first pose + <tsep>; 19 poses follow, and the trailing "<te>" segment
        # has no match so it repeats the last pose (utils.py:91-93) -> 20 rows
        assert p.shape == (20, 6) and np.array_equal(p[-1], p[-2]) and np.array_equal(p, gm["tf_poses"][b])


def test_sampling_filter_matches_hf_processors():
    """oracle/sampling.py vs the real HF processor classes the reference's generate() drives (model_arch.py:94-108)."""
    import torch
    from transformers.generation.logits_process import (RepetitionPenaltyLogitsProcessor, TemperatureLogitsWarper,
                                                        TopKLogitsWarper, TopPLogitsWarper)
    from oracle.sampling import hf_filter
    g = torch.Generator().manual_seed(11)
    for V, k, p, temp, pen in [(32262, 50, 0.95, 1.0, 1.0), (32262, 50, 0.95, 0.7, 1.3), (1000, 5, 0.5, 1.0, 1.0),
                               (4096, 200, 0.99, 1.5, 1.1)]:
        scores = (torch.randn(4, V, generator=g) * 3).bfloat16().float()        # bf16-valued like the reference's logits
        ids = torch.randint(0, V, (4, 37), generator=g)
        ref = scores.clone()
        if pen != 1.0:
            ref = RepetitionPenaltyLogitsProcessor(pen)(ids, ref)
        if temp != 1.0:
            ref = TemperatureLogitsWarper(temp)(ids, ref)
        ref = TopKLogitsWarper(k)(ids, ref)
        ref = TopPLogitsWarper(p)(ids, ref)
        mine = hf_filter(scores, ids, temp, k, p, pen)
        # value-ties across the top-p boundary are ordered arbitrarily by torch.sort(stable=False): compare per value
        assert torch.equal(torch.sort(ref, dim=-1)[0], torch.sort(mine, dim=-1)[0])
        same = (ref == mine) | (torch.isinf(ref) & torch.isinf(mine))
        for b in range(4):                                   # any disagreement is a tie with the smallest kept value
            kept_min = ref[b][torch.isfinite(ref[b])].min()
            pre = hf_filter(scores, ids, temp, k, 1.0, pen)[b]          # before top-p
            assert bool((pre[~same[b]] == kept_min).all())


# ------------------------------------------------------------------ other recipes (fixtures: make_golden.py variants)
def _variant_check(name, point_cfg, vlm_cfg, unfreeze_lm, min_grads, llm=None):
    gv = np.load(os.path.join(GOLDEN, name + ".npz"))
    torch.manual_seed(0)
    m = TrajVLMOracle(llm or LlmCfg.tiny(), point_cfg)
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    m.configure_trainable(unfreeze_language_model=unfreeze_lm)
    m.eval()
    bt = synth.make_batch(vlm_cfg, 2, seed=1)
    P = int(gv["prompt_len"])
    start = torch.from_numpy(gv["fps_start"])
    logits = m(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=start, training=True)
    loss = train_loss(logits, bt["tokens"], P, 0)
    loss.backward()
    sub = logits.detach()[:, gv["logit_rows"]][:, :, gv["logit_cols"]].numpy()
    np.testing.assert_allclose(sub, gv["logits_sub"], atol=3e-4, rtol=1e-4)
    assert abs(loss.item() - float(gv["loss"])) < 1e-4
    n = 0
    for k, p in m.named_parameters():
        if f"gradnorm/{k}" not in gv.files:
            continue
        g = p.grad.flatten()
        ref = gv[f"gradval/{k}"]
        got = g[torch.from_numpy(gv[f"gradidx/{k}"])].numpy()
        np.testing.assert_allclose(got, ref, atol=1e-6 + 2e-3 * np.abs(ref).max(), rtol=2e-3, err_msg=k)
        n += 1
    assert n >= min_grads
    with torch.no_grad():
        pf = m.model.point_backbone(bt["pcrgbs"], start)
    np.testing.assert_allclose(pf[:, ::16, ::8].numpy(), gv["point_feats_sub"], atol=2e-4, rtol=1e-4)
    T = gv["greedy_ids"].shape[1]
    ids = greedy_decode(m, bt["prompts"], bt["prompt_masks"], bt["pcrgbs"], T, fps_start=start)
    gen = ids[:, P:].numpy()
    gap = gv["greedy_step_top2gap"]
    k = min(gen.shape[1], gap.shape[1])
    agree = gen[:, :k] == gv["greedy_ids"][:, :k]
    first_eos = [(list(r).index(2) if 2 in r else k) for r in gv["greedy_ids"][:, :k]]
    for b in range(2):                                        # the fixture loop does not special-case EOS rows
        upto = min(k, first_eos[b] + 1)
        assert agree[b, :upto][gap[b, :upto] > 1e-3].all()


def test_oracle_unfrozen_language_model_matches_reference():
    """--unfreeze_language_model (model_arch.py:40-45): gradients of every LLM-layer tensor."""
    _variant_check("traj_vlm_tiny_unfrozen_lm", PointCfg(), TrajVLMConfig.tiny(), True, 9 + 2 * 9)


def test_oracle_hd128_unfrozen_language_model_matches_reference():
    """The full-size head width (2 heads x 128) at parity-test size: every LLM-layer gradient vs the real reference."""
    _variant_check("traj_vlm_hd128_unfrozen_lm", PointCfg(), TrajVLMConfig.tiny_hd128(), True, 9 + 2 * 9, llm=LlmCfg.tiny_hd128())


def test_oracle_hd128_trained_greedy_trajectory_matches_reference():
    """traj_vlm_hd128_trained.npz: the over-fitted reference model decodes a REAL trajectory; the restatement must emit
    the same 152 ids (the reference's smallest top-2 margin is > 8), the same logits, and the same poses as the
    reference's own str_to_float on the decoded string."""
    g = np.load(os.path.join(GOLDEN, "traj_vlm_hd128_trained.npz"))
    m = TrajVLMOracle(LlmCfg.tiny_hd128(), PointCfg())
    m.load_state_dict(apply_trained_delta(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), g), strict=True)
    m.configure_trainable()
    m.eval()
    bt = synth.make_batch(TrajVLMConfig.tiny_hd128(), 2, seed=1)
    assert np.array_equal(bt["tokens"].numpy(), g["tokens"])
    P = int(g["prompt_len"])
    start = torch.from_numpy(g["fps_start"])
    logits = m(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=start, training=True)
    loss = train_loss(logits, bt["tokens"], P, 0)
    sub = logits.detach()[:, g["logit_rows"]][:, :, g["logit_cols"]].numpy()
    np.testing.assert_allclose(sub, g["logits_sub"], atol=2e-3, rtol=1e-4)
    assert abs(loss.item() - float(g["loss"])) < 1e-5
    ref = g["greedy_ids"]
    assert len(np.unique(ref)) > 100 and float(g["greedy_step_top2gap"].min()) > 1.0        # a real, decisive trajectory
    ids, scores = greedy_decode(m, bt["prompts"], bt["prompt_masks"], bt["pcrgbs"], 152, fps_start=start, return_scores=True)
    gen = ids[:, P:].numpy()
    assert np.array_equal(gen[:, :ref.shape[1]], ref)
    assert np.array_equal(ref, g["tokens"][:, P:P + ref.shape[1]])            # ... and it IS the training trajectory
    sl = torch.stack(scores, 1)[:, :, g["logit_cols"]].numpy()
    np.testing.assert_allclose(sl, g["greedy_step_logits_sub"][:, :sl.shape[1]], atol=2e-3, rtol=1e-4)
    for b in range(2):
        raw = tt.ids_to_poses(gen[b].tolist())
        assert np.array_equal(raw, g[f"greedy_poses_raw/{b}"]) and raw.shape == (20, 6)
        assert np.array_equal(tt.rt2_scaler(raw.copy(), g["maxmin"].tolist()), g[f"greedy_poses_rt2/{b}"])


def test_oracle_base_encoder_matches_reference():
    """PointTransformer_base_8192point (trans 1152, 12 x 96, k=48, enc 512, Linear projector)."""
    _variant_check("traj_vlm_tiny_base_encoder", PointCfg.base(), TrajVLMConfig.tiny_base(), False, 5)


@pytest.mark.parametrize("name,base", [("traj_vlm_tiny_unfrozen_pc", False), ("traj_vlm_tiny_base_unfrozen_pc", True)])
def test_oracle_unfrozen_point_encoder_matches_reference(name, base):
    """--unfreeze_pc_encoder (model_arch.py:36-38,121-122): encoder in TRAIN mode = BatchNorm batch statistics
    (dvae.py:196,202) + DropPath (point_encoder.py:62,74-75,134); the fixture stores the reference's DropPath draws."""
    gv = np.load(os.path.join(GOLDEN, name + ".npz"))
    torch.manual_seed(0)
    m = TrajVLMOracle(LlmCfg.tiny(), PointCfg.base() if base else PointCfg())
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    m.configure_trainable(unfreeze_pc_encoder=True)
    m.eval()
    m.model.point_backbone.train()
    m.model.point_backbone.drop_u = torch.from_numpy(gv["drop_u"])
    bt = synth.make_batch(TrajVLMConfig.tiny_base() if base else TrajVLMConfig.tiny(), 2, seed=1)
    P = int(gv["prompt_len"])
    logits = m(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=torch.from_numpy(gv["fps_start"]), training=True)
    loss = train_loss(logits, bt["tokens"], P, 0)
    loss.backward()
    sub = logits.detach()[:, gv["logit_rows"]][:, :, gv["logit_cols"]].numpy()
    np.testing.assert_allclose(sub, gv["logits_sub"], atol=3e-4, rtol=1e-4)
    assert abs(loss.item() - float(gv["loss"])) < 1e-4
    n = 0
    for k, p in m.named_parameters():
        if f"gradnorm/{k}" not in gv.files:
            continue
        ref = gv[f"gradval/{k}"]
        got = p.grad.flatten()[torch.from_numpy(gv[f"gradidx/{k}"])].numpy()
        np.testing.assert_allclose(got, ref, atol=1e-6 + 3e-3 * np.abs(ref).max(), rtol=3e-3, err_msg=k)
        n += 1
    assert n > 150                                            # all encoder tensors + the default trainable set
    enc = m.model.point_backbone.encoder
    for nm, bn in (("first_conv.1", enc.first_conv[1]), ("second_conv.1", enc.second_conv[1])):
        np.testing.assert_allclose(bn.running_mean.numpy(), gv[f"running_mean/{nm}"], atol=1e-5, rtol=1e-4)
        np.testing.assert_allclose(bn.running_var.numpy(), gv[f"running_var/{nm}"], atol=1e-5, rtol=1e-4)


def test_metrics_oracle_matches_reference():
    """oracle/metrics.py vs outputs of the real egoscaler/models/utils/metrics.py (tests/golden/traj_metrics.npz)."""
    from oracle import metrics as om
    g = np.load(os.path.join(GOLDEN, "traj_metrics.npz"))
    for b in range(g["gen"].shape[0]):
        gen, gt = g["gen"][b, :g["gen_len"][b]], g["gt"][b, :g["gt_len"][b]]
        got = [om.ade(gen, gt), om.fde(gen, gt), om.ide(gen, gt), om.angular_distance(gen[:, 3:], gt[:, 3:])]
        np.testing.assert_allclose(got, g["ade_fde_ide_gd_all6"][b], rtol=1e-6, atol=1e-7)
        got3 = [om.ade(gen[:, :3], gt[:, :3]), om.fde(gen[:, :3], gt[:, :3]), om.ide(gen[:, :3], gt[:, :3])]
        np.testing.assert_allclose(got3, g["ade_fde_ide_xyz"][b], rtol=1e-6, atol=1e-7)


RGBD_CASES = ["plain", "boxes_thres", "aria_crop", "all_boxed"]


def test_rgbd_frontend_oracle_matches_reference():
    """oracle/rgbd_frontend.py vs outputs of the real get_points_colors (pcm_tools.py:68-96) and pc_norm
    (pointllm/data/utils.py:146-157) stored in tests/golden/rgbd_frontend.npz: bit-exact."""
    from oracle import rgbd_frontend as rf
    g = np.load(os.path.join(GOLDEN, "rgbd_frontend.npz"))
    for name in RGBD_CASES:
        c, fx, fy, thr = g[f"{name}/cam"]
        pts, cols = rf.get_points_colors(g[f"{name}/rgbd"], g[f"{name}/boxes"].tolist(), c, fx, fy, None if np.isnan(thr) else thr)
        assert pts.dtype == np.float64 and cols.dtype == np.float32
        assert np.array_equal(pts, g[f"{name}/points"]) and np.array_equal(cols, g[f"{name}/colors"]), name
        if f"{name}/pc_norm" in g.files:
            idx = g[f"{name}/sample_idx"]
            pc = np.concatenate([pts[idx], cols[idx].astype(np.float64)], axis=1)
            assert np.array_equal(rf.pc_norm(pc), g[f"{name}/pc_norm"]), name
    assert g["all_boxed/points"].shape == (0, 3)


def test_traj_preprocess_oracle_matches_reference():
    """oracle smoothing_traj / preprocess_traj vs outputs of the real traj_utils.py (tests/golden/traj_preprocess.npz)."""
    from oracle import traj_tokenizer as ot
    g = np.load(os.path.join(GOLDEN, "traj_preprocess.npz"))
    for ns in (20, 7):
        for smooth in (0, 1):
            for b, n in enumerate(g["lens"]):
                t = g["traj"][b, :n]
                if smooth:
                    t = ot.smoothing_traj(t)
                r, m = ot.preprocess_traj(t, ns)
                assert np.array_equal(r, g[f"out/{ns}/{smooth}"][b]), (ns, smooth, n)
                assert np.array_equal(m, g[f"mask/{ns}/{smooth}"][b])
