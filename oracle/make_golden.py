"""Generate tests/golden/*.npz from the REAL reference (dev container only; TEST INFRASTRUCTURE).

    python -m oracle.make_golden

Imports the unmodified reference from /root/reference through ``oracle.ref_shim`` and records
small fixtures that pin the restatement in ``oracle/`` (and, on the GPU box where the reference
tree does not exist, the CUDA path):

* ``traj_tokenizer.npz`` -- utils/utils.py:13-104 (discretize_action / token_to_action /
  str_to_float(rt2=True) / rt2_scaler / simple_scaler) and dataset.py:126-148 (denorm).
* ``traj_vlm_tiny.npz``  -- PointLLMLlamaForCausalLM (pointllm.py:181-253) on BASELINE config A
  (tiny, B=2, L=696): logits (sub-sampled), loss (train.py:174-181), gradients of the default
  trainable set (model_arch.py:33-51), manual KV-cache greedy ids and poses.

Weights come from ``egoscaler_b200.synth.det_state_dict`` (pure function of the state_dict keys),
so the fixture stays small.
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_shim, traj_tokenizer as tt          # noqa: E402
from oracle.traj_vlm import LlmCfg                          # noqa: E402
from egoscaler_b200 import synth                            # noqa: E402
from egoscaler_b200.config import TrajVLMConfig             # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
FPS_SEED = 1234          # torch.manual_seed right before each forward (misc.py:52 draws the FPS start)


def golden_tokenizer():
    _, _, ru = ref_shim.load()
    rng = np.random.default_rng(7)
    edges = np.linspace(-1, 1, 256)
    x32 = np.concatenate([
        rng.uniform(-1.05, 1.05, 4096).astype(np.float32),
        edges.astype(np.float32),                                # float32-rounded edges
        np.nextafter(edges.astype(np.float32), np.float32(2)),
        np.nextafter(edges.astype(np.float32), np.float32(-2)),
        np.array([np.nan, np.inf, -np.inf, -1.0, 1.0, 0.0, -0.0, 1e-45, -1e-45, 3.4e38, -3.4e38], np.float32),
    ])
    x64 = np.concatenate([rng.uniform(-1.05, 1.05, 1024), edges, np.nextafter(edges, 2), np.nextafter(edges, -2)])
    out = dict(x32=x32, q32=np.array(ru.discretize_action(x32, 256), np.int64),
               x64=x64, q64=np.array(ru.discretize_action(x64, 256), np.int64),
               q32_64bins=np.array(ru.discretize_action(x32, 64), np.int64))
    ids = np.arange(256)
    out["deq256"] = np.array(ru.token_to_action(ids.tolist(), 256), np.float64)
    out["deq64"] = np.array(ru.token_to_action(np.arange(64).tolist(), 64), np.float64)
    # str_to_float(rt2=True): well-formed, with a missing segment, with garbage, empty
    strs = []
    b = rng.integers(0, 256, (20, 6))
    strs.append("<ts>" + "".join(" ".join(f"<p{v}>" for v in row) + "<tsep>" for row in b) + "<te>")
    s1 = [" ".join(f"<p{v}>" for v in row) for row in b[:6]]
    s1[2] = "<p3> <p4>"                       # malformed -> repeats previous pose
    s1[4] = ""
    strs.append("<tsep>".join(s1))
    strs.append("hello <p1> <p2> <p3> <p4> <p5> <p6> world<tsep>junk<tsep><p255> <p0> <p128> <p127> <p1> <p254>")
    strs.append("nothing here")
    maxmin = [2.5, 0.1]
    for i, s in enumerate(strs):
        r = ru.str_to_float(s, maxmin, "val", rt2=True, num_bins=256)
        out[f"s2f_{i}"] = np.zeros((0, 6), np.float32) if r is None else r
    out["s2f_strings"] = np.array(strs)
    out["s2f_maxmin"] = np.array(maxmin)
    t = rng.uniform(-1, 1, (33, 6)).astype(np.float32)
    out["scaler_in"] = t
    out["rt2_out"] = ru.rt2_scaler(t.copy(), maxmin, "val")
    t2 = rng.uniform(0, 100, (33, 6)).astype(np.float32)
    out["simple_in"] = t2
    out["simple_out"] = ru.simple_scaler(t2.copy(), maxmin)
    # denorm (dataset.py:126-148): the method only reads self.do_norm/do_standard/mean/std
    import importlib.util
    cwd = os.getcwd()
    os.chdir(os.path.join(ref_shim.REF_ROOT, "egoscaler", "models", "pointllm"))
    sys.path.insert(0, os.getcwd())
    try:
        sys.modules.setdefault("utils.utils", ru)
        src = open("dataset.py").read()
    finally:
        os.chdir(cwd)
    # exec only the denorm method body against the reference's DatasetConfig
    from egoscaler.configs.dataset import DatasetConfig as dataset_cfg
    start = src.index("    def denorm(")
    end = src.index("    def collate_fn(")
    ns = {"np": np, "torch": torch, "dataset_cfg": dataset_cfg}
    exec("class _D:\n" + src[start:end], ns)
    D = ns["_D"]
    tr = torch.from_numpy(rng.uniform(-1, 1, (3, 20, 6)).astype(np.float32))
    d = D(); d.do_norm, d.do_standard = True, False
    out["denorm_in"] = tr.numpy().copy()
    out["denorm_norm_out"] = d.denorm(tr.clone(), None)
    d = D(); d.do_norm, d.do_standard = False, True
    d.mean = rng.normal(size=6); d.std = rng.uniform(0.5, 2, 6)
    ma = rng.uniform(0.5, 3, (3, 6))
    out["denorm_mean"], out["denorm_std"], out["denorm_max_abs"] = d.mean, d.std, ma
    out["denorm_std_out"] = d.denorm(tr.clone(), ma)
    # real trajectory from the reference's demo asset, normalised like denorm's inverse
    import pickle
    with open(os.path.join(ref_shim.REF_ROOT, "assets", "demo", "trajectory.pkl"), "rb") as f:
        demo = pickle.load(f)
    rv = np.asarray(demo["traj_rotvec"], np.float64)
    out["demo_rotvec"] = rv
    n = rv.copy()
    n[:, 0] = (n[:, 0] - dataset_cfg.min_x) / (dataset_cfg.max_x - dataset_cfg.min_x) * 2 - 1
    n[:, 1] = (n[:, 1] - dataset_cfg.min_y) / (dataset_cfg.max_y - dataset_cfg.min_y) * 2 - 1
    n[:, 2] = (n[:, 2] - dataset_cfg.min_z) / (dataset_cfg.max_z - dataset_cfg.min_z) * 2 - 1
    n[:, 3:] /= np.pi
    out["demo_norm"] = n
    out["demo_q"] = np.array(ru.discretize_action(n, 256), np.int64)
    np.savez_compressed(os.path.join(GOLD, "traj_tokenizer.npz"), **out)
    print("traj_tokenizer.npz:", {k: v.shape for k, v in out.items()})


def _fps_start(n_points: int, batch: int) -> torch.Tensor:
    """The draw misc.py:52 makes right after ``torch.manual_seed(FPS_SEED)``."""
    torch.manual_seed(FPS_SEED)
    return torch.randint(0, n_points, (batch,), dtype=torch.long)


def _ref_forward(m, ids, mask, pc, **kw):
    torch.manual_seed(FPS_SEED)
    return m(input_ids=ids, attention_mask=mask, point_clouds=pc, return_dict=True, **kw)


def golden_model(name="traj_vlm_tiny", batch=2, new_tokens=152):
    llm = LlmCfg.tiny()
    m = ref_shim.build_reference_model(llm)
    sd = synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True)
    m.load_state_dict(sd, strict=True)
    # model_arch.py:33-51 + :110-124 : default freezing, frozen parts in eval
    for p in m.model.point_backbone.parameters():
        p.requires_grad = False
    for p in m.model.layers.parameters():
        p.requires_grad = False
    m.train()
    m.model.layers.eval()
    m.model.point_backbone.eval()

    cfg = TrajVLMConfig.tiny()
    bt = synth.make_batch(cfg, batch, seed=1)
    ids, mask, pc = bt["tokens"], bt["attention_masks"], bt["pcrgbs"]
    P = bt["prompts"].shape[1]
    out = dict(fps_seed=np.int64(FPS_SEED), fps_start=_fps_start(pc.shape[1], batch).numpy(),
               tokens=ids.numpy(), prompt_len=np.int64(P),
               torch_version=np.array(torch.__version__), transformers_version=np.array(__import__("transformers").__version__))

    # ---- forward + loss + grads (train.py:166-183)
    res = _ref_forward(m, ids, mask, pc)
    logits = res.logits
    lg = logits[:, P - 1:-1, :]
    tg = ids[:, P:]
    loss = torch.nn.functional.cross_entropy(lg.reshape(-1, lg.shape[-1]), tg.flatten(), ignore_index=llm.pad_token_id)
    loss.backward()
    L = logits.detach()
    rng = np.random.default_rng(3)
    cols = np.sort(np.concatenate([rng.choice(32000, 256, replace=False), np.arange(32000, 32262)]))
    rows = np.sort(np.concatenate([np.arange(0, 8), np.arange(510, 530), np.arange(P - 4, L.shape[1])]))
    out.update(logit_rows=rows, logit_cols=cols,
               logits_sub=L[:, rows][:, :, cols].numpy(),
               logits_argmax=L.argmax(-1).numpy(), logits_max=L.max(-1).values.numpy(),
               logits_lse=torch.logsumexp(L, -1).numpy(),
               loss=np.float64(loss.item()))
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        g = p.grad.detach()
        out[f"gradnorm/{k}"] = np.float64(g.norm().item())
        flat = g.flatten()
        sel = torch.from_numpy(rng.choice(flat.numel(), min(4096, flat.numel()), replace=False)).long()
        out[f"gradidx/{k}"] = sel.numpy()
        out[f"gradval/{k}"] = flat[sel].numpy()
    # embedding-grad rows that are touched (token rows) -- useful to check the scatter
    eg = m.model.embed_tokens.weight.grad
    out["embed_grad_rownorm_nonzero"] = torch.nonzero(eg.norm(dim=1) > 0).flatten().numpy()

    # ---- intermediate: point tokens after projector (for encoder parity)
    with torch.no_grad():
        torch.manual_seed(FPS_SEED)
        pf = m.model.point_backbone(pc)
        out["point_feats_sub"] = pf[:, ::16, ::8].numpy()
        out["point_proj_sub"] = m.model.point_proj(pf)[:, ::16, ::8].numpy()

    # ---- manual KV-cache greedy decode (generate() is unusable under transformers 5.5)
    m.eval()
    with torch.no_grad():
        seq, msk = bt["prompts"].clone(), bt["prompt_masks"].clone()
        r = _ref_forward(m, seq, msk, pc, use_cache=True)
        past = r.past_key_values
        step_logits = [r.logits[:, -1].float()]
        unfinished = torch.ones(batch, dtype=torch.long)
        gen = []
        for t in range(new_tokens):
            lgt = step_logits[-1]
            nxt = lgt.argmax(-1)
            nxt = nxt * unfinished + llm.pad_token_id * (1 - unfinished)
            gen.append(nxt)
            unfinished = unfinished & (nxt != llm.eos_token_id).long()
            msk = torch.cat((msk, torch.ones(batch, 1, dtype=torch.bool)), 1)
            if t == new_tokens - 1 or unfinished.max() == 0:
                break
            r = m(input_ids=nxt[:, None], attention_mask=msk, past_key_values=past, use_cache=True, return_dict=True)
            past = r.past_key_values
            step_logits.append(r.logits[:, -1].float())
        gen = torch.stack(gen, 1)
        out["greedy_ids"] = gen.numpy()
        sl = torch.stack(step_logits, 1)
        out["greedy_step_logits_sub"] = sl[:, :, cols].numpy()
        out["greedy_step_top2gap"] = (sl.topk(2, -1).values[..., 0] - sl.topk(2, -1).values[..., 1]).numpy()

        # teacher-forced decode of the ground-truth trajectory tokens: ids -> poses path
        tf_ids = ids[:, P:]
        poses = []
        for b in range(batch):
            p_ = tt.ids_to_poses(tf_ids[b].tolist())
            poses.append(p_)
        out["tf_poses"] = np.stack(poses)
    np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), **out)
    print(f"{name}.npz: loss={out['loss']:.6f} greedy[0,:12]={out['greedy_ids'][0, :12]}",
          f"size={os.path.getsize(os.path.join(GOLD, name + '.npz')) / 1e6:.2f} MB")


def golden_variant(name, point_cfg_name="PointTransformer_8192point_2layer", unfreeze_lm=False, batch=2, new_tokens=24,
                   point="v12", unfreeze_pc=False, hd128=False):
    """Smaller fixtures for the other recipes of SURVEY 8d: (a) ``--unfreeze_language_model`` (model_arch.py:40-45):
    gradients of every LLM-layer tensor; (b) config E: the base PointBERT encoder yaml (trans 1152, 12 heads x 96,
    k=48, enc 512, single-Linear projector) with the default freezing: logits / loss / trainable grads / greedy ids."""
    llm = LlmCfg.tiny_hd128() if hd128 else LlmCfg.tiny()
    m = ref_shim.build_reference_model(llm, point_cfg_name)
    sd = synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True)
    m.load_state_dict(sd, strict=True)
    for p in m.model.point_backbone.parameters():
        p.requires_grad = bool(unfreeze_pc)
    for p in m.model.layers.parameters():
        p.requires_grad = bool(unfreeze_lm)
    m.train()
    if not unfreeze_lm:
        m.model.layers.eval()
    if not unfreeze_pc:
        m.model.point_backbone.eval()                      # model_arch.py:121-122
    cfg = TrajVLMConfig.tiny() if point == "v12" else TrajVLMConfig.tiny_base()
    if hd128:
        cfg = TrajVLMConfig.tiny_hd128()
    bt = synth.make_batch(cfg, batch, seed=1)
    ids, mask, pc = bt["tokens"], bt["attention_masks"], bt["pcrgbs"]
    P = bt["prompts"].shape[1]
    out = dict(fps_seed=np.int64(FPS_SEED), fps_start=_fps_start(pc.shape[1], batch).numpy(), prompt_len=np.int64(P),
               torch_version=np.array(torch.__version__), transformers_version=np.array(__import__("transformers").__version__))
    if unfreeze_pc:
        # the CPU-generator draws the train-mode encoder makes after torch.manual_seed(FPS_SEED): the FPS start
        # (misc.py:52) and then one torch.rand((B,1,1)) per DropPath with rate > 0 (blocks 1..11, attention then MLP)
        torch.manual_seed(FPS_SEED)
        torch.randint(0, pc.shape[1], (batch,), dtype=torch.long)
        depth = len(m.model.point_backbone.blocks.blocks)
        du = torch.zeros(depth, 2, batch)
        for j in range(1, depth):
            for br in range(2):
                du[j, br] = torch.rand((batch, 1, 1)).flatten()
        out["drop_u"] = du.numpy()
    res = _ref_forward(m, ids, mask, pc)
    logits = res.logits
    lg = logits[:, P - 1:-1, :]
    loss = torch.nn.functional.cross_entropy(lg.reshape(-1, lg.shape[-1]), ids[:, P:].flatten(), ignore_index=llm.pad_token_id)
    loss.backward()
    L = logits.detach()
    rng = np.random.default_rng(3)
    cols = np.sort(np.concatenate([rng.choice(32000, 256, replace=False), np.arange(32000, 32262)]))
    rows = np.sort(np.concatenate([np.arange(0, 8), np.arange(510, 530), np.arange(P - 4, L.shape[1])]))
    out.update(logit_rows=rows, logit_cols=cols, logits_sub=L[:, rows][:, :, cols].numpy(), loss=np.float64(loss.item()))
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        g = p.grad.detach()
        out[f"gradnorm/{k}"] = np.float64(g.norm().item())
        flat = g.flatten()
        sel = torch.from_numpy(rng.choice(flat.numel(), min(2048, flat.numel()), replace=False)).long()
        out[f"gradidx/{k}"] = sel.numpy()
        out[f"gradval/{k}"] = flat[sel].numpy()
    if unfreeze_pc:
        enc = m.model.point_backbone.encoder
        for nm, bn in (("first_conv.1", enc.first_conv[1]), ("second_conv.1", enc.second_conv[1])):
            out[f"running_mean/{nm}"] = bn.running_mean.numpy().copy()       # after ONE train-mode forward (momentum 0.1)
            out[f"running_var/{nm}"] = bn.running_var.numpy().copy()
        np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), **out)
        print(f"{name}.npz: loss={out['loss']:.6f} size={os.path.getsize(os.path.join(GOLD, name + '.npz')) / 1e6:.2f} MB")
        return
    with torch.no_grad():
        torch.manual_seed(FPS_SEED)
        pf = m.model.point_backbone(pc)
        out["point_feats_sub"] = pf[:, ::16, ::8].numpy()
        out["point_proj_sub"] = m.model.point_proj(pf)[:, ::16, ::8].numpy()
    m.eval()
    with torch.no_grad():
        msk = bt["prompt_masks"].clone()
        r = _ref_forward(m, bt["prompts"].clone(), msk, pc, use_cache=True)
        past = r.past_key_values
        step_logits = [r.logits[:, -1].float()]
        gen = []
        for t in range(new_tokens):
            nxt = step_logits[-1].argmax(-1)
            gen.append(nxt)
            msk = torch.cat((msk, torch.ones(batch, 1, dtype=torch.bool)), 1)
            if t == new_tokens - 1:
                break
            r = m(input_ids=nxt[:, None], attention_mask=msk, past_key_values=past, use_cache=True, return_dict=True)
            past = r.past_key_values
            step_logits.append(r.logits[:, -1].float())
        sl = torch.stack(step_logits, 1)
        out["greedy_ids"] = torch.stack(gen, 1).numpy()           # EOS not special-cased: compare with suppress... see test
        out["greedy_step_logits_sub"] = sl[:, :, cols].numpy()
        out["greedy_step_top2gap"] = (sl.topk(2, -1).values[..., 0] - sl.topk(2, -1).values[..., 1]).numpy()
    np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), **out)
    print(f"{name}.npz: loss={out['loss']:.6f} size={os.path.getsize(os.path.join(GOLD, name + '.npz')) / 1e6:.2f} MB")


def golden_trained(name="traj_vlm_hd128_trained", batch=2, new_tokens=152, steps=400, lr=2e-2):
    """A fixture whose greedy decode is a REAL trajectory (every earlier greedy golden emitted 1-2
    distinct ids), on a trunk with the full-size head width (2 heads x 128) so the head_dim-128 tcgen05 attention, its
    backward and the decode attention are pinned to the reference, not only to op tests.

    Recipe (all of it reproducible from this file): det_state_dict weights; the REAL reference model is over-fitted on the
    synthetic batch for ``steps`` Adam steps, training only rows 32000..32261 of ``embed_tokens`` / ``lm_head`` (point +
    trajectory vocabulary) and the RMSNorm weights - small enough to store.  The trained values are rounded to bf16, loaded
    back, and every golden quantity is produced by the reference from the ROUNDED weights."""
    llm = LlmCfg.tiny_hd128()
    m = ref_shim.build_reference_model(llm)
    sd = synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True)
    m.load_state_dict(sd, strict=True)
    cfg = TrajVLMConfig.tiny_hd128()
    bt = synth.make_batch(cfg, batch, seed=1)
    ids, mask, pc = bt["tokens"], bt["attention_masks"], bt["pcrgbs"]
    P = bt["prompts"].shape[1]
    lo, hi = 32000, llm.vocab_size
    for p_ in m.parameters():
        p_.requires_grad = False
    norm_names = [k for k, _ in m.named_parameters() if k.endswith("layernorm.weight") or k == "model.norm.weight"]
    named = dict(m.named_parameters())
    train = [named["model.embed_tokens.weight"], named["lm_head.weight"]] + [named[k] for k in norm_names]
    for p_ in train:
        p_.requires_grad = True
    m.eval()                                                   # no dropout anywhere; encoder on running stats
    opt = torch.optim.Adam(train, lr=lr)
    rowmask = torch.zeros(llm.vocab_size, 1)
    rowmask[lo:hi] = 1.0

    def loss_of(logits):
        lg = logits[:, P - 1:-1, :]
        return torch.nn.functional.cross_entropy(lg.reshape(-1, lg.shape[-1]), ids[:, P:].flatten(), ignore_index=llm.pad_token_id)

    for it in range(steps):
        opt.zero_grad(set_to_none=True)
        loss = loss_of(_ref_forward(m, ids, mask, pc).logits)
        loss.backward()
        named["model.embed_tokens.weight"].grad.mul_(rowmask)
        named["lm_head.weight"].grad.mul_(rowmask)
        opt.step()
        if it % 25 == 0 or it == steps - 1:
            print(f"  overfit step {it}: loss {loss.item():.5f}", flush=True)
    with torch.no_grad():
        for p_ in train:
            p_.copy_(p_.to(torch.bfloat16).float())
    delta = {"rows_lo": np.int64(lo), "embed_rows": named["model.embed_tokens.weight"][lo:hi].detach().numpy().copy(),
             "lm_head_rows": named["lm_head.weight"][lo:hi].detach().numpy().copy()}
    for k in norm_names:
        delta[f"norm/{k}"] = named[k].detach().numpy().copy()
    # ---- goldens from the rounded weights, default freezing of model_arch.py:33-51
    for p_ in m.parameters():
        p_.requires_grad = True
    for p_ in m.model.point_backbone.parameters():
        p_.requires_grad = False
    for p_ in m.model.layers.parameters():
        p_.requires_grad = False
    m.train(); m.model.layers.eval(); m.model.point_backbone.eval()
    m.zero_grad(set_to_none=True)
    out = dict(fps_seed=np.int64(FPS_SEED), fps_start=_fps_start(pc.shape[1], batch).numpy(), tokens=ids.numpy(),
               prompt_len=np.int64(P), torch_version=np.array(torch.__version__),
               transformers_version=np.array(__import__("transformers").__version__), **delta)
    logits = _ref_forward(m, ids, mask, pc).logits
    loss = loss_of(logits)
    loss.backward()
    L = logits.detach()
    rng = np.random.default_rng(3)
    cols = np.sort(np.concatenate([rng.choice(32000, 256, replace=False), np.arange(32000, 32262)]))
    rows = np.sort(np.concatenate([np.arange(0, 8), np.arange(510, 530), np.arange(P - 4, L.shape[1])]))
    out.update(logit_rows=rows, logit_cols=cols, logits_sub=L[:, rows][:, :, cols].numpy(), logits_lse=torch.logsumexp(L, -1).numpy(),
               loss=np.float64(loss.item()))
    for k, p_ in m.named_parameters():
        if p_.grad is None:
            continue
        g = p_.grad.detach()
        out[f"gradnorm/{k}"] = np.float64(g.norm().item())
        flat = g.flatten()
        sel = torch.from_numpy(rng.choice(flat.numel(), min(4096, flat.numel()), replace=False)).long()
        out[f"gradidx/{k}"] = sel.numpy()
        out[f"gradval/{k}"] = flat[sel].numpy()
    # ---- manual KV-cache greedy decode with EOS handling of HF greedy search
    m.eval()
    _, _, ru = ref_shim.load()
    with torch.no_grad():
        msk = bt["prompt_masks"].clone()
        r = _ref_forward(m, bt["prompts"].clone(), msk, pc, use_cache=True)
        past = r.past_key_values
        step_logits = [r.logits[:, -1].float()]
        unfinished = torch.ones(batch, dtype=torch.long)
        gen = []
        for t in range(new_tokens):
            nxt = step_logits[-1].argmax(-1)
            nxt = nxt * unfinished + llm.pad_token_id * (1 - unfinished)
            gen.append(nxt)
            unfinished = unfinished & (nxt != llm.eos_token_id).long()
            msk = torch.cat((msk, torch.ones(batch, 1, dtype=torch.bool)), 1)
            if t == new_tokens - 1 or unfinished.max() == 0:
                break
            r = m(input_ids=nxt[:, None], attention_mask=msk, past_key_values=past, use_cache=True, return_dict=True)
            past = r.past_key_values
            step_logits.append(r.logits[:, -1].float())
        gen = torch.stack(gen, 1)
        sl = torch.stack(step_logits, 1)
        out["greedy_ids"] = gen.numpy()
        out["greedy_step_logits_sub"] = sl[:, :, cols].numpy()
        t2 = sl.topk(2, -1).values
        out["greedy_step_top2gap"] = (t2[..., 0] - t2[..., 1]).numpy()
        # post-processing of train.py:240-246 with the REAL str_to_float on the decoded string (tokens spelled as the
        # tokenizer of builder.py:33-46 would: "<pK>" joined by spaces, "<tsep>" separators)
        maxmin = [2.5, 0.1]
        names = {32003: "<ts>", 32004: "<tsep>", 32005: "<te>"}
        for b in range(batch):
            toks = gen[b].tolist()
            if llm.eos_token_id in toks:
                toks = toks[:toks.index(llm.eos_token_id)]
            s, prev_bin = "", False
            for t_ in toks:
                if 32006 <= t_ < 32262:
                    s += (" " if prev_bin else "") + f"<p{t_ - 32006}>"
                    prev_bin = True
                else:
                    s += names.get(t_, "?")
                    prev_bin = False
            rr = ru.str_to_float(s, maxmin, "val", rt2=True, num_bins=256)
            out[f"greedy_poses_rt2/{b}"] = np.zeros((0, 6), np.float32) if rr is None else rr
            o = tt.ids_to_poses(gen[b].tolist())
            out[f"greedy_poses_raw/{b}"] = np.zeros((0, 6), np.float32) if o is None else o
        out["maxmin"] = np.array(maxmin)
    agree = (gen == ids[:, P:P + gen.shape[1]]).float().mean().item()
    np.savez_compressed(os.path.join(GOLD, f"{name}.npz"), **out)
    print(f"{name}.npz: loss={out['loss']:.6f} greedy==training sequence on {agree:.3f} of steps, "
          f"distinct ids {len(np.unique(out['greedy_ids']))}, min top2 gap {out['greedy_step_top2gap'].min():.3f}, "
          f"size={os.path.getsize(os.path.join(GOLD, name + '.npz')) / 1e6:.2f} MB")


def golden_metrics():
    """Outputs of the REAL egoscaler/models/utils/metrics.py (fastdtw stubbed: not installed, DTW is not on the path) on
    seeded [T, 6] trajectory pairs of unequal lengths, incl. tiny rotation angles and identical rotations."""
    import importlib.util
    import types
    sys_mod = __import__("sys")
    if "fastdtw" not in sys_mod.modules:
        st = types.ModuleType("fastdtw"); st.fastdtw = lambda *a, **k: (0.0, [])
        sys_mod.modules["fastdtw"] = st
    spec = importlib.util.spec_from_file_location("ref_metrics", os.path.join(ref_shim.REF_ROOT, "egoscaler/models/utils/metrics.py"))
    rm = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rm)
    rng = np.random.default_rng(11)
    B, TG, TT = 24, 20, 20
    gen = rng.uniform(-1, 1, (B, TG, 6)).astype(np.float32)
    gt = rng.uniform(-1, 1, (B, TT, 6)).astype(np.float32)
    gen[:, :, 3:] *= np.float32(np.pi)
    gt[:, :, 3:] *= np.float32(np.pi)
    gen[0, :, 3:] = gt[0, :, 3:]                        # identical rotations -> dot = 1 (clip / arccos edge)
    gen[1, :, 3:] *= np.float32(1e-4); gt[1, :, 3:] *= np.float32(1e-4)      # small-angle branch of from_rotvec
    gen[2, :, 3:] = 0; gt[2, :, 3:] = 0
    gen_len = rng.integers(1, TG + 1, B).astype(np.int32)
    gt_len = rng.integers(1, TT + 1, B).astype(np.int32)
    gen_len[:4] = [TG, 5, 20, 1]; gt_len[:4] = [TT, 20, 7, 20]
    out = dict(gen=gen, gt=gt, gen_len=gen_len, gt_len=gt_len)
    res = np.zeros((B, 4), np.float64)
    for b in range(B):
        g, t = gen[b, :gen_len[b]], gt[b, :gt_len[b]]
        res[b] = [rm.average_displacement_error(g, t), rm.final_displacement_error(g, t), rm.initial_displacement_error(g, t),
                  rm.anglar_distance(g[:, 3:], t[:, 3:])]
    out["ade_fde_ide_gd_all6"] = res
    res3 = np.zeros((B, 3), np.float64)
    for b in range(B):
        g, t = gen[b, :gen_len[b], :3], gt[b, :gt_len[b], :3]
        res3[b] = [rm.average_displacement_error(g, t), rm.final_displacement_error(g, t), rm.initial_displacement_error(g, t)]
    out["ade_fde_ide_xyz"] = res3
    np.savez_compressed(os.path.join(GOLD, "traj_metrics.npz"), **out)
    print("traj_metrics.npz", res[:3])


def _ref_function(relpath, name, **ns):
    """Compile ONE function of a reference file in isolation (its module imports open3d / transformers at file scope, which
    are absent or irrelevant here): the function body that runs is the reference's own, read from REF_ROOT at run time."""
    import ast
    path = os.path.join(ref_shim.REF_ROOT, relpath)
    tree = ast.parse(open(path).read(), path)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == name]
    assert len(fn) == 1, (relpath, name)
    g = dict(np=np, **ns)
    exec(compile(ast.Module(body=fn, type_ignores=[]), path, "exec"), g)
    return g[name]


def golden_rgbd_frontend():
    """Outputs of the REAL get_points_colors (data/tools/pcm_tools.py:68-96) and pc_norm (pointllm/data/utils.py:146-157) on
    seeded RGB-D frames: with / without bboxes and depth threshold, a non-square frame, an all-invalid frame."""
    from oracle.rgbd_frontend import synth_rgbd
    gpc = _ref_function("egoscaler/data/tools/pcm_tools.py", "get_points_colors")
    pcn = _ref_function("egoscaler/models/pointllm/pointllm/data/utils.py", "pc_norm")
    cases = [  # name, H, W, boxes, c, fx, fy, d_thres
        ("plain", 40, 40, None, 19.5, 30.25, 30.25, None),
        ("boxes_thres", 64, 48, [(5, 7, 20, 30), (30, 0, 48, 10), (10, 50, 12, 64)], 23.5, 41.0, 39.5, 4.0),
        ("aria_crop", 33, 57, [(0, 0, 57, 3)], 703.5, 605.343, 605.343, 3.0),
        ("all_boxed", 16, 16, [(0, 0, 16, 16)], 7.5, 10.0, 10.0, None),
    ]
    out = {}
    rng = np.random.default_rng(3)
    for k, (name, H, W, boxes, c, fx, fy, thr) in enumerate(cases):
        rgbd = synth_rgbd(H, W, seed=100 + k)
        bb = None if boxes is None else [{"box": dict(xmin=b[0], ymin=b[1], xmax=b[2], ymax=b[3])} for b in boxes]
        pts, cols = gpc(rgbd=rgbd, bbox=bb, width=W, height=H, principal_p=c, focal_len_x=fx, focal_len_y=fy, d_thres=thr)
        out[f"{name}/rgbd"] = rgbd
        out[f"{name}/boxes"] = np.zeros((0, 4), np.int32) if boxes is None else np.asarray(boxes, np.int32)
        out[f"{name}/cam"] = np.asarray([c, fx, fy, np.nan if thr is None else thr], np.float64)
        out[f"{name}/points"] = pts
        out[f"{name}/colors"] = cols
        assert pts.dtype == np.float64 and cols.dtype == np.float32, (pts.dtype, cols.dtype)
        if len(pts) >= 64:
            idx = rng.choice(len(pts), size=min(512, len(pts)), replace=False)
            pc = np.concatenate([pts[idx], cols[idx].astype(np.float64)], axis=1)
            out[f"{name}/sample_idx"] = idx.astype(np.int64)
            out[f"{name}/pc_norm"] = pcn(pc)
        print(name, rgbd.dtype, pts.shape, cols.shape)
    np.savez_compressed(os.path.join(GOLD, "rgbd_frontend.npz"), **out)


def golden_traj_preprocess():
    """Outputs of the REAL egoscaler/models/utils/traj_utils.py (pure numpy) on seeded float32 trajectories of every edge
    length (1..6) and longer ones, for num_steps 20 (the reference's) and 7."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_traj_utils", os.path.join(ref_shim.REF_ROOT, "egoscaler/models/utils/traj_utils.py"))
    tu = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(tu)
    rng = np.random.default_rng(21)
    lens = np.asarray([1, 2, 3, 4, 5, 6, 7, 19, 20, 21, 33, 80, 81, 64], np.int32)
    T = int(lens.max())
    traj = rng.uniform(-2, 2, (len(lens), T, 6)).astype(np.float32)
    out = dict(traj=traj, lens=lens)
    for ns in (20, 7):
        for smooth in (0, 1):
            res = np.zeros((len(lens), ns, 6), np.float32)
            msk = np.zeros((len(lens), ns), np.int32)
            for b, n in enumerate(lens):
                t = traj[b, :n]
                if smooth:
                    t = tu.smoothing_traj(t)
                    assert t.dtype == np.float32
                r, m = tu.preprocess_traj(t, ns, return_padding_mask=True)
                res[b], msk[b] = r, m
            out[f"out/{ns}/{smooth}"] = res
            out[f"mask/{ns}/{smooth}"] = msk
    np.savez_compressed(os.path.join(GOLD, "traj_preprocess.npz"), **out)
    print("traj_preprocess.npz", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    import sys
    os.makedirs(GOLD, exist_ok=True)
    torch.set_num_threads(os.cpu_count())
    if len(sys.argv) > 1 and sys.argv[1] == "trained":
        golden_trained()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "hd128":
        # 2 heads x 128: forward / loss / gradients of EVERY LLM-layer tensor / greedy ids through the head_dim-128 kernels
        golden_variant("traj_vlm_hd128_unfrozen_lm", unfreeze_lm=True, hd128=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "metrics":
        golden_metrics()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "traj_preprocess":
        golden_traj_preprocess()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "rgbd":
        golden_rgbd_frontend()
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "unfrozen_pc":
        golden_variant("traj_vlm_tiny_unfrozen_pc", unfreeze_pc=True)
        golden_variant("traj_vlm_tiny_base_unfrozen_pc", point_cfg_name="PointTransformer_base_8192point", point="base",
                       unfreeze_pc=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "variants":
        golden_variant("traj_vlm_tiny_unfrozen_lm", unfreeze_lm=True)
        golden_variant("traj_vlm_tiny_base_encoder", point_cfg_name="PointTransformer_base_8192point", point="base")
        sys.exit(0)
    torch.set_num_threads(os.cpu_count())
    golden_tokenizer()
    golden_model()
