"""GPU: the module as a DROP-IN for the reference's entry points (VERDICT r1 missing #1):

* ``build_model(args)`` (builder.py:49-56) and the ``TrajPointLLMForCausalLM(args, config, model_name)`` constructor
  (model_arch.py:13), including loading a saved ``state_dict`` from ``model_name``;
* the LITERAL loop body of train.py:165-184 - ``outputs = model(input_ids=..., attention_mask=..., point_clouds=...)``,
  ``F.cross_entropy(outputs.logits[:, P-1:-1] ...)``, ``loss.backward()``, ``torch.optim.AdamW(...).step()`` - must fill
  ``.grad`` of exactly the trainable parameters with the gradients the hand-written fast path (``Trainer``) computes and
  that the real reference produced (tests/golden), and an optimizer step through ``torch.optim`` must be picked up by the
  next forward (derived operand copies are rebuilt).
"""
import argparse
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from egoscaler_b200 import synth
from egoscaler_b200.builder import build_model
from egoscaler_b200.config import TrajVLMConfig
from egoscaler_b200.model import TrajPointLLMForCausalLM
from egoscaler_b200 import decode  # noqa: F401
from egoscaler_b200.trainer import Trainer
from conftest import GOLDEN


def _rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _literal_step(model, batch, pad_token_id, fps_start=None, optimizer=None):
    """train.py:157-184, verbatim except for the DeepSpeed engine calls (``model_engine.backward/step``)."""
    pcrgbs = batch['pcrgbs']
    tokens = batch['tokens']
    attention_masks = batch['attention_masks']
    prompts = batch['prompts']
    if optimizer is not None:
        optimizer.zero_grad()
    with torch.autocast("cuda", dtype=torch.bfloat16):
        outputs = model(input_ids=tokens, attention_mask=attention_masks, point_clouds=pcrgbs, return_dict=True,
                        **({} if fps_start is None else {"fps_start": fps_start}))
    logits = outputs.logits[:, prompts.shape[1] - 1:-1, :]
    tokens = tokens[:, prompts.shape[1]:]
    loss = F.cross_entropy(logits.reshape(-1, logits.shape[-1]), tokens.flatten(), ignore_index=pad_token_id)
    loss.backward()
    if optimizer is not None:
        optimizer.step()
    return loss


def _det_model(cfg, **kw):
    m = TrajPointLLMForCausalLM(cfg, "cuda", **kw)
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    return m


@pytest.mark.parametrize("recipe", ["default", "unfreeze_lm"])
def test_literal_training_loop_fills_grads_like_trainer_and_reference(recipe):
    gname = "traj_vlm_tiny.npz" if recipe == "default" else "traj_vlm_tiny_unfrozen_lm.npz"
    gm = np.load(os.path.join(GOLDEN, gname))
    cfg = TrajVLMConfig.tiny()
    kw = dict(unfreeze_language_model=(recipe == "unfreeze_lm"))
    bt = {k: v.cuda() for k, v in synth.make_batch(cfg, 2, seed=1).items()}
    start = torch.from_numpy(gm["fps_start"]).cuda()
    m = _det_model(cfg, **kw)
    m.train()
    loss = _literal_step(m, bt, cfg.pad_token_id, fps_start=start)
    assert abs(float(loss) - float(gm["loss"])) <= 2e-2 * float(gm["loss"])
    trainable = {n for n, p in m.named_parameters() if p.requires_grad}
    for n, p in m.named_parameters():
        assert (p.grad is not None) == (n in trainable), n                  # exactly the recipe's trainable set
    # (1) vs the real reference's gradients
    for n in sorted(trainable):
        ref = gm[f"gradval/{n}"]
        got = dict(m.named_parameters())[n].grad.float().flatten().cpu()[torch.from_numpy(gm[f"gradidx/{n}"])].numpy()
        cos = float((got * ref).sum() / (np.linalg.norm(got) * np.linalg.norm(ref) + 1e-30))
        assert cos >= 0.998, (n, cos)
    # (2) vs the fast path on the same weights / batch: same kernels except the head runs on all rows and torch's own
    #     cross-entropy backward produced dlogits (bf16 rounding of softmax - onehot differs from the fused kernel)
    m2 = _det_model(cfg, **kw)
    m2.train()
    tr = Trainer(m2, lr=2e-5)
    tr.zero_grad()
    loss2 = tr.forward_backward(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], bt["prompts"].shape[1], start)
    assert abs(float(loss) - float(loss2)) <= 1e-2 * float(loss2)        # the literal loop holds the loss in bf16
    named = dict(m.named_parameters())
    for n in tr.names:
        assert _rel(named[n].grad.float(), tr.grad[n]) <= 1.5e-2, (n, _rel(named[n].grad.float(), tr.grad[n]))
    if recipe == "unfreeze_lm":
        d = cfg.hidden_size
        assert _rel(named["model.layers.1.self_attn.k_proj.weight"].grad.float(), tr.wgrad["wqkv"][1][d:2 * d]) <= 2e-2
        assert _rel(named["model.layers.0.mlp.down_proj.weight"].grad.float(), tr.wgrad["wdown"][0]) <= 2e-2
    # a second backward through the same graph must fail loudly, not silently reuse freed activations
    out = m(input_ids=bt["tokens"], attention_mask=bt["attention_masks"], point_clouds=bt["pcrgbs"], fps_start=start)
    s = out.logits.float().sum()
    s.backward(retain_graph=True)
    with pytest.raises(RuntimeError):
        s.backward()


def test_torch_optimizer_steps_are_picked_up_by_the_next_forward():
    cfg = TrajVLMConfig.tiny()
    bt = {k: v.cuda() for k, v in synth.make_batch(cfg, 2, seed=1).items()}
    start = torch.tensor([17, 4242]).cuda()
    m = _det_model(cfg, unfreeze_language_model=True)
    m.train()
    opt = torch.optim.AdamW([p for p in m.parameters() if p.requires_grad], lr=1e-3)
    losses = [float(_literal_step(m, bt, cfg.pad_token_id, start, opt)) for _ in range(5)]
    assert losses[-1] < losses[0] - 0.05, losses                        # torch.optim.AdamW trains the CUDA path
    # derived copies follow the parameters: transposed trunk weights, projector biases, lm_head^T
    lw = m.trunk.layers[0]
    m.sync_derived()
    assert lw.woT is None and lw.wqkvT is None                             # NN dgrad: no contraction-major trunk copies to go stale
    assert torch.equal(m._packed_proj.b[0], dict(m.named_parameters())[m.projector_params()[0][1]].detach().float())
    # eval forward == fresh module loaded from the trained state_dict
    m.eval()
    with torch.no_grad():
        a = m(input_ids=bt["tokens"], attention_mask=bt["attention_masks"], point_clouds=bt["pcrgbs"], fps_start=start).logits
    m3 = TrajPointLLMForCausalLM(cfg, "cuda")
    m3.load_state_dict(m.state_dict(), strict=True)
    m3.eval()
    with torch.no_grad():
        b = m3(input_ids=bt["tokens"], attention_mask=bt["attention_masks"], point_clouds=bt["pcrgbs"], fps_start=start).logits
    assert torch.equal(a, b)


def test_build_model_and_reference_constructor(tmp_path):
    args = argparse.Namespace(model_name="synthetic:tiny", num_bins=256, unfreeze_pc_encoder=False, unfreeze_language_model=False)
    model, tokenizer, pbc, use_se = build_model(args)
    assert isinstance(model, TrajPointLLMForCausalLM) and use_se is True
    assert len(tokenizer) == 32262 == model.cfg.vocab_size and tokenizer.pad_token_id == 0
    assert pbc["point_token_len"] == 513 and pbc["point_patch_token"] == 32000 and pbc["point_start_token"] == 32001
    assert tokenizer.convert_tokens_to_ids(["<ts>", "<tsep>", "<te>", "<p0>", "<p255>"]) == [32003, 32004, 32005, 32006, 32261]
    assert tokenizer.decode([32003, 32006, 32007, 32004, 32005, 2]) == "<ts><p0> <p1><tsep><te></s>"
    assert not any(p.requires_grad for n, p in model.named_parameters() if n.startswith("model.layers."))
    assert model.model.embed_tokens.weight.requires_grad and model.lm_head.weight.requires_grad
    # the reference constructor with a checkpoint path: weights come from the saved state_dict (model_arch.py:26-31)
    ck = tmp_path / "pytorch_model.bin"
    torch.save({k: v.cpu() for k, v in model.state_dict().items()}, ck)
    args2 = argparse.Namespace(unfreeze_pc_encoder=False, unfreeze_language_model=True)
    m2 = TrajPointLLMForCausalLM(args2, model.cfg, str(tmp_path))
    for k, v in model.state_dict().items():
        assert torch.equal(m2.state_dict()[k], v), k
    assert all(p.requires_grad for n, p in m2.named_parameters() if n.startswith("model.layers."))
    # generate() through the built model (evaluate.py:114-124 call form)
    bt = synth.make_batch(model.cfg, 2, seed=1)
    model.eval()
    out = model.generate(input_ids=bt["prompts"].cuda(), attention_mask=bt["prompt_masks"].cuda(), point_clouds=bt["pcrgbs"].cuda(),
                         max_length=8, temperature=1.0, top_k=50, top_p=0.95, repetition_penalty=1.0, do_sample=True,
                         num_return_sequences=1)
    assert out.sequences.shape == (2, bt["prompts"].shape[1] + 8) and len(out.scores) == 8
