"""GPU: the multi-rank Trainer path end to end: two processes share cuda:0 and talk over gloo (NCCL refuses two
ranks on one device; the N-GPU NCCL run of the same check is ``bench.py --verify-dp`` under torchrun, recorded in
profiles/).  Covers Trainer._sync_initial_state (ranks are seeded DIFFERENTLY on purpose), the async all-reduce of the
default recipe, and for --unfreeze_language_model the ZeRO-1 pieces: _reduce_layer_async, owner-only masters, _trunk_step,
broadcast_from_owners.  Reference semantics: DeepSpeed ZeRO-1 data parallelism, train.py:92-125,183-184:
post-exchange gradients == single-process gradients on the concatenated batch, weights identical on all ranks after the step."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _build(unfreeze_lm, seed):
    from egoscaler_b200 import synth
    from egoscaler_b200.config import TrajVLMConfig
    from egoscaler_b200.model import TrajPointLLMForCausalLM
    cfg = TrajVLMConfig.tiny_hd128()
    m = TrajPointLLMForCausalLM(cfg, "cuda", unfreeze_language_model=unfreeze_lm)
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=seed, round_bf16=True), strict=True)
    m.train()
    return cfg, m


def _grads(tr):
    out = {n: tr.grad[n].clone() for n in tr.names}
    for u, (k, li, _) in enumerate(tr.trunk_units):
        if tr.dp.owns(u):
            out[f"unit{u}"] = tr.wgrad[k][li].clone()
    return out


def _worker(rank, world, port, unfreeze_lm, q):
    import torch.distributed as dist
    from egoscaler_b200 import synth
    from egoscaler_b200.trainer import Trainer
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.cuda.set_device(0)
        cfg, m = _build(unfreeze_lm, seed=rank * 7)             # rank 1 starts from DIFFERENT weights: the initial sync must fix it
        tr = Trainer(m, lr=1e-3, world_size=world)
        bt = synth.make_batch(cfg, 2 * world, seed=3)
        P = bt["prompts"].shape[1]
        start = torch.tensor([5, 77, 1000, 4000])[:2 * world]
        sl = slice(2 * rank, 2 * rank + 2)
        tr.zero_grad()
        # two micro-batches of one sample each: the first without exchange (grad accumulation), the second with
        tr.forward_backward(bt["tokens"][sl][:1], bt["attention_masks"][sl][:1], bt["pcrgbs"][sl][:1], P, start[sl][:1], sync=False)
        loss = tr.forward_backward(bt["tokens"][sl][1:], bt["attention_masks"][sl][1:], bt["pcrgbs"][sl][1:], P, start[sl][1:], sync=True)
        tr._wait_reduce()
        g = {k: v.cpu().numpy() for k, v in _grads(tr).items()}        # numpy: pickled by value (torch tensors travel as shm handles)
        tr.optimizer_step()
        torch.cuda.synchronize()
        w = {n: tr.param[n].detach().float().cpu().numpy() for n in tr.names}
        for u, (k, li, t) in enumerate(tr.trunk_units):
            w[f"unit{u}"] = t.detach().float().cpu().numpy()
        q.put((rank, float(loss), g, w))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(900)
@pytest.mark.parametrize("unfreeze_lm", [False, True])
def test_two_rank_trainer_matches_single_process(unfreeze_lm):
    from egoscaler_b200 import synth
    from egoscaler_b200.trainer import Trainer
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, unfreeze_lm, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=800) for _ in range(world)], key=lambda r: r[0])
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    # single process, world 1, rank 0's weights, the concatenated batch as 4 accumulated micro-batches of one sample
    cfg, m = _build(unfreeze_lm, seed=0)
    tr = Trainer(m, lr=1e-3)
    bt = synth.make_batch(cfg, 4, seed=3)
    P = bt["prompts"].shape[1]
    start = torch.tensor([5, 77, 1000, 4000])
    tr.zero_grad()
    for i in range(4):
        tr.forward_backward(bt["tokens"][i:i + 1], bt["attention_masks"][i:i + 1], bt["pcrgbs"][i:i + 1], P, start[i:i + 1])
    ref = {n: tr.grad[n].cpu() for n in tr.names}
    for u, (k, li, _) in enumerate(tr.trunk_units):
        ref[f"unit{u}"] = tr.wgrad[k][li].cpu()
    n_cmp = 0
    for r in range(world):
        for name, g in res[r][2].items():
            den = float(ref[name].double().norm())
            if den == 0.0:
                continue
            rel = float((torch.from_numpy(g).double() - ref[name].double()).norm()) / den
            assert rel <= 2e-2, (r, name, rel)
            n_cmp += 1
    assert n_cmp >= (2 * 9 if not unfreeze_lm else 9 + 9 + 12)
    owned = [set(k for k in res[r][2] if k.startswith("unit")) for r in range(world)]
    if unfreeze_lm:
        assert owned[0] and owned[1] and not (owned[0] & owned[1])        # ZeRO-1: disjoint ownership, both ranks own units
        assert len(owned[0] | owned[1]) == len(tr.trunk_units)
    # after the step every rank holds bit-identical weights (all-reduce / broadcast-from-owner) ...
    for name in res[0][3]:
        assert np.array_equal(res[0][3][name], res[1][3][name]), name
    # ... equal to the single-process step on the same (summed, 1/world-scaled vs 4-sample-summed) gradients up to bf16
    tr.world = 1
    for g in list(tr.grad.values()) + [x for lst in (tr.wgrad or {}).values() for x in lst]:
        g.mul_(0.5)                                                         # ranks average 2 shards; each shard SUMMED 2 micro-batches
    tr.optimizer_step()
    for n in tr.names:
        a, b = torch.from_numpy(res[0][3][n]), tr.param[n].detach().float().cpu()
        assert float((a - b).abs().max()) <= 4e-3 * max(1.0, float(b.abs().max())), n
