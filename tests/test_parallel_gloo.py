"""CPU, world_size 2 over gloo: the host-side data-parallel logic of the training step (egoscaler_b200/parallel.py),
exercised with CPU tensors and with the oracle's gradients standing in for the CUDA backward.
Reference semantics: DeepSpeed ZeRO-1 bf16 DP, train.py:92-125 (gradients averaged over ranks, states partitioned)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from egoscaler_b200.parallel import GradExchange, shard_batch, shard_bounds
from egoscaler_b200 import synth
from egoscaler_b200.config import TrajVLMConfig


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_bounds_cover_and_balance():
    for n in (1, 2, 7, 8, 64, 257):
        for w in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, w) for r in range(w)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1
    bt = synth.make_batch(TrajVLMConfig.tiny(), 4, seed=5)
    parts = [shard_batch(bt, r, 2) for r in range(2)]
    for k in ("tokens", "attention_masks", "pcrgbs", "prompts"):
        assert torch.equal(torch.cat([p[k] for p in parts]), bt[k])


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(100 + rank)
        dp = GradExchange(world)
        assert dp.rank == rank and dp.grad_scale == 0.5
        # (1) default recipe: async sum all-reduce of several gradient tensors
        g = [torch.full((5, 3), float(rank + 1)), torch.arange(7.0) * (rank + 1)]
        dp.allreduce_async(g)
        dp.wait()
        assert torch.equal(g[0], torch.full((5, 3), 3.0)) and torch.equal(g[1], torch.arange(7.0) * 3)
        # (1b) named all-reduces (the optimizer waits per tensor / per bucket), asynchronous gathers (embedding rows + ids)
        a, bkt = torch.full((3,), float(rank + 1)), torch.arange(6.0) + rank
        dp.allreduce_async_named([("lm_head.weight", a), ("bucket:point_proj", bkt)])
        rows, w_rows = dp.all_gather_async(torch.full((2, 4), float(rank)))
        ids, w_ids = dp.all_gather_async(torch.tensor([10 + rank, 20 + rank]))
        assert set(dp._named) == {"lm_head.weight", "bucket:point_proj"}
        dp.wait_named("lm_head.weight")
        assert torch.equal(a, torch.full((3,), 3.0)) and "lm_head.weight" not in dp._named
        dp.wait_named("not pending")                                                 # no-op
        dp.wait()                                                                    # joins whatever is left, named included
        assert torch.equal(bkt, 2 * torch.arange(6.0) + 1) and not dp._named
        for w in (w_rows, w_ids):
            if w is not None:
                w.wait()
        assert rows.shape == (2, 2, 4) and torch.equal(rows[1], torch.ones(2, 4)) and torch.equal(rows[0], torch.zeros(2, 4))
        assert ids.tolist() == [[10, 20], [11, 21]]
        # (2) ZeRO-1: reduce to owner, owner "steps", broadcast the updated weights
        units = [torch.full((4,), float(10 * u + rank)) for u in range(5)]          # gradients
        weights = [torch.zeros(4) for _ in range(5)]
        for u, t in enumerate(units):
            dp.reduce_to_owner_async(u, t)
        dp.wait()
        for u, t in enumerate(units):
            if dp.owns(u):
                assert torch.equal(t, torch.full((4,), float(20 * u + 1)))           # sum over the two ranks
                weights[u] -= dp.grad_scale * t                                      # owner-only update
        dp.broadcast_from_owners(list(enumerate(weights)))
        for u, w in enumerate(weights):
            assert torch.equal(w, torch.full((4,), -0.5 * (20 * u + 1))), (rank, u)
        assert [dp.owner(u) for u in range(5)] == [0, 1, 0, 1, 0]
        # (3) DP semantics on the real model arithmetic (oracle gradients): mean over ranks of per-rank mean losses
        from oracle.traj_vlm import LlmCfg, PointCfg, TrajVLMOracle, train_loss
        torch.manual_seed(0)
        m = TrajVLMOracle(LlmCfg.tiny(), PointCfg())
        m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
        m.configure_trainable()
        m.eval()
        bt = synth.make_batch(TrajVLMConfig.tiny(), world, seed=1)
        P = bt["prompts"].shape[1]
        start = torch.zeros(world, dtype=torch.long)
        mine = shard_batch({**bt, "start": start}, rank, world)
        loss = train_loss(m(mine["tokens"], mine["attention_masks"], mine["pcrgbs"], fps_start=mine["start"], training=True),
                          mine["tokens"], P, 0)
        loss.backward()
        names = ["model.norm.weight", "lm_head.weight"]
        grads = {n: dict(m.named_parameters())[n].grad.clone() for n in names}
        dp.allreduce_async(grads.values())
        dp.wait()
        q.put((rank, float(loss), {n: (g * dp.grad_scale).numpy() for n, g in grads.items()}))
    except Exception:                                            # report instead of leaving the parent to time out
        import traceback
        q.put((rank, None, traceback.format_exc()))
        raise
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_grad_exchange_world2_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=500) for _ in range(world)]
    for r in res:
        assert r[1] is not None, f"rank {r[0]} failed:\n{r[2]}"
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort(key=lambda r: r[0])
    # every rank holds the same averaged gradient ...
    for n in res[0][2]:
        assert np.array_equal(res[0][2][n], res[1][2][n])
    # ... which equals the single-process gradient of the mean of the per-rank losses on the concatenated batch
    from oracle.traj_vlm import LlmCfg, PointCfg, TrajVLMOracle, train_loss
    torch.manual_seed(0)
    m = TrajVLMOracle(LlmCfg.tiny(), PointCfg())
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    m.configure_trainable()
    m.eval()
    bt = synth.make_batch(TrajVLMConfig.tiny(), world, seed=1)
    P = bt["prompts"].shape[1]
    tot = 0.0
    for r in range(world):
        sl = slice(r, r + 1)
        l = train_loss(m(bt["tokens"][sl], bt["attention_masks"][sl], bt["pcrgbs"][sl], fps_start=torch.zeros(1, dtype=torch.long),
                         training=True), bt["tokens"][sl], P, 0)
        assert abs(float(l) - res[r][1]) < 1e-5
        tot = tot + l / world
    tot.backward()
    for n in res[0][2]:
        ref = dict(m.named_parameters())[n].grad.numpy()
        np.testing.assert_allclose(res[0][2][n], ref, atol=1e-7 + 1e-4 * np.abs(ref).max(), rtol=1e-4)
