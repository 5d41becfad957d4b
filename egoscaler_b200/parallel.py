"""Data-parallel exchange of the training step (host logic; device-agnostic so it is testable with gloo on CPU).

Reference: DeepSpeed ZeRO-1 bf16 data parallelism configured at train.py:92-125 (gradients averaged over ranks, optimizer
states partitioned, updated parameters gathered).  Here: one process per GPU, ``torch.distributed`` (NCCL over NVLink on
the GPU box), batch sharded by rank (the reference forgot the DistributedSampler, train.py:72-77 / SURVEY 9.6).

* default recipe (275 M trainable parameters): every gradient is sum-all-reduced, asynchronously, as soon as it is final
  (head gradients before the 32-layer trunk backward starts) and scaled by 1/world inside the AdamW kernel;
* ``--unfreeze_language_model``: each LLM-layer tensor ("unit") has an OWNER rank ``u % world`` that alone keeps its fp32
  master / moments (ZeRO-1); its gradient is sum-REDUCED to the owner when the layer's backward has been queued and the
  updated bf16 weight is BROADCAST from the owner after the step (reduce-scatter + all-gather at tensor granularity).
"""
from __future__ import annotations

from typing import Iterable, List, Optional, Sequence

import torch


def shard_bounds(n: int, rank: int, world: int) -> "tuple[int, int]":
    """Contiguous shard [lo, hi) of ``n`` samples for ``rank`` (sizes differ by at most one, earlier ranks larger)."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(batch: dict, rank: int, world: int) -> dict:
    """Rank's slice of a collate_fn-shaped global batch (dataset.py:184-194): tensors are split on dim 0."""
    n = next(v.shape[0] for v in batch.values() if torch.is_tensor(v))
    lo, hi = shard_bounds(n, rank, world)
    return {k: (v[lo:hi] if torch.is_tensor(v) and v.dim() > 0 and v.shape[0] == n else v) for k, v in batch.items()}


class GradExchange:
    def __init__(self, world_size: int = 1, process_group=None):
        self.world = max(1, int(world_size))
        self.pg = process_group
        self.rank = 0
        self._pending: List = []
        self._named: dict = {}                   # name -> pending work (allreduce_async_named / wait_named)
        self._reduce_is_allreduce = False
        if self.world > 1:
            import torch.distributed as dist
            self.dist = dist
            self.rank = dist.get_rank(process_group)
            # gloo has no device-tensor ``reduce`` (NCCL has): the single-GPU two-process Trainer test runs the owner
            # reduction as an all-reduce there - same values on the owner, only more traffic
            self._reduce_is_allreduce = dist.get_backend(process_group) == "gloo"

    # ---- ownership (ZeRO-1)
    def owner(self, unit: int) -> int:
        return unit % self.world

    def owns(self, unit: int) -> bool:
        return self.owner(unit) == self.rank

    def _global(self, r: int) -> int:
        return self.dist.get_global_rank(self.pg, r) if self.pg is not None else r

    # ---- collectives (all asynchronous; ``wait`` joins them)
    def allreduce_async(self, tensors: Iterable[torch.Tensor]) -> None:
        if self.world == 1:
            return
        for t in tensors:
            self._pending.append(self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.pg, async_op=True))

    def allreduce_async_named(self, named) -> None:
        """Like ``allreduce_async`` for (name, tensor) pairs; ``wait_named(name)`` joins one of them, so the optimizer can start on
        the tensors whose exchange has finished while later ones are still in flight."""
        if self.world == 1:
            return
        for name, t in named:
            self._named.setdefault(name, []).append(self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.pg, async_op=True))

    def wait_named(self, name: str) -> None:
        for w in self._named.pop(name, []):
            w.wait()

    def all_gather_async(self, t: torch.Tensor):
        """(out [world, *t.shape], work or None): ``all_gather`` without the wait (``work.wait()`` before ``out`` is read)."""
        if self.world == 1 or (self._reduce_is_allreduce and t.is_cuda):
            return self.all_gather(t), None
        out = torch.empty((self.world,) + tuple(t.shape), device=t.device, dtype=t.dtype)
        return out, self.dist.all_gather_into_tensor(self._gather_view(out, t), t.contiguous(), group=self.pg, async_op=True)

    def reduce_to_owner_async(self, unit: int, tensor: torch.Tensor) -> None:
        if self.world == 1:
            return
        if self._reduce_is_allreduce and tensor.is_cuda:
            self._pending.append(self.dist.all_reduce(tensor, op=self.dist.ReduceOp.SUM, group=self.pg, async_op=True))
            return
        self._pending.append(self.dist.reduce(tensor, dst=self._global(self.owner(unit)), op=self.dist.ReduceOp.SUM,
                                              group=self.pg, async_op=True))

    def all_gather(self, t: torch.Tensor) -> torch.Tensor:
        """[world, *t.shape] with every rank's ``t`` (blocking; same shape on all ranks).  gloo cannot all-gather device
        tensors: there the gather is an all-reduce of a zero-padded buffer (single-GPU two-process test path only)."""
        if self.world == 1:
            return t.unsqueeze(0)
        out = torch.zeros((self.world,) + tuple(t.shape), device=t.device, dtype=t.dtype)
        if self._reduce_is_allreduce and t.is_cuda:
            out[self.rank].copy_(t)
            self.dist.all_reduce(out, op=self.dist.ReduceOp.SUM, group=self.pg)
        else:
            self.dist.all_gather_into_tensor(self._gather_view(out, t), t.contiguous(), group=self.pg)
        return out

    def _gather_view(self, out: torch.Tensor, t: torch.Tensor) -> torch.Tensor:
        """gloo (CPU tensors, the host-logic tests) only accepts the CONCATENATED output form [world * n, ...]; NCCL takes the
        stacked [world, n, ...] buffer as it is."""
        if t.is_cuda or t.dim() == 0:
            return out
        return out.view((self.world * t.shape[0],) + tuple(t.shape[1:]))

    def broadcast_from_owners(self, units: Sequence["tuple[int, torch.Tensor]"]) -> None:
        if self.world == 1:
            return
        hs = [self.dist.broadcast(t, src=self._global(self.owner(u)), group=self.pg, async_op=True) for u, t in units]
        for h in hs:
            h.wait()

    def wait(self) -> None:
        for w in self._pending:
            w.wait()
        self._pending = []
        for name in list(self._named):
            self.wait_named(name)

    @property
    def grad_scale(self) -> float:
        """DDP / DeepSpeed average: gradients are summed by the collective and scaled by 1/world in the optimizer."""
        return 1.0 / self.world
