"""Training step of the trajectory VLM on the CUDA path (forward + hand-written backward + fused AdamW), with
data-parallel gradient all-reduce over NCCL overlapped with the backward.

Reference: train.py:153-197 -- forward under autocast(bf16) (:166-172), loss on ``logits[:, P-1:-1]`` vs
``tokens[:, P:]`` with ``ignore_index=pad`` (:174-181), ``model_engine.backward`` / ``.step`` (:183-184) with
torch.optim.AdamW defaults (:113), linear warm-up schedule (:114-117), DeepSpeed ZeRO-1 bf16 data parallelism
(:92-125).  Default trainable set (model_arch.py:33-51): embed_tokens, lm_head, point_proj, model.norm.

Only the T = L - P trajectory rows go through lm_head (the loss never reads the others).  Data parallelism: one
process per GPU, batch sharded by rank, trainable gradients summed with ``torch.distributed.all_reduce`` (NCCL over
NVLink); the lm_head/norm gradients -- half of the bytes -- are reduced while the 32-layer backward is still running.
"""
from __future__ import annotations

import os

from typing import Dict, List, Optional

import torch

from . import ops
from .autograd import backward_from_hidden, forward_hidden_train
from .model import TrajPointLLMForCausalLM
from .parallel import GradExchange


def linear_warmup_lr(base_lr: float, step: int, warmup: int, total: int) -> float:
    """transformers.get_linear_schedule_with_warmup (train.py:114-117); ``step`` counts completed optimizer steps."""
    if total <= 0:
        return base_lr
    if step < warmup:
        return base_lr * step / max(1, warmup)
    return base_lr * max(0.0, (total - step) / max(1, total - warmup))


class Trainer:
    def __init__(self, model: TrajPointLLMForCausalLM, lr: float = 2e-5, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, warmup_steps: int = 0, total_steps: int = 0, process_group=None,
                 world_size: int = 1):
        self.model = model
        self.enc_train = model.enc_train()  # --unfreeze_pc_encoder (model_arch.py:36-38,121-122): encoder_train.py
        self.lr, self.betas, self.eps, self.wd = lr, betas, eps, weight_decay
        self.warmup, self.total = warmup_steps, total_steps
        self.pg, self.world = process_group, world_size
        self.step_count = 0
        if world_size > 1:
            self._sync_initial_state()
        named = dict(model.named_parameters())
        # LLM-layer tensors are optimised as the PACKED kernel operands (q|k|v and gate|up share one buffer each; AdamW is
        # element-wise, so this equals stepping the reference's separate parameters); everything else by parameter name.
        self.names: List[str] = [n for n in model.trainable_names() if not n.startswith("model.layers.")]
        self.dp = GradExchange(world_size, process_group)
        self.rank = self.dp.rank
        self.trunk_units = []               # (kind, layer, bf16 tensor); --unfreeze_language_model (model_arch.py:40-45)
        self.wgrad = None
        if model.unfreeze_language_model:
            kinds = (("wqkv", "wqkv"), ("wo", "wo"), ("wgu", "wgu"), ("wdown", "wdown"), ("ln1", "ln1"), ("ln2", "ln2"))
            self.wgrad = {k: [] for k, _ in kinds}
            for li, lw in enumerate(model.trunk.layers):
                for k, attr in kinds:
                    t = getattr(lw, attr)
                    self.trunk_units.append((k, li, t))
                    self.wgrad[k].append(torch.zeros(t.shape, device=t.device, dtype=torch.float32))
            # ZeRO-1 (train.py:99-103): optimizer states of unit u live on rank u % world only
            self.t_master, self.t_m, self.t_v = {}, {}, {}
            for u, (k, li, t) in enumerate(self.trunk_units):
                if self.dp.owns(u):
                    self.t_master[u] = t.detach().float().clone()
                    self.t_m[u] = torch.zeros_like(self.t_master[u])
                    self.t_v[u] = torch.zeros_like(self.t_master[u])
        self.param: Dict[str, torch.Tensor] = {n: named[n] for n in self.names}
        # fp32 master weights + moments (DeepSpeed bf16 optimizer keeps the same)
        self.master = {n: p.detach().float().clone() for n, p in self.param.items()}
        self.m = {n: torch.zeros_like(t) for n, t in self.master.items()}
        self.v = {n: torch.zeros_like(t) for n, t in self.master.items()}
        self.grad = {n: torch.zeros_like(t) for n, t in self.master.items()}
        # the many small tensors of the projector and of the point encoder (150 with --unfreeze_pc_encoder) live in ONE flat fp32
        # buffer per group, so that a data-parallel step exchanges each group with a single all-reduce instead of one ~50 us
        # collective per tensor at the very end of the backward
        self._bucket_of: Dict[str, str] = {}
        self._buckets: Dict[str, torch.Tensor] = {}
        self._bucket_names: Dict[str, List[str]] = {}
        for key, prefix in (("bucket:point_proj", "model.point_proj."), ("bucket:point_backbone", "model.point_backbone.")):
            group = [n for n in self.names if n.startswith(prefix)]
            if len(group) < 2:
                continue
            sizes = [(self.master[n].numel() + 7) // 8 * 8 for n in group]          # 32-byte aligned views
            flat = torch.zeros(sum(sizes), device=self.master[group[0]].device, dtype=torch.float32)
            off = 0
            for n, sz in zip(group, sizes):
                self.grad[n] = flat[off:off + self.master[n].numel()].view(self.master[n].shape)
                self._bucket_of[n] = key
                off += sz
            self._buckets[key], self._bucket_names[key] = flat, group
        self.last_loss: Optional[torch.Tensor] = None
        self._sync = world_size > 1
        self._wgrad_fresh = False           # True right after zero_grad(): the next backward overwrites the trunk weight gradients
        self.sparse_embed_exchange = True   # (id, row) all-gather + deterministic scatter instead of the dense embed_tokens all-reduce
        self._exchange_rows = -1
        self._embed_pending = []

    def _sync_initial_state(self):
        """What ``deepspeed.initialize`` does at train.py:119-125: every rank starts from rank 0's parameters and buffers
        (ranks seeded differently would otherwise train diverging replicas silently).  Packed q|k|v / gate|up parameters are
        views of one buffer each, so the packed buffers are sent instead of their three / two views."""
        import torch.distributed as dist
        m = self.model
        src = dist.get_global_rank(self.pg, 0) if self.pg is not None else 0
        sent = []
        for t in list(m.wqkv) + list(m.wgu):
            sent.append(t)
        skip = (".self_attn.q_proj.", ".self_attn.k_proj.", ".self_attn.v_proj.", ".mlp.gate_proj.", ".mlp.up_proj.")
        for n, t in list(m.named_parameters()) + list(m.named_buffers()):
            if any(k in n for k in skip):
                continue
            sent.append(t.detach())
        hs = [dist.broadcast(t, src=src, group=self.pg, async_op=True) for t in sent]
        for h in hs:
            h.wait()
        m.repack()
        if self.enc_train is not None:
            self.enc_train["pk"].refresh()

    def zero_grad(self):
        self._embed_pending = []
        for n, g in self.grad.items():
            if n != "lm_head.weight":                      # overwritten by the first head backward after this call
                g.zero_()
        self._wgrad_fresh = True
        if self.wgrad is not None:
            # the four big units of a layer are OVERWRITTEN by the first backward after this call (wgrad_accumulate=False), so only
            # the norm-weight gradients (atomically accumulated) are cleared: no 27 GB memset + re-read per step
            for k in ("ln1", "ln2"):
                for g in self.wgrad[k]:
                    g.zero_()
            self._wgrad_fresh = True

    # ------------------------------------------------------------------ forward + backward
    def forward_backward(self, tokens: torch.Tensor, attention_masks: torch.Tensor, pcrgbs: torch.Tensor, prompt_len: int,
                         fps_start: Optional[torch.Tensor] = None, drop_u: Optional[torch.Tensor] = None,
                         sync: bool = True) -> torch.Tensor:
        """Forward + backward of one micro-batch; gradients ACCUMULATE into ``self.grad`` / ``self.wgrad``.  ``sync=False``
        (DDP's ``no_sync``): queue no gradient exchange - for every micro-batch but the last of a --grad_accum_steps group
        (train.py:183 lets DeepSpeed do the same), so already-summed gradients are never reduced twice."""
        m, cfg = self.model, self.model.cfg
        self._sync = bool(sync) and self.world > 1
        dev = m.dev
        P = int(prompt_len)
        h, ctx = forward_hidden_train(m, tokens, attention_masks, pcrgbs, fps_start, drop_u, out_rows=(P - 1, tokens.shape[1] - 1))
        tokens = ctx.tokens
        B, L = ctx.B, ctx.L
        d, V = cfg.hidden_size, cfg.vocab_size
        Tn = L - P
        # ---- head: rows P-1 .. L-2 only (train.py:174); the trunk may already have returned just those rows
        rows_only = h.shape[0] == B * Tn and Tn != L
        if rows_only:
            hs2 = h
        else:
            hs = torch.empty(B, Tn, d, device=dev, dtype=torch.bfloat16)
            ops.copy_rows(h.view(B, L, d)[:, P - 1:L - 1, :], hs)
            hs2 = hs.view(B * Tn, d)
        norm_w = m.model.norm.weight.detach()
        hn, rstd = ops.rmsnorm(hs2, norm_w, cfg.rms_norm_eps, want_rstd=True)
        lbuf = torch.empty(B * Tn, m.v_pad, device=dev, dtype=torch.bfloat16)
        logits = lbuf[:, :V]
        ops.gemm(hn, m.lm_head.weight.detach(), out=logits)
        targets = tokens[:, P:].contiguous().view(-1)
        loss = ops.cross_entropy(logits, V, targets, cfg.pad_token_id, write_grad=True)      # logits <- dlogits
        # ---- head backward
        ops.gemm_tn(logits, hn, out=self.grad["lm_head.weight"], accumulate=not self._wgrad_fresh)   # dW = dlogits^T . hn
        dhn = ops.gemm_nn(logits, m.lm_head.weight.detach()) if ops.gemm_nn_enabled else ops.gemm(logits, m.lm_head_T())
        dhs = ops.rmsnorm_bwd(dhn, hs2, norm_w, rstd, dw=self.grad["model.norm.weight"])
        self._reduce_async(["lm_head.weight", "model.norm.weight"])       # overlaps with the trunk backward below
        # NCCL's CTAs need SMs of their own to make progress next to the persistent 1-CTA-per-SM GEMMs: keep a few SMs out of the
        # GEMM grids while the head gradients (528 MB at full size) are in flight - the first layers of the trunk backward for the
        # default recipe, the whole backward when every layer's gradients are exchanged (--unfreeze_language_model)
        reserve = self.nccl_sm_reserve if self._sync else 0
        if reserve > 0:
            ops.set_gemm_sm_reserve(reserve)
        n_layers = cfg.num_hidden_layers
        release_at = -1 if self.wgrad is not None else n_layers - self.nccl_reserve_layers      # layers run top (n-1) to bottom (0)

        def layer_done(li):
            if self.wgrad is not None and self._sync:
                self._reduce_layer_async(li)
            if reserve > 0 and li == release_at:
                ops.set_gemm_sm_reserve(0)
        if rows_only:
            dh = dhs
        else:
            dh = torch.zeros(B, L, d, device=dev, dtype=torch.bfloat16)
            ops.copy_rows(dhs.view(B, Tn, d), dh[:, P - 1:L - 1, :])
            dh = dh.view(B * L, d)
        # ---- trunk backward (data gradients through the frozen layers), embedding scatter, projector, encoder
        backward_from_hidden(m, ctx, dh, self.grad, self.wgrad,
                             on_layer_done=layer_done if (reserve > 0 or (self.wgrad is not None and self._sync)) else None,
                             on_ready=self._reduce_async, wgrad_accumulate=not self._wgrad_fresh,
                             embed_exchange=self._embed_exchange if (self.world > 1 and self.sparse_embed_exchange) else None)
        if reserve > 0:
            ops.set_gemm_sm_reserve(0)
        self._wgrad_fresh = False
        self.last_loss = loss
        return loss

    # ------------------------------------------------------------------ data parallel (parallel.GradExchange)
    def _reduce_async(self, names: List[str]):
        if self._sync:
            names = list(names)
            whole = [k for k, group in self._bucket_names.items() if all(n in names for n in group)]
            for k in whole:
                self.dp.allreduce_async_named([(k, self._buckets[k])])
            covered = {n for k in whole for n in self._bucket_names[k]}
            self.dp.allreduce_async_named((n, self.grad[n]) for n in names if n not in covered)

    def _embed_exchange(self, tokens: torch.Tensor, dx: torch.Tensor, splice_pos, n_point: int):
        """embed_tokens gradient of a data-parallel step WITHOUT the dense [V, d] fp32 all-reduce (528 MB at full size, issued
        after the last backward kernel and therefore fully exposed): at most B * (L - n_point) rows are non-zero per rank, so
        the ranks all-gather (token id, bf16 gradient row) pairs - 12 MB per rank - and every rank adds ALL of them into its
        dense fp32 gradient with the deterministic scatter (esb_embed_scatter_rows): same summands in the same order on every
        rank, hence bit-identical replicas.  Micro-batches run with ``sync=False`` only stash their pairs; the exchange of the
        whole group happens on the synchronising micro-batch.  Assumes the collate_fn shapes (B, L) and the number of
        micro-batches are equal on all ranks (dataset.py:150-194 pads to fixed lengths; checked once)."""
        cfg = self.model.cfg
        B, L = tokens.shape
        all_spliced = splice_pos is not None and bool(self.model.last_splice_all)
        rmax = L - n_point if all_spliced else L
        self._embed_pending.append(ops.embed_rows_pack(tokens, dx, splice_pos, n_point, rmax, cfg.pad_token_id))
        if not self._sync:
            return
        rows = torch.cat([r for r, _ in self._embed_pending]) if len(self._embed_pending) > 1 else self._embed_pending[0][0]
        ids = torch.cat([i for _, i in self._embed_pending]) if len(self._embed_pending) > 1 else self._embed_pending[0][1]
        self._embed_pending = []
        if self._exchange_rows != rows.shape[0]:                    # first step / shape change: agree before a fixed-size gather
            n = torch.tensor([rows.shape[0]], device=dx.device, dtype=torch.int64)
            alln = self.dp.all_gather(n)
            if not bool((alln == n).all()):
                raise RuntimeError("sparse embedding-gradient exchange needs equal batch shapes and micro-batch counts on all "
                                   f"ranks, got row counts {alln.flatten().tolist()} (set Trainer.sparse_embed_exchange = False)")
            self._exchange_rows = rows.shape[0]
        # the gathers run behind the projector / encoder backward and the AdamW of the other tensors; the scatter waits for them
        # right before the embedding's own optimizer step (_finish_embed_exchange)
        self._embed_inflight = (self.dp.all_gather_async(rows), self.dp.all_gather_async(ids), rows.shape[1])

    def _finish_embed_exchange(self):
        if self._embed_inflight is None:
            return
        (all_rows, w_rows), (all_ids, w_ids), d = self._embed_inflight
        self._embed_inflight = None
        ev = self._mark("embed_exchange")
        for w in (w_rows, w_ids):
            if w is not None:
                w.wait()
        all_rows, all_ids = all_rows.view(-1, d), all_ids.view(-1)
        ops.embed_scatter_sorted(all_ids, all_rows, self.grad["model.embed_tokens.weight"], self.model.cfg.pad_token_id)
        self._mark_end(ev)

    def _reduce_layer_async(self, li: int):
        """ZeRO-1 gradient exchange of one LLM layer, started as soon as its backward has been queued: every unit is
        sum-reduced to its owner rank only, overlapping the backward of the remaining layers."""
        for u, (k, l2, _) in enumerate(self.trunk_units):
            if l2 == li:
                self.dp.reduce_to_owner_async(u, self.wgrad[k][li])

    def _wait_reduce(self):
        """All gradients of this step are final in ``self.grad`` / ``self.wgrad`` once the current stream reaches this point."""
        ev = self._mark("wait_reduce")
        self.dp.wait()
        self._finish_embed_exchange()
        self._mark_end(ev)

    _embed_inflight = None

    # optional CUDA-event spans around the exposed parts of the exchange (bench.py --gpus N: dp_timeline)
    trace_spans = None          # None: off; dict name -> list of (start event, end event)
    # SMs kept out of the persistent GEMM grids while gradients are being exchanged, and for how many layers of the trunk
    # backward after the head gradients were handed to NCCL (default recipe)
    nccl_sm_reserve = int(os.environ.get("ESB_NCCL_SM_RESERVE", "0"))       # measured at N = 2 and 8: 0 / 8 / 16 within noise -> off
    nccl_reserve_layers = int(os.environ.get("ESB_NCCL_RESERVE_LAYERS", "4"))

    def _mark(self, name):
        if self.trace_spans is None or self.world == 1:
            return None
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        self.trace_spans.setdefault(name, []).append((e0, e1))
        return e1

    @staticmethod
    def _mark_end(ev):
        if ev is not None:
            ev.record()

    # ------------------------------------------------------------------ optimizer
    def optimizer_step(self):
        lr = linear_warmup_lr(self.lr, self.step_count, self.warmup, self.total) if self.total > 0 else self.lr
        self.step_count += 1
        scale = 1.0 / self.world                                   # mean over ranks (DDP semantics)
        # tensors in the order their exchange was started (head first, embedding last): each AdamW launch waits only for its own
        # collective, so the tail of the exchange (projector all-reduce, embedding-row gathers) hides behind the earlier updates
        emb = "model.embed_tokens.weight"
        order = [n for n in self.names if n.startswith("lm_head.") or n == "model.norm.weight"]
        order += [n for n in self.names if n not in order and n != emb] + [n for n in self.names if n == emb]
        for n in order:
            p = self.param[n]
            for key in (n, self._bucket_of.get(n)):
                if key is not None and key in self.dp._named:
                    ev = self._mark("wait:" + key)
                    self.dp.wait_named(key)
                    self._mark_end(ev)
            if n == emb:
                self._finish_embed_exchange()
            ops.adamw_(self.master[n], self.grad[n], self.m[n], self.v[n], p.detach() if p.dtype == torch.bfloat16 else None,
                       lr, self.betas[0], self.betas[1], self.eps, self.wd, self.step_count, scale)
            if p.dtype != torch.bfloat16:
                p.detach().copy_(self.master[n])
        self._wait_reduce()                                         # LLM-layer units (ZeRO-1 reductions), anything left
        if self.trunk_units:
            self._trunk_step(lr, scale)
        # refresh derived operands of the trainable tensors (the fused AdamW kernel writes the parameters through raw
        # pointers, so autograd's version counters - what model.sync_derived watches - do not move)
        self.model._lm_head_T = None
        self.model._lm_head_tiled = None
        self.model._packed_proj.refresh()
        if self.enc_train is not None:
            self.enc_train["pk"].refresh()
            self.model.repack_encoder()

    def _trunk_step(self, lr: float, scale: float):
        """AdamW on the LLM-layer units this rank owns, then (N > 1) broadcast of the updated bf16 weights from their
        owners (the all-gather half of ZeRO-1) and refresh of the contraction-major copies."""
        for u, (k, li, t) in enumerate(self.trunk_units):
            if u in self.t_master:
                ops.adamw_(self.t_master[u].view(-1), self.wgrad[k][li].view(-1), self.t_m[u].view(-1), self.t_v[u].view(-1),
                           t.view(-1) if t.is_contiguous() else None, lr, self.betas[0], self.betas[1], self.eps, self.wd,
                           self.step_count, scale)
                if not t.is_contiguous():
                    t.copy_(self.t_master[u])
        self.dp.broadcast_from_owners([(u, t) for u, (k, li, t) in enumerate(self.trunk_units)])
        for lw in self.model.trunk.layers:
            lw.refresh_derived()

    def train_step(self, batch: dict, fps_start: Optional[torch.Tensor] = None, drop_u: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One iteration of train.py:157-184 on a collate_fn-shaped batch (dataset.py:184-194)."""
        ev = self._mark("zero_grad")
        self.zero_grad()
        self._mark_end(ev)
        ev = self._mark("forward_backward")
        loss = self.forward_backward(batch["tokens"], batch["attention_masks"], batch["pcrgbs"], batch["prompts"].shape[1],
                                     fps_start, drop_u)
        self._mark_end(ev)
        ev = self._mark("optimizer_step")
        self.optimizer_step()
        self._mark_end(ev)
        return loss
