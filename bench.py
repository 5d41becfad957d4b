#!/usr/bin/env python
"""Benchmark of the trajectory-VLM training step (BASELINE.json configs[1]: full-size traj-VLM bf16 training step,
LLaMA-7B dims + PointBERT v1.2, per-GPU batch 8, L=696, P=544, default freezing) on N B200s of one node.

    python bench.py --gpus N --steps K --warmup W            # our CUDA path (one JSON line on rank 0)
    python bench.py --impl reference --gpus N --steps K ...  # the CPU oracle port on the host cores (rank 0 only)

A "step" = forward + backward + gradient all-reduce (N>1) + AdamW on one synthetic collate_fn-shaped batch.
`value` = samples/s with the batch already resident in HBM; `e2e` = the same step fed from pinned HOST buffers
(H2D of tokens / masks / clouds inside the timed region) with the loss read back (D2H) every step.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time
import time as time_mod

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "train samples/s"
UNIT = "samples/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="full", choices=["full", "tiny"])
    ap.add_argument("--batch", type=int, default=8, help="per-GPU batch (train.py --bs 8)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-baseline", action="store_true", help="skip the stock eager-PyTorch-on-this-GPU baseline")
    ap.add_argument("--no-traj-tokens", action="store_true", help="skip the quantize/dequantize kernel measurement")
    ap.add_argument("--no-decode", action="store_true", help="skip the batched greedy-decode measurement")
    ap.add_argument("--decode-only", action="store_true", help="profiling aid: build the model and run only the decode measurement")
    ap.add_argument("--decode-batch", type=int, default=32, help="sequences per GPU (BASELINE configs[3]: 256 / 8 GPUs)")
    ap.add_argument("--cpu-budget-s", type=float, default=200.0)
    ap.add_argument("--encoder", default="v12", choices=["v12", "base"],
                    help="point encoder yaml: v12 = PointTransformer_8192point_2layer (default), base = PointTransformer_base_8192point "
                         "(BASELINE configs[4], SURVEY 8d config E)")
    ap.add_argument("--verify-dp", action="store_true",
                    help="N>1: only run the numeric check of the gradient exchange (dp_check) and print it")
    ap.add_argument("--dp-check-batch", type=int, default=1, help="samples per rank in the dp_check")
    ap.add_argument("--no-dp-check", action="store_true")
    ap.add_argument("--recipe", default="default", choices=["default", "unfreeze_lm", "unfreeze_pc"],
                    help="freezing recipe of model_arch.py:33-51: default, --unfreeze_language_model, --unfreeze_pc_encoder")
    return ap.parse_args()


def workload_name(cfgname, B, encoder="v12", recipe="default"):
    if cfgname == "full":
        enc = "PointBERT v1.2" if encoder == "v12" else "PointBERT base (1152-d, k=48)"
        rec = {"default": "default freezing", "unfreeze_lm": "--unfreeze_language_model", "unfreeze_pc": "--unfreeze_pc_encoder"}[recipe]
        return f"full-size traj-VLM bf16 training step (LLaMA-7B dims, {enc}, B={B}/GPU, L=696, P=544, {rec})"
    return f"tiny traj-VLM training step (d=256, 2 layers, B={B}/GPU, L=696)"


def config_dict(args, B, world, L=696, P=544):
    """The `config` object of the JSON line - identical for our arm and the reference arm (the driver compares them)."""
    return {"workload": workload_name(args.config, B, args.encoder, args.recipe), "per_gpu_batch": B, "global_batch": B * world,
            "seq_len": L, "prompt_len": P, "parallelism": f"dp{world}", "l2": "working set (13.5 GB of weights per pass) >> 126 MB L2"}


def flops_per_sample_train(cfg, L=696, T=152, recipe="default"):
    """forward + data-gradients through the trunk/attention, head on T rows fwd+dgrad+wgrad, projector
    fwd+dgrad+wgrad, encoder forward only."""
    d, f, n, V = cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers, cfg.vocab_size
    lin = 2 * L * n * (4 * d * d + 3 * d * f)
    att = 2 * L * L * d * n
    head = 2 * T * d * V
    p = cfg.point
    Np = p.point_token_len
    dims = [p.trans_dim] + list(p.projection_hidden_dim) + [d]
    proj = sum(2 * Np * a * b for a, b in zip(dims[:-1], dims[1:]))
    enc = 2 * Np * p.depth * 12 * p.trans_dim ** 2 + 4 * Np * Np * p.trans_dim * p.depth \
        + 2 * p.num_group * p.group_size * (p.point_dims * 128 + 128 * 256 + 512 * 512 + 512 * p.encoder_dims)
    if recipe == "unfreeze_lm":                   # + weight gradients of the trunk (attention backward counted 2x forward)
        return 3 * (lin + att) + 3 * (head + proj) + enc
    if recipe == "unfreeze_pc":                   # + data and weight gradients of the encoder
        return 2 * (lin + att) + 3 * (head + proj) + 3 * enc
    return 2 * (lin + att) + 3 * (head + proj) + enc


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz, self._halt = index, [], set(), None, threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap", nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake"}
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, nm in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.1)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


def gemm_traffic():
    """dram read+write bytes of the largest GEMM launch from the committed `ncu --set full` capture (profiles/)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "gemm_traffic.json")))
    except Exception:
        return None


class GemmTimer:
    """CUDA-event pairs around every GEMM launch of the timed region (same stream as the launches)."""

    def __init__(self, torch):
        self.torch, self.pairs, self.flops, self.n = torch, [], 0.0, 0
        self.pool = []

    def before(self, M, N, K):
        t = self.torch
        a = t.cuda.Event(enable_timing=True)
        b = t.cuda.Event(enable_timing=True)
        a.record()
        self.pairs.append((a, b))
        self.flops += 2.0 * M * N * K
        self.n += 1

    def after(self):
        self.pairs[-1][1].record()

    def total_ms(self):
        return sum(a.elapsed_time(b) for a, b in self.pairs)


# ------------------------------------------------------------------------------------------------ our arm
def build_model(torch, cfg, device, **model_kw):
    """Random-init weights of the named architecture (no checkpoint is available offline): N(0, 0.02) for the LLM /
    embeddings / head (HF initializer_range), fan-in scaled for the point encoder and projector, identity norms."""
    from egoscaler_b200.model import TrajPointLLMForCausalLM
    from egoscaler_b200 import decode  # noqa: F401
    m = TrajPointLLMForCausalLM(cfg, device, **model_kw)
    g = torch.Generator(device=device).manual_seed(0)
    with torch.no_grad():
        for name, p in list(m.named_parameters()) + list(m.named_buffers()):
            leaf = name.rsplit(".", 1)[-1]
            if leaf == "num_batches_tracked":
                continue
            if leaf == "running_var" or ("norm" in name and leaf == "weight") or (".1.weight" in name and "conv" in name):
                p.fill_(1.0)
            elif leaf in ("running_mean", "bias"):
                p.zero_()
            elif p.dim() >= 2 and ("point_backbone" in name or "point_proj" in name):
                p.normal_(0.0, 1.0 / (p.shape[1] ** 0.5), generator=g)
            else:
                p.normal_(0.0, 0.02, generator=g)
    m.repack()
    return m


def host_batches(torch, cfg, B, n, seed0):
    from egoscaler_b200 import synth
    out = []
    for i in range(n):
        bt = synth.make_batch(cfg, B, seed=seed0 + i, random_desc_len=True)
        out.append({k: v.pin_memory() for k, v in bt.items()})
    return out


def run_b200(args):
    import torch
    import torch.distributed as dist
    from egoscaler_b200 import lib as esb_lib, ops
    from egoscaler_b200.config import TrajVLMConfig
    from egoscaler_b200.trainer import Trainer

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.config == "full":
        cfg = TrajVLMConfig.full() if args.encoder == "v12" else TrajVLMConfig.full_base()
    else:
        cfg = TrajVLMConfig.tiny() if args.encoder == "v12" else TrajVLMConfig.tiny_base()
    B = args.batch
    model = build_model(torch, cfg, dev, unfreeze_language_model=args.recipe == "unfreeze_lm",
                        unfreeze_pc_encoder=args.recipe == "unfreeze_pc")
    if args.decode_only:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        d = decode_bench(torch, dist, model, cfg, args, dev, world, rank, peaks)
        if rank == 0:
            print(json.dumps(d), flush=True)
        return
    model.train()
    tr = Trainer(model, lr=2e-5, world_size=world)
    dpc = None
    if world > 1 and (args.verify_dp or not args.no_dp_check):
        dpc = dp_check(torch, dist, tr, cfg, args, dev, world, rank)
        if args.verify_dp:
            if rank == 0:
                print(json.dumps({"dp_check": dpc, "n_gpus": world}), flush=True)
            dist.destroy_process_group()
            return
    nb = 4
    hb = host_batches(torch, cfg, B, nb, seed0=1000 * (rank + 1))          # different data per rank, same weights
    P = hb[0]["prompts"].shape[1]
    L = hb[0]["tokens"].shape[1]
    dbatches = [{k: v.to(dev) for k, v in b.items()} for b in hb]
    n_pts = hb[0]["pcrgbs"].shape[1]
    starts = [torch.randint(0, n_pts, (B,), generator=torch.Generator().manual_seed(7 + i)).pin_memory() for i in range(nb)]
    dstarts = [s.to(dev) for s in starts]

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def step_resident(i):
        return tr.train_step(dbatches[i % nb], dstarts[i % nb])

    stage = {k: torch.empty_like(v, device=dev) for k, v in hb[0].items()}
    stage_start = torch.empty(B, device=dev, dtype=torch.long)

    def step_e2e(i):
        b = hb[i % nb]
        for k in ("tokens", "attention_masks", "pcrgbs", "prompts"):
            stage[k].copy_(b[k], non_blocking=True)
        stage_start.copy_(starts[i % nb], non_blocking=True)
        loss = tr.train_step(stage, stage_start)
        return float(loss.cpu())                                           # D2H read of the loss, every step

    def timed(fn, steps):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        last = None
        for i in range(steps):
            last = fn(i)
        e1.record()
        sync_all()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms), last

    for i in range(max(args.warmup, 3)):
        step_resident(i)
    sampler = ClockSampler(local)
    sampler.start()
    l0 = esb_lib.launch_count
    ms_res, loss_res = timed(step_resident, args.steps)                    # headline: no per-kernel events in this pass
    launches = (esb_lib.launch_count - l0) // args.steps
    for i in range(2):
        step_e2e(i)
    ms_e2e, loss_e2e = timed(step_e2e, args.steps)
    # roofline pass: the same resident step with a CUDA-event pair around every GEMM launch (644 events per step cost
    # ~1 % of host time, so this pass is NOT the headline)
    timer = GemmTimer(torch)
    ops.gemm_timer = timer
    gsteps = max(2, min(args.steps, 8))
    ms_gpass, _ = timed(step_resident, gsteps)
    ops.gemm_timer = None
    torch.cuda.synchronize()
    gemm_ms = timer.total_ms() / gsteps * args.steps                       # scaled to K steps: formulas below stay per-K
    timer.flops = timer.flops / gsteps * args.steps
    timer.n = timer.n // gsteps * args.steps
    clocks = sampler.stop()
    # N > 1: the same step WITHOUT any gradient exchange, timed on every rank by itself - separates a slow GPU of the box
    # (all ranks wait for it in every collective) from exchange time that the overlap does not hide
    dp_timeline = None
    spans = None
    if world > 1:
        tr.trace_spans = {}
        sync_all()
        for i in range(4):
            step_resident(i)
        sync_all()
        spans = {k: round(sum(a.elapsed_time(b) for a, b in v) / 4, 3) for k, v in tr.trace_spans.items()}
        tr.trace_spans = None
    if world > 1 and args.recipe != "unfreeze_lm":
        def step_local(i):
            b = dbatches[i % nb]
            ev = tr._mark("zero_grad")
            tr.zero_grad()
            tr._mark_end(ev)
            ev = tr._mark("forward_backward")
            loss = tr.forward_backward(b["tokens"], b["attention_masks"], b["pcrgbs"], b["prompts"].shape[1], dstarts[i % nb], None,
                                       sync=False)
            tr._mark_end(ev)
            tr._embed_pending = []
            ev = tr._mark("optimizer_step")
            tr.optimizer_step()
            tr._mark_end(ev)
            return loss
        step_local(0)
        tr.trace_spans = {}
        sync_all()
        for i in range(4):
            step_local(i)
        sync_all()
        spans_local = {k: round(sum(a.elapsed_time(b) for a, b in v) / 4, 3) for k, v in tr.trace_spans.items()}
        tr.trace_spans = None
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(4):
            step_local(i)
        e1.record()
        torch.cuda.synchronize()
        mine = torch.tensor([e0.elapsed_time(e1) / 4], device=dev)
        allms = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allms, mine)
        per_rank = [round(float(t), 2) for t in allms]
        dp_timeline = {"local_step_ms_per_rank_no_exchange": per_rank, "slowest_rank_ms": max(per_rank), "fastest_rank_ms": min(per_rank),
                       "step_ms_with_exchange": round(ms_res / args.steps, 2),
                       "exchange_not_hidden_ms": round(ms_res / args.steps - max(per_rank), 2),
                       "rank0_ms_per_step_inside": spans, "rank0_ms_per_step_inside_no_exchange": spans_local}

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak_tf = float(peaks.get("bf16_tflops_sustained", 1400.0))
    peak_src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)" if "bf16_tflops_sustained" in peaks else "fallback 1.4 PFLOP/s sustained"
    achieved_tf = timer.flops / (gemm_ms * 1e-3) / 1e12 if gemm_ms > 0 else 0.0
    samples = B * world * args.steps
    value = samples / (ms_res * 1e-3)
    e2e = samples / (ms_e2e * 1e-3)
    h2d = sum(hb[0][k].numel() * hb[0][k].element_size() for k in ("tokens", "attention_masks", "pcrgbs", "prompts")) + B * 8
    line = {
        "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": round(ms_res / args.steps, 3), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic (collate_fn-shaped batches, random-init weights)",
        "config": config_dict(args, B, world, L, P),
        "e2e": {"value": round(e2e, 3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": 4,
                "ms_per_step": round(ms_e2e / args.steps, 3)},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": {"bound": "tensor", "kernel": "gemm2_bf16_tcgen05_kernel (CTA pairs, large shapes) + gemm_bf16_tcgen05_kernel", "achieved": round(achieved_tf, 1), "peak": peak_tf,
                     "unit": "TFLOP/s", "frac": round(achieved_tf / peak_tf, 4), "traffic": gemm_traffic(), "peak_source": peak_src,
                     "traffic_source": "committed ncu --set full capture (profiles/gemm_traffic.json), not measured in this run",
                     "timing": "CUDA events around every GEMM launch in a separate pass of the same resident step "
                               f"({gsteps} steps, {round(ms_gpass / gsteps, 3)} ms/step with events)",
                     "launches_per_step": timer.n // args.steps, "gemm_ms_per_step": round(gemm_ms / args.steps, 3),
                     "gemm_share_of_step": round(gemm_ms / args.steps / (ms_gpass / gsteps), 4),
                     "step_model_tflops": round(flops_per_sample_train(cfg, L, L - P, args.recipe) * B / (ms_res / args.steps * 1e-3) / 1e12, 1)},
        "loss": round(float(loss_res), 4),
    }
    if dpc is not None:
        line["dp_check"] = dpc
    if dp_timeline is not None:
        line["dp_timeline"] = dp_timeline
    tr = None                                                          # frees optimizer state + gradient buffers
    torch.cuda.empty_cache()
    if not args.no_decode:
        line["decode"] = decode_bench(torch, dist, model, cfg, args, dev, world, rank, peaks)
    if not args.no_traj_tokens and args.recipe == "default":
        try:
            line["traj_tokens"] = traj_tokens_bench(torch, dev, peaks, with_cpu=(rank == 0 and world == 1))
        except Exception as e:
            line["traj_tokens"] = {"error": repr(e)[:300]}
    if rank == 0 and world == 1 and not args.no_gpu_baseline and args.recipe == "default":
        try:
            model = None
            torch.cuda.empty_cache()
            line["gpu_baseline"] = gpu_eager_baseline(torch, cfg, args, dev, B)
            line["gpu_baseline"]["speedup_of_value"] = round(value / line["gpu_baseline"]["value"], 3)
        except Exception as e:
            line["gpu_baseline"] = {"value": None, "error": repr(e)[:300]}
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_reference_step(args, budget_s=min(args.cpu_budget_s, 45.0), max_steps=2, warmup=1)
                if line["cpu_baseline"].get("value"):
                    line["vs_cpu_port"] = round(e2e / line["cpu_baseline"]["value"], 1)      # e2e arm / CPU port (vs_baseline stays null: BASELINE.md publishes no number)
            except Exception as e:  # the baseline is informative; the GPU numbers stand without it
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "error": repr(e)[:200]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def dp_check(torch, dist, tr, cfg, args, dev, world, rank):
    """SURVEY 8d config C correctness: "post-all-reduce grads == 1-GPU grads on the concatenated batch" (what DeepSpeed's
    ZeRO-1 engine guarantees at train.py:92-125,183-184).  Every rank takes its shard of ONE global batch, runs
    forward/backward with the real NCCL exchange, then recomputes the whole global batch locally without any exchange and
    compares: default recipe - every all-reduced gradient; unfreeze_lm - each LLM-layer unit on its ZeRO-1 OWNER rank
    (reduce-to-owner); unfreeze_pc - the local recomputation runs shard by shard, because the train-mode BatchNorm of the
    point encoder uses per-rank batch statistics exactly like the reference (no SyncBN, model_arch.py:121-122).  Then one
    optimizer step is taken and the updated bf16 weights must be bit-identical on all ranks (broadcast-from-owner)."""
    from egoscaler_b200 import synth
    b = args.dp_check_batch
    gb = synth.make_batch(cfg, b * world, seed=4242, random_desc_len=True)
    gb = {k: v.to(dev) for k, v in gb.items()}
    P = gb["prompts"].shape[1]
    starts = torch.randint(0, gb["pcrgbs"].shape[1], (b * world,), generator=torch.Generator().manual_seed(99)).to(dev)
    drop_u = None
    if tr.enc_train is not None:
        depth = len(tr.enc_train["pk"].blocks)
        drop_u = torch.rand(depth, 2, b * world, generator=torch.Generator().manual_seed(98)).to(dev)
    ntok = (gb["tokens"][:, P:] != cfg.pad_token_id).sum(1)
    equal_counts = bool((ntok == ntok[0]).all())                         # mean-of-means == global mean only then

    def fb(lo, hi, sync):
        return tr.forward_backward(gb["tokens"][lo:hi], gb["attention_masks"][lo:hi], gb["pcrgbs"][lo:hi], P, starts[lo:hi],
                                   None if drop_u is None else drop_u[:, :, lo:hi].contiguous(), sync=sync)

    lo, hi = rank * b, (rank + 1) * b
    tr.zero_grad()
    loss_shard = fb(lo, hi, True)
    tr._wait_reduce()
    torch.cuda.synchronize()
    got = {n: tr.grad[n].clone() for n in tr.names}
    for u, (k, li, _) in enumerate(tr.trunk_units):
        if tr.dp.owns(u):
            got[f"unit{u}:{k}.{li}"] = tr.wgrad[k][li].clone()
    tr.zero_grad()
    sparse = tr.sparse_embed_exchange
    tr.sparse_embed_exchange = False                                      # local recomputation: plain local scatter, no exchange
    if tr.enc_train is not None:
        losses = [fb(r * b, (r + 1) * b, False) for r in range(world)]
        loss_ref = sum(float(x) for x in losses) / world
        ref_scale = 1.0                                                   # sum over shards of per-shard means, like the exchange
    else:
        loss_ref = float(fb(0, b * world, False))
        ref_scale = float(world)                                          # global mean * world == sum of equal-count shard means
    tr.sparse_embed_exchange = sparse
    worst, worst_name, n_cmp = 0.0, "", 0
    rels = []
    # conv biases directly in front of a train-mode BatchNorm have a structurally zero gradient (the batch mean removes them):
    # both sides hold round-off only, so a relative comparison is meaningless (tests/test_variants_gpu.py bounds their size)
    skipped = [n for n in got if tr.enc_train is not None and n.endswith(("encoder.first_conv.0.bias", "encoder.second_conv.0.bias"))]
    for name, g in got.items():
        if name in skipped:
            continue
        if name.startswith("unit"):
            u = int(name[4:name.index(":")])
            k, li, _ = tr.trunk_units[u]
            ref = tr.wgrad[k][li]
        else:
            ref = tr.grad[name]
        ref = ref.double() * ref_scale
        den = float(ref.norm())
        if name.endswith(".bias") and name[:-5] + ".weight" in tr.grad:
            # a bias in front of a train-mode BatchNorm has a structurally zero gradient (round-off only): measure it against
            # the scale of its module's weight gradient instead of against itself
            den = max(den, 5e-3 * float(tr.grad[name[:-5] + ".weight"].double().norm()) * ref_scale)
        if den == 0.0:
            continue
        rel = float((g.double() - ref).norm()) / den
        n_cmp += 1
        rels.append((round(rel, 6), name))
        if rel > worst:
            worst, worst_name = rel, name
    del got
    # one real step: weights must stay bit-identical across ranks
    tr.zero_grad()
    fb(lo, hi, True)
    tr.optimizer_step()
    sums = []
    bits = lambda t: t.detach().contiguous().view(torch.int16 if t.element_size() == 2 else torch.int32).long().sum().double()
    for n in tr.names:
        sums.append(bits(tr.param[n]))
    for _, _, t in tr.trunk_units:
        sums.append(bits(t))
    chk = torch.stack(sums)
    mx, mn = chk.clone(), chk.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    dist.all_reduce(mn, op=dist.ReduceOp.MIN)
    identical = bool((mx == mn).all())
    w = torch.tensor([worst, float(loss_shard)], device=dev, dtype=torch.float64)
    wmax = w.clone()
    dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
    wsum = w.clone()
    dist.all_reduce(wsum, op=dist.ReduceOp.SUM)
    shard_losses = [None] * world
    dist.all_gather_object(shard_losses, round(float(loss_shard), 5))
    names = [None] * world
    dist.all_gather_object(names, (worst, worst_name))
    worst_all = max(names, key=lambda t: t[0])
    return {"ranks": world, "per_rank_batch": b, "recipe": args.recipe, "encoder": args.encoder,
            "max_rel_l2": round(float(wmax[0]), 6), "worst_tensor": worst_all[1], "tensors_compared_rank0": n_cmp,
            "tolerance": 2e-2, "ok": bool(float(wmax[0]) <= 2e-2 and identical),
            "loss_mean_over_ranks": round(float(wsum[1]) / world, 5), "loss_single_gpu_global_batch": round(loss_ref, 5),
            "loss_per_rank": shard_losses,
            "loss_recomputed_per_shard_rank0": [round(float(x), 5) for x in losses] if tr.enc_train is not None else None,
            "worst5_rank0": sorted(rels, reverse=True)[:5],
            "equal_token_counts": equal_counts, "weights_bit_identical_after_step": identical, "skipped_structurally_zero": skipped,
            "reference": ("per-shard recomputation on one GPU (per-rank BatchNorm statistics, no SyncBN)" if tr.enc_train is not None
                          else "one GPU on the concatenated global batch")}


def traj_tokens_bench(torch, dev, peaks, with_cpu=True):
    """north_star's trajectory-token kernel pair on SURVEY 8d's workload: 2^24 float32 values uniform in [-1.05, 1.05] with every
    exact bin edge and NaN / +-inf specials mixed in.  quantize = discretize_action (utils/utils.py:13-16): 4 B in + 8 B (int64
    id) out = 12 B/value; dequantize = token_to_action + float32 cast (:18-21,:96): 8 B in + 4 B out = 12 B/value.  Three rotating
    buffer sets (600 MB) so no launch finds its data in the 126 MB L2.  CPU baseline: the NumPy oracle round trip on the same
    values on the host."""
    import numpy as np
    from egoscaler_b200 import ops
    n = 1 << 24
    g = torch.Generator(device=dev).manual_seed(5)
    nset = 3
    xs = []
    edges = torch.linspace(-1, 1, 256, dtype=torch.float64).float().to(dev)
    special = torch.tensor([float("nan"), float("inf"), float("-inf"), -1.0, 1.0, 0.0], device=dev)
    for _ in range(nset):
        x = torch.rand(n, device=dev, generator=g) * 2.1 - 1.05
        x[:256] = edges
        x[256:262] = special
        xs.append(x[torch.randperm(n, device=dev, generator=g)].contiguous())
    n6 = n // 6 * 6                                                        # de-quantisation works on whole 6-DoF poses
    ids = [ops.traj_quantize(x, 256, 32006) for x in xs]
    outs = [ops.traj_dequantize(i[:n6].view(-1, 6), 256, 32006) for i in ids]             # warm-up + outputs for the check
    reps = 30

    def time(fn):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for r in range(reps):
            fn(r % nset)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    obuf = [torch.empty_like(i) for i in ids]
    fbuf = [torch.empty(n // 6, 6, device=dev) for _ in ids]
    lq = lambda k: ops._l.call("esb_traj_quantize", xs[k].data_ptr(), 0, n, 256, 32006, obuf[k].data_ptr(), ops._stream())
    ld = lambda k: ops._l.call("esb_traj_dequantize", ids[k].data_ptr(), n // 6, 256, 32006, 0, 0.0, 0.0, fbuf[k].data_ptr(), ops._stream())
    lq(0); ld(0)
    ms_q, ms_d = time(lq), time(ld)
    hbm = float(peaks.get("hbm_gbs", 6400.0))
    gq = 12.0 * n / (ms_q * 1e-3) / 1e9
    gd = 12.0 * n6 / (ms_d * 1e-3) / 1e9
    out = {"values": n, "bytes_per_value": 12,
           "quantize": {"us": round(ms_q * 1e3, 2), "achieved": round(gq, 1), "peak": hbm, "unit": "GB/s", "frac": round(gq / hbm, 4)},
           "dequantize": {"us": round(ms_d * 1e3, 2), "achieved": round(gd, 1), "peak": hbm, "unit": "GB/s", "frac": round(gd / hbm, 4)},
           "l2": "3 rotating buffer sets of 201 MB each > 126 MB L2"}
    if with_cpu:
        from oracle import traj_tokenizer as tt
        xh = xs[0].cpu().numpy()
        t0 = time_mod.time()
        q = np.asarray(tt.discretize_action(xh, 256))
        t1 = time_mod.time()
        dq = np.asarray(tt.token_to_action(q, 256)).astype(np.float32)
        t2 = time_mod.time()
        ok = bool(np.array_equal(q + 32006, ids[0].cpu().numpy()))
        valid = (q >= 0)
        ok_d = bool(np.array_equal(dq.reshape(-1)[:n6][valid[:n6]], outs[0].cpu().numpy().reshape(-1)[valid[:n6]]))
        out["cpu_baseline"] = {"kind": "port", "cores": 1, "sample": "the same 2^24 values, NumPy np.digitize / table lookup (oracle/traj_tokenizer.py)",
                               "quantize_gbs": round(12.0 * n / (t1 - t0) / 1e9, 3), "dequantize_gbs": round(12.0 * n / (t2 - t1) / 1e9, 3)}
        out["bit_exact_vs_cpu"] = ok and ok_d
    return out


def gpu_eager_baseline(torch, cfg, args, dev, B, steps=5, warmup=3):
    """BASELINE.md 3b / SURVEY 8d config B "reference timing": the same module in stock eager PyTorch on the SAME GPU - the oracle
    restatement (plain nn.Linear / nn.Embedding / F.cross_entropy) moved to cuda, bf16 weights for the frozen parts like the
    DeepSpeed bf16 engine (train.py:97-98), autocast(bf16) (train.py:166), attention through F.scaled_dot_product_attention with
    the 4-D padding+causal mask HF builds, torch.optim.AdamW (train.py:113).  cuBLASLt + SDPA + ~3k eager launches: what
    train.py:166-184 runs on a GPU today.  Informative only; never on the product path."""
    import torch.nn.functional as F
    from egoscaler_b200 import synth
    from oracle import traj_vlm as ov

    def sdpa_forward(self, x, cos, sin, mask, cache):
        Bq, Lq, _ = x.shape
        q = self.q_proj(x).view(Bq, Lq, self.heads, self.hd).transpose(1, 2)
        k = self.k_proj(x).view(Bq, Lq, self.heads, self.hd).transpose(1, 2)
        v = self.v_proj(x).view(Bq, Lq, self.heads, self.hd).transpose(1, 2)
        cos, sin = cos.to(q.dtype), sin.to(q.dtype)
        q = q * cos + ov._rot_half(q) * sin
        k = k * cos + ov._rot_half(k) * sin
        o = F.scaled_dot_product_attention(q, k, v, attn_mask=mask.expand(Bq, 1, Lq, Lq))
        return self.o_proj(o.transpose(1, 2).reshape(Bq, Lq, -1))

    llm = ov.LlmCfg.full() if args.config == "full" else ov.LlmCfg.tiny()
    pc = ov.PointCfg() if args.encoder == "v12" else ov.PointCfg.base()
    old_fwd = ov._SelfAttn.forward
    ov._SelfAttn.forward = sdpa_forward
    try:
        with torch.device("meta"):
            m = ov.TrajVLMOracle(llm, pc)
        m = m.to_empty(device=dev)
        gen = torch.Generator(device=dev).manual_seed(0)
        with torch.no_grad():
            for n_, p in list(m.named_parameters()) + list(m.named_buffers()):
                if not p.dtype.is_floating_point:
                    p.zero_()
                elif n_.endswith("running_var") or (p.dim() == 1 and not n_.endswith("bias") and not n_.endswith("running_mean")):
                    p.fill_(1.0)
                elif p.dim() == 1:
                    p.zero_()
                else:
                    p.normal_(0.0, 0.02, generator=gen)
        m.configure_trainable()
        m.to(torch.bfloat16)                                            # DeepSpeed bf16 engine: every parameter in bf16 (train.py:97-98)
        m.train()
        m.model.layers.eval()
        m.model.point_backbone.eval()
        opt = torch.optim.AdamW([q for q in m.parameters() if q.requires_grad], lr=2e-5)
        bts = [{k: v.to(dev) for k, v in synth.make_batch(cfg, B, seed=500 + i, random_desc_len=True).items()} for i in range(2)]
        P = bts[0]["prompts"].shape[1]
        start = torch.zeros(B, dtype=torch.long, device=dev)

        def step(i):
            bt = bts[i % 2]
            opt.zero_grad(set_to_none=True)
            with torch.autocast("cuda", dtype=torch.bfloat16):
                logits = m(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=start, training=True)
                loss = ov.train_loss(logits, bt["tokens"], P, cfg.pad_token_id)
            loss.backward()
            opt.step()
            return loss

        for i in range(warmup):
            step(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(steps):
            loss = step(i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        res = {"value": round(B / (ms * 1e-3), 3), "unit": UNIT, "ms_per_step": round(ms, 2), "steps": steps, "warmup": warmup,
               "per_gpu_batch": B, "loss": round(float(loss.detach()), 4),
               "what": "stock eager PyTorch on this GPU: oracle module on cuda, bf16 trunk, autocast(bf16), SDPA with the HF 4-D mask, "
                       f"torch.optim.AdamW; torch {torch.__version__}"}
    finally:
        ov._SelfAttn.forward = old_fwd
    del m, opt
    torch.cuda.empty_cache()
    return res


def decode_bench(torch, dist, model, cfg, args, dev, world, rank, peaks):
    """BASELINE configs[3]: batched greedy trajectory generation with the static KV cache, batch-sharded (no
    collective): `--decode-batch` prompts per GPU of P=544 tokens (point block + description + first pose), T=152 new
    tokens, EOS suppressed so every sequence runs full length.  traj/s = sequences / wall of
    generate() (prefill incl. point encoder + T cached steps), timed with CUDA events, max over ranks."""
    from egoscaler_b200 import synth
    Bd = args.decode_batch
    bt = synth.make_batch(cfg, Bd, seed=77 + rank)
    prompts, pmask, clouds = bt["prompts"].to(dev), bt["prompt_masks"].to(dev), bt["pcrgbs"].to(dev)
    P = prompts.shape[1]
    T = bt["tokens"].shape[1] - P
    start = torch.randint(0, clouds.shape[1], (Bd,), generator=torch.Generator().manual_seed(3)).to(dev)
    model.eval()

    def run(t_new):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = model.generate(input_ids=prompts, attention_mask=pmask, point_clouds=clouds, max_length=t_new, do_sample=False,
                             fps_start=start, suppress_eos=True)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms), out

    run(4)                                                                  # warm-up (allocator, lazy operands)
    ms1, _ = run(1)
    ms_full = min(run(T)[0] for _ in range(5))               # ~1 s each; the pool shows multi-second slow episodes
    d, f, n, V = cfg.hidden_size, cfg.intermediate_size, cfg.num_hidden_layers, cfg.vocab_size
    per_tok_ms = (ms_full - ms1) / max(1, T - 1)
    ctx_mean = P + T / 2.0
    step_bytes = 2.0 * (n * (4 * d * d + 3 * d * f) + d * V) + Bd * 2 * n * d * 2 * ctx_mean     # weights + mean KV
    hbm = float(peaks.get("hbm_gbs", 6400.0))
    ach = step_bytes / (per_tok_ms * 1e-3) / 1e9 if per_tok_ms > 0 else 0.0
    return {"metric": "traj decode/s", "value": round(Bd * world / (ms_full * 1e-3), 3), "unit": "traj/s",
            "tok_per_s": round(Bd * world * T / (ms_full * 1e-3), 1), "batch_per_gpu": Bd, "global_batch": Bd * world,
            "prompt_len": P, "new_tokens": T, "mode": "greedy, KV cache, EOS suppressed", "ms_total": round(ms_full, 2),
            "ms_prefill_plus_first_token": round(ms1, 2), "ms_per_token_step": round(per_tok_ms, 3),
            "roofline": {"bound": "hbm", "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s", "frac": round(ach / hbm, 4),
                         "bytes_per_step": int(step_bytes)}}


# ------------------------------------------------------------------------------------------------ CPU oracle port
def cpu_reference_step(args, budget_s: float, max_steps: int, warmup: int = 1):
    """Times the reference algorithm's training step (oracle/traj_vlm.py, plain fp32 PyTorch restatement of
    pointllm.py:181-253 + train.py:166-184) on the host cores, on a bounded sample of the workload: the largest per-step batch
    b in {B, B/2, ..., 1} whose fp32 activations fit in host RAM next to the 27 GB of fp32 weights (the GPU arm's B = 8 needs
    ~115 GB).  Same init recipe as the GPU arm (N(0, 0.02) LLM / head, fan-in scaled encoder).  kind="port": /root/reference
    does not exist on the GPU box."""
    import psutil
    import torch
    from egoscaler_b200 import synth
    from egoscaler_b200.config import TrajVLMConfig
    from oracle.traj_vlm import LlmCfg, PointCfg, TrajVLMOracle, train_loss

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    enc = getattr(args, "encoder", "v12")
    if args.config == "full":
        cfg = TrajVLMConfig.full() if enc == "v12" else TrajVLMConfig.full_base()
        llm = LlmCfg.full()
    else:
        cfg = TrajVLMConfig.tiny() if enc == "v12" else TrajVLMConfig.tiny_base()
        llm = LlmCfg.tiny()
    d, f, nl, V, L = llm.hidden_size, llm.intermediate_size, llm.num_layers, llm.vocab_size, 696
    w_gb = 4e-9 * (nl * (4 * d * d + 3 * d * f) + 2 * V * d)
    # fp32 autograd state per sample: ~(10 d + 5 f) floats per token per layer + 2 L^2 H attention maps + logits (x3 for CE)
    act_gb = 4e-9 * (nl * (L * (10 * d + 5 * f) + 2 * L * L * llm.num_heads) + 3 * L * V)
    avail_gb = psutil.virtual_memory().available / 1e9
    note = ""
    b = max(1, int(getattr(args, "batch", 1)))
    while b > 1 and w_gb * 1.1 + 6 + 1.3 * b * act_gb > 0.8 * avail_gb:     # conservative: never drive the host out of memory
        b //= 2
    if w_gb * 1.1 + 6 + act_gb > avail_gb:       # not even one sample next to all layers: fewer layers, say so
        llm.num_layers = max(1, int(nl * (avail_gb - 8) / (w_gb * 1.1 + act_gb)))
        note = f" (host RAM {avail_gb:.0f} GB: {llm.num_layers} of {nl} layers)"
    with torch.device("meta"):
        m = TrajVLMOracle(llm, PointCfg() if enc == "v12" else PointCfg.base())
    m = m.to_empty(device="cpu")
    g = torch.Generator().manual_seed(0)
    with torch.no_grad():
        for n_, p in list(m.named_parameters()) + list(m.named_buffers()):
            if not p.dtype.is_floating_point:
                p.zero_()
            elif n_.endswith("running_var") or (p.dim() == 1 and not n_.endswith("bias") and not n_.endswith("running_mean")):
                p.fill_(1.0)
            elif p.dim() == 1:
                p.zero_()
            elif "point_backbone" in n_ or "point_proj" in n_:
                p.normal_(0.0, 1.0 / (p.shape[1] ** 0.5), generator=g)
            else:
                p.normal_(0.0, 0.02, generator=g)
    m.configure_trainable()
    m.train()
    m.model.layers.eval()
    m.model.point_backbone.eval()
    opt = torch.optim.AdamW([q for q in m.parameters() if q.requires_grad], lr=2e-5)
    bt = synth.make_batch(cfg, b, seed=5, random_desc_len=True)
    P = bt["prompts"].shape[1]
    times = []
    t_begin = time.time()
    for i in range(warmup + max_steps):
        t0 = time.time()
        opt.zero_grad(set_to_none=True)
        logits = m(bt["tokens"], bt["attention_masks"], bt["pcrgbs"], fps_start=torch.zeros(b, dtype=torch.long), training=True)
        loss = train_loss(logits, bt["tokens"], P, cfg.pad_token_id)
        loss.backward()
        opt.step()
        dt = time.time() - t0
        if i >= warmup:
            times.append(dt)
        if time.time() - t_begin > budget_s and times:
            break
    t = sum(times) / len(times)
    return {"value": round(b / t, 5), "unit": UNIT, "cores": cores, "kind": "port", "steps_timed": len(times), "warmup": warmup,
            "s_per_step": round(t, 3), "batch_per_step": b,
            "sample": f"{b} sample(s) per step of the same workload (host RAM {avail_gb:.0f} GB bounds the batch), fp32, "
                      f"torch {torch.__version__} CPU, {torch.get_num_threads()} threads{note}"}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # keep the whole run within a few minutes whatever K/W the driver asks for: the budget stops the loop early
    warm = 1 if args.warmup > 0 else 0
    cb = cpu_reference_step(args, budget_s=args.cpu_budget_s, max_steps=max(1, args.steps), warmup=warm)
    world = int(os.environ.get("WORLD_SIZE", "1"))
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": world,
        "steps": cb["steps_timed"], "steps_requested": args.steps, "warmup": warm, "warmup_requested": args.warmup,
        "ms_per_step": round(cb["s_per_step"] * 1e3, 1), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic (collate_fn-shaped batches, random-init weights)",
        "config": config_dict(args, args.batch, world),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
