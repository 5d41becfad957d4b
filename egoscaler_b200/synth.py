"""Synthetic inputs and deterministic random-init weights.

The EgoScaler dataset and the PointLLM checkpoint are not released/available offline
(reference README.md:46; checkpoints submodule missing), so benchmarks and parity tests use
synthetic batches with exactly the layout ``CustomDataset.collate_fn`` produces
(reference dataset.py:150-194): ``tokens = [point block | description (padded to
max_desc_token) | sep | trajectory tokens (padded to max_traj_token)]``, ``attention_masks``
= non-pad, ``prompts`` = tokens up to and including the first ``<tsep>`` (dataset.py:180-182),
``pcrgbs`` = pc_norm'ed xyz + rgb (pointllm/data/utils.py:146-157).
"""
from __future__ import annotations

import zlib
from typing import Dict, Optional

import torch

from .config import (TrajVLMConfig, POINT_PATCH_TOKEN_ID, POINT_START_TOKEN_ID, POINT_END_TOKEN_ID,
                     TS_TOKEN_ID, TSEP_TOKEN_ID, TE_TOKEN_ID, TRAJ_TOKEN_BASE)

SEP_TOKEN_ID = 13          # "\n" in the LLaMA vocabulary (stand-in for dataset.sep_token_id)
MAX_DESC_TOKEN = 20        # train.py --max_desc_token
NUM_STEPS = 20             # train.py --num_steps
MAX_TRAJ_TOKEN = 160       # train.py --max_traj_token
NUM_BINS = 256             # train.py --num_bins


def pc_norm(xyz: torch.Tensor) -> torch.Tensor:
    """Centre on the centroid and scale to the unit sphere (pointllm/data/utils.py:146-157)."""
    xyz = xyz - xyz.mean(dim=0, keepdim=True)
    return xyz / xyz.pow(2).sum(-1).sqrt().max()


def make_batch(cfg: TrajVLMConfig, batch: int, seed: int = 1, npoints: Optional[int] = None,
               random_desc_len: bool = True) -> Dict[str, torch.Tensor]:
    """One collate_fn-shaped batch on the CPU (L = 696, prompt P = 544, T = 152 for the default
    point encoder)."""
    g = torch.Generator().manual_seed(seed)
    n = npoints or cfg.point.npoints
    Np = cfg.point.point_token_len
    clouds = []
    for _ in range(batch):
        xyz = pc_norm(torch.rand(n, 3, generator=g) * 2 - 1)
        rgb = torch.rand(n, 3, generator=g)
        clouds.append(torch.cat((xyz, rgb), dim=1))
    pcrgbs = torch.stack(clouds).float()

    pad = cfg.pad_token_id
    rows = []
    for _ in range(batch):
        dl = int(torch.randint(8, MAX_DESC_TOKEN + 1, (1,), generator=g)) if random_desc_len else MAX_DESC_TOKEN
        desc = torch.randint(3, 32000, (dl,), generator=g).tolist() + [pad] * (MAX_DESC_TOKEN - dl)
        bins = torch.randint(0, NUM_BINS, (NUM_STEPS, 6), generator=g) + TRAJ_TOKEN_BASE
        traj = [TS_TOKEN_ID]
        for s in range(NUM_STEPS):
            traj += bins[s].tolist() + [TSEP_TOKEN_ID]
        traj += [TE_TOKEN_ID, cfg.eos_token_id]
        traj += [pad] * (MAX_TRAJ_TOKEN - len(traj))
        row = [POINT_START_TOKEN_ID] + [POINT_PATCH_TOKEN_ID] * Np + [POINT_END_TOKEN_ID] + desc + [SEP_TOKEN_ID] + traj
        rows.append(row)
    tokens = torch.tensor(rows, dtype=torch.long)
    masks = tokens != pad
    pos = int(torch.where(tokens[0] == TSEP_TOKEN_ID)[0][0])          # dataset.py:180
    return {
        "pcrgbs": pcrgbs,
        "tokens": tokens,
        "attention_masks": masks,
        "prompts": tokens[:, :pos + 1].contiguous(),
        "prompt_masks": masks[:, :pos + 1].contiguous(),
    }


def det_tensor(key: str, shape, seed: int = 0) -> torch.Tensor:
    """Deterministic fp32 tensor for a ``state_dict`` key (CPU generator seeded by crc32(key)).

    Both the real reference (oracle/make_golden.py) and the tests fill their modules with this
    recipe, so golden files need not store the 43 M tiny-config weights.
    """
    g = torch.Generator().manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    shape = tuple(shape)
    leaf = key.rsplit(".", 1)[-1]
    in_point = ("point_backbone" in key) or ("point_proj" in key)
    if leaf == "num_batches_tracked":
        return torch.zeros(shape, dtype=torch.long)
    if leaf == "running_mean":
        return 0.1 * torch.randn(shape, generator=g)
    if leaf == "running_var":
        return 0.5 + torch.rand(shape, generator=g)
    if leaf in ("cls_token", "cls_pos"):
        return 0.5 * torch.randn(shape, generator=g)
    if len(shape) == 1:
        if leaf == "bias":
            return 0.05 * torch.randn(shape, generator=g)
        return 1.0 + 0.1 * torch.randn(shape, generator=g)           # norm / BN scale
    fan_in = shape[1]
    std = (1.0 / fan_in ** 0.5) if in_point else 0.02
    return std * torch.randn(shape, generator=g)


def det_state_dict(template: Dict[str, torch.Tensor], seed: int = 0, round_bf16: bool = True) -> Dict[str, torch.Tensor]:
    """Fill every entry of ``template`` (a state_dict) deterministically.  With ``round_bf16`` the
    values are rounded to bf16-representable numbers so an fp32 oracle and the bf16 CUDA path hold
    bit-identical weights."""
    out = {}
    for k, v in template.items():
        t = det_tensor(k, v.shape, seed)
        if t.dtype.is_floating_point:
            if round_bf16:
                t = t.to(torch.bfloat16).float()
            t = t.to(v.dtype) if v.dtype.is_floating_point else t
        out[k] = t
    return out
