"""Data side of the hot path: the EgoScaler on-disk formats -> a ``collate_fn``-shaped batch, with the
per-sample NumPy work of the reference's (un-shipped) ``CustomDataset.__getitem__`` done as batched device kernels.

On-disk formats (what the dataset-construction pipeline writes):
  * ``{data_dir}/{split}.json`` - COCO-like ``{"images": [{id, dataset_name, video_uid, file_name, ...}], "annotations":
    [{image_id, action_description, ...}]}`` (egoscaler/models/utils/dataset_base.py:31-40,76-88; egoscaler/data/README.md:11-31);
  * ``{root_dir}/trajs/{dataset_name}/{video_uid}/{file_name}.pkl`` - pickle ``{"init_bbox", "traj_quat" [n,7], "traj_rotvec"
    [n,6]}`` (egoscaler/data/train/7_get_object_trajectory.py:324-328); ``traj_rotvec`` is what the model trains on
    (dataset_base.py:97-101);
  * ``{root_dir}/pcrgbs/{dataset_name}/{video_uid}/{file_name}.npy`` - coloured point cloud [N, 6] (README.md:33-37).

Batch layout = ``CustomDataset.collate_fn`` (egoscaler/models/pointllm/dataset.py:150-194): ``tokens = [point block | description
padded to max_desc_token | <sep> | trajectory tokens padded to max_traj_token]``, ``attention_masks``, ``prompts`` = tokens up to
and including the first ``<tsep>`` (:180-182), ``pcrgbs``, ``trajectories``, ``trajectory_masks``, ``max_abs``, ``image_ids``.

Per-sample pipeline restated from the shipped pieces (the reference's ``__getitem__`` itself is missing): description
template ``DESC2TRAJ`` (dataset.py:16-19), ``preprocess_traj`` then ``smoothing_traj`` in the order ``compute_mean_std`` applies
them (dataset.py:91-94), the normaliser as the exact inverse of ``denorm`` (dataset.py:126-148), ``discretize_action``
(utils/utils.py:13-16), token templates (constant.py), ``pc_norm`` (pointllm/data/utils.py:146-157).  Device kernels:
``esb_traj_preprocess``, ``esb_traj_poses_to_ids``, ``esb_traj_quantize``, ``esb_pc_sample_norm``.
"""
from __future__ import annotations

import json
import os
import pickle
import zlib
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import ops
from .config import (TrajVLMConfig, POINT_PATCH_TOKEN_ID, POINT_START_TOKEN_ID, POINT_END_TOKEN_ID, TS_TOKEN_ID, TSEP_TOKEN_ID,
                     TE_TOKEN_ID, TRAJ_TOKEN_BASE)

DESC2TRAJ = {"desc": "Action description: {desc}", "traj": "To execute the description, action trajectory will be {traj}"}
SEP_TOKEN = "<sep>"
# egoscaler/configs/dataset.py:1-7
XYZ_RANGES = (-2.0, 2.0, -2.0, 2.0, 0.0, 2.5)


def encode_text(tokenizer, text: str) -> List[int]:
    """Token ids of ``text`` without special tokens.  A HF tokenizer is called like dataset.py:52; the offline stand-in
    (builder.SyntheticTokenizer has no vocabulary) maps every whitespace-separated word to a stable id in [3, 32000)."""
    if callable(tokenizer) and not hasattr(tokenizer, "special"):
        return list(tokenizer(text, add_special_tokens=False).input_ids)
    return [3 + zlib.crc32(w.encode()) % (32000 - 3) for w in text.split()]


class TrajDataset(torch.utils.data.Dataset):
    """File access of ``DatasetBase`` (dataset_base.py:8-103) + the point cloud of the README layout.  Items are raw host
    objects; all arithmetic happens batched in :class:`DeviceCollator`."""

    def __init__(self, args, split: str, tokenizer):
        if split not in ("train", "val", "test"):
            raise ValueError(f"Invalid split: {split}. Expected 'train', 'val', or 'test'.")          # dataset_base.py:35
        self.args, self.split, self.tokenizer = args, split, tokenizer
        self.root_dir = args.root_dir
        with open(f"{args.data_dir}/{split}.json", "r") as f:
            dataset = json.load(f)
        self.id2data = {item["id"]: item for item in dataset["images"]}
        self.annotations = dataset["annotations"]

    def __len__(self) -> int:
        return len(self.annotations)

    def paths(self, data: dict) -> Dict[str, str]:
        sub = f"{data['dataset_name']}/{data['video_uid']}/{data['file_name']}"
        return {"traj": f"{self.root_dir}/trajs/{sub}.pkl", "pcrgb": f"{self.root_dir}/pcrgbs/{sub}.npy"}

    def __getitem__(self, item: int) -> dict:
        annot = self.annotations[item]
        data = self.id2data[annot["image_id"]]
        desc = annot["action_description"]
        desc = desc.lower() if isinstance(desc, str) else str(desc)                               # dataset_base.py:85-88
        p = self.paths(data)
        with open(p["traj"], "rb") as f:
            traj = np.asarray(pickle.load(f)["traj_rotvec"], dtype=np.float32)                    # dataset_base.py:97-101
        pcrgb = np.load(p["pcrgb"]).astype(np.float32)
        return {"image_id": int(annot["image_id"]), "desc": desc, "traj": traj, "pcrgb": pcrgb}


class DeviceCollator:
    """``collate_fn`` (dataset.py:150-194) over raw :class:`TrajDataset` items, arithmetic on ``device``."""

    def __init__(self, cfg: TrajVLMConfig, tokenizer, device="cuda", num_steps: int = 20, max_traj_token: int = 160,
                 max_desc_token: int = 20, num_bins: int = 256, do_norm: bool = True, do_standard: bool = False,
                 smooth_traj: bool = False, mean: Optional[Sequence[float]] = None, std: Optional[Sequence[float]] = None,
                 npoints: Optional[int] = None, seed: int = 0):
        assert not (do_norm and do_standard), "Cannot enable both normalization methods."         # dataset.py:44
        if 3 + 7 * num_steps > max_traj_token:
            raise ValueError(f"max_traj_token={max_traj_token} cannot hold {num_steps} steps ({3 + 7 * num_steps} tokens)")
        self.cfg, self.tok, self.dev = cfg, tokenizer, torch.device(device)
        self.num_steps, self.max_traj_token, self.max_desc_token, self.num_bins = num_steps, max_traj_token, max_desc_token, num_bins
        self.do_norm, self.do_standard, self.smooth = do_norm, do_standard, smooth_traj
        self.npoints = npoints or cfg.point.npoints
        self.gen = torch.Generator().manual_seed(seed)
        sep = encode_text(tokenizer, SEP_TOKEN) if not hasattr(tokenizer, "special") else [13]     # dataset.py:52
        self.sep_ids = torch.tensor(sep, dtype=torch.long)
        self.ranges = torch.tensor(XYZ_RANGES, dtype=torch.float64, device=self.dev)
        if do_standard:
            if mean is None or std is None:
                raise ValueError("do_standard needs the mean / std of norm_param.json (dataset.py:103-124)")
            self.mean = torch.tensor(np.asarray(mean, np.float32), device=self.dev)
            self.std = torch.tensor(np.asarray(std, np.float32), device=self.dev)
        Np = cfg.point.point_token_len
        self.point_block = torch.tensor([POINT_START_TOKEN_ID] + [POINT_PATCH_TOKEN_ID] * Np + [POINT_END_TOKEN_ID], dtype=torch.long) \
            if cfg.mm_use_point_start_end else torch.tensor([POINT_PATCH_TOKEN_ID] * Np, dtype=torch.long)

    # ---- host: text
    def encode_desc(self, desc: str):
        ids = encode_text(self.tok, DESC2TRAJ["desc"].format(desc=desc))[: self.max_desc_token]
        pad = self.cfg.pad_token_id
        n = len(ids)
        return ids + [pad] * (self.max_desc_token - n), [True] * n + [False] * (self.max_desc_token - n)

    # ---- device: trajectories
    def tokenize_trajs(self, trajs: List[np.ndarray]):
        """raw [n_i, 6] float32 trajectories -> (token ids [B, max_traj_token] int64, resampled physical poses [B, ns, 6],
        padding mask [B, ns] int32, max_abs [B, 6] float32)."""
        B = len(trajs)
        Tmax = max(int(t.shape[0]) for t in trajs)
        host = torch.zeros(B, Tmax, 6, dtype=torch.float32)
        lens = torch.tensor([int(t.shape[0]) for t in trajs], dtype=torch.int32)
        for b, t in enumerate(trajs):
            host[b, : t.shape[0]] = torch.from_numpy(np.ascontiguousarray(t[:, :6], dtype=np.float32))
        raw, lens_d = host.to(self.dev), lens.to(self.dev)
        poses, mask = ops.traj_preprocess(raw, lens_d, self.num_steps, smooth=False)              # preprocess_traj
        if self.smooth:
            poses, _ = ops.traj_preprocess(poses, None, self.num_steps, smooth=True)              # then smoothing_traj (dataset.py:91-94)
        max_abs = torch.ones(B, 6, device=self.dev, dtype=torch.float32)
        if self.do_standard:                     # inverse of dataset.py:146-148: standardise, then scale each dimension into [-1, 1]
            z = (poses - self.mean) / self.std
            max_abs = z.abs().amax(dim=1).clamp_min(1e-8)
            ids, _ = ops.traj_poses_to_ids(z / max_abs[:, None, :], None, self.num_bins, base_id=TRAJ_TOKEN_BASE, ts_id=TS_TOKEN_ID,
                                           tsep_id=TSEP_TOKEN_ID, te_id=TE_TOKEN_ID, eos_id=self.cfg.eos_token_id,
                                           pad_id=self.cfg.pad_token_id, out_len_total=self.max_traj_token)
        else:                                    # inverse of dataset.py:139-145 inside the kernel (ranges of configs/dataset.py)
            ids, _ = ops.traj_poses_to_ids(poses, None, self.num_bins, base_id=TRAJ_TOKEN_BASE, ts_id=TS_TOKEN_ID, tsep_id=TSEP_TOKEN_ID,
                                           te_id=TE_TOKEN_ID, eos_id=self.cfg.eos_token_id, pad_id=self.cfg.pad_token_id,
                                           ranges=self.ranges if self.do_norm else None, out_len_total=self.max_traj_token)
        return ids, poses, mask, max_abs

    # ---- device: point clouds
    def sample_clouds(self, clouds: List[np.ndarray]) -> torch.Tensor:
        """[N_i, 6] clouds -> [B, npoints, 6]: ``npoints`` rows drawn without replacement (with, when the cloud is smaller) from
        the seeded host generator, then pc_norm (pointllm/data/utils.py:146-157) on device."""
        B, S = len(clouds), self.npoints
        Nmax = max(int(c.shape[0]) for c in clouds)
        host = torch.zeros(B, Nmax, 6, dtype=torch.float32)
        idx = torch.empty(B, S, dtype=torch.long)
        for b, c in enumerate(clouds):
            n = int(c.shape[0])
            if n < 1 or c.shape[1] < 6:
                raise ValueError(f"point cloud {b} has shape {tuple(c.shape)}; need [N >= 1, 6] (xyz + rgb)")
            host[b, :n] = torch.from_numpy(np.ascontiguousarray(c[:, :6], dtype=np.float32))
            idx[b] = torch.randperm(n, generator=self.gen)[:S] if n >= S else torch.randint(0, n, (S,), generator=self.gen)
        return ops.pc_sample_norm(host.to(self.dev), idx.to(self.dev))

    def __call__(self, samples: List[dict]) -> Dict[str, torch.Tensor]:
        B = len(samples)
        pad = self.cfg.pad_token_id
        desc = [self.encode_desc(s["desc"]) for s in samples]
        desc_ids = torch.tensor([d[0] for d in desc], dtype=torch.long)
        desc_mask = torch.tensor([d[1] for d in desc], dtype=torch.bool)
        head = torch.cat((self.point_block.expand(B, -1), desc_ids, self.sep_ids.expand(B, -1)), dim=1).to(self.dev)
        head_mask = torch.cat((torch.ones(B, self.point_block.numel(), dtype=torch.bool), desc_mask,
                               torch.ones(B, self.sep_ids.numel(), dtype=torch.bool)), dim=1).to(self.dev)
        traj_ids, poses, tmask, max_abs = self.tokenize_trajs([s["traj"] for s in samples])
        tokens = torch.cat((head, traj_ids), dim=1)                                                  # dataset.py:177
        masks = torch.cat((head_mask, traj_ids != pad), dim=1)                                      # dataset.py:178
        pos = int(torch.where(tokens[0] == TSEP_TOKEN_ID)[0][0])                                   # dataset.py:180
        return {
            "image_ids": torch.tensor([s["image_id"] for s in samples], dtype=torch.long),
            "pcrgbs": self.sample_clouds([s["pcrgb"] for s in samples]),
            "prompts": tokens[:, : pos + 1].contiguous(),
            "prompt_masks": masks[:, : pos + 1].contiguous(),
            "tokens": tokens,
            "attention_masks": masks,
            "trajectories": poses,
            "trajectory_masks": traj_ids != pad,                   # dataset.py:192 hands out the TOKEN mask under this name
            "trajectory_step_masks": tmask.bool(),                 # preprocess_traj's padding mask (1 = real step)
            "max_abs": max_abs,
        }
