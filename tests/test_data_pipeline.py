"""The data side of the hot path (egoscaler_b200/data.py): on-disk EgoScaler formats -> collate_fn-shaped batch.

CPU part: the file formats (dataset_base.py:31-40,76-101; 7_get_object_trajectory.py:324-328; data/README.md:11-37) and the
host-side text handling.  GPU part: the batched device collate against the NumPy oracle chain pinned to the real reference
(oracle/traj_tokenizer.py: preprocess_traj -> smoothing_traj -> norm_poses -> discretize_action -> token layout), the batch
layout of dataset.py:150-194, and the round trip through ids -> poses -> denorm."""
import argparse
import json
import os
import pickle

import numpy as np
import pytest
import torch

from egoscaler_b200.builder import SyntheticTokenizer
from egoscaler_b200.config import TrajVLMConfig
from egoscaler_b200.data import DESC2TRAJ, DeviceCollator, TrajDataset, encode_text
from oracle import traj_tokenizer as tt


def _make_dataset(root, n=5, seed=0):
    rng = np.random.default_rng(seed)
    data_dir = os.path.join(root, "data")
    os.makedirs(data_dir, exist_ok=True)
    images, annots = [], []
    lens = [3, 20, 47, 11, 80][:n]
    for i in range(n):
        sub = f"egoexo/take_{i % 2}"
        name = f"frame_{i:03d}"
        images.append({"id": 100 + i, "dataset_name": "egoexo", "video_uid": f"take_{i % 2}", "file_name": name})
        annots.append({"image_id": 100 + i, "id": 100 + i, "action_description": f"C Picks up the cup number {i} with the left hand"})
        os.makedirs(os.path.join(root, "trajs", sub), exist_ok=True)
        os.makedirs(os.path.join(root, "pcrgbs", sub), exist_ok=True)
        T = lens[i]
        rot = rng.uniform(-3.0, 3.0, (T, 3))
        pos = np.stack([rng.uniform(-1.9, 1.9, T), rng.uniform(-1.9, 1.9, T), rng.uniform(0.05, 2.4, T)], 1)
        traj = {"init_bbox": rng.normal(size=(8, 3)), "traj_quat": rng.normal(size=(T, 7)), "traj_rotvec": np.hstack([pos, rot])}
        with open(os.path.join(root, "trajs", sub, name + ".pkl"), "wb") as f:
            pickle.dump(traj, f)
        npts = 9000 if i != 3 else 5000                                  # one cloud smaller than 8192 points
        pc = np.concatenate([rng.normal(size=(npts, 3)) * 0.4 + 0.3, rng.uniform(0, 1, (npts, 3))], 1).astype(np.float32)
        np.save(os.path.join(root, "pcrgbs", sub, name + ".npy"), pc)
    for split in ("train", "val"):
        with open(os.path.join(data_dir, f"{split}.json"), "w") as f:
            json.dump({"images": images, "annotations": annots}, f)
    return argparse.Namespace(root_dir=root, data_dir=data_dir)


def test_dataset_reads_the_reference_layout(tmp_path):
    args = _make_dataset(str(tmp_path))
    ds = TrajDataset(args, "train", SyntheticTokenizer())
    assert len(ds) == 5
    it = ds[2]
    assert it["image_id"] == 102 and it["desc"] == "c picks up the cup number 2 with the left hand"       # lower-cased
    assert it["traj"].shape == (47, 6) and it["traj"].dtype == np.float32 and it["pcrgb"].shape == (9000, 6)
    with pytest.raises(ValueError):
        TrajDataset(args, "dev", SyntheticTokenizer())
    a, b = encode_text(SyntheticTokenizer(), "pick up the cup"), encode_text(SyntheticTokenizer(), "pick up the cup")
    assert a == b and len(a) == 4 and all(3 <= t < 32000 for t in a)
    assert DESC2TRAJ["desc"].format(desc="x") == "Action description: x"


@pytest.mark.gpu
@pytest.mark.parametrize("smooth", [False, True])
def test_device_collate_matches_numpy_oracle_and_reference_layout(tmp_path, smooth):
    from egoscaler_b200 import ops
    args = _make_dataset(str(tmp_path))
    tok = SyntheticTokenizer()
    tok.add_tokens(["<ts>", "<tsep>", "<te>"]); tok.add_tokens([f"<p{i}>" for i in range(256)])
    ds = TrajDataset(args, "train", tok)
    cfg = TrajVLMConfig.tiny()
    col = DeviceCollator(cfg, tok, "cuda", num_steps=20, max_traj_token=160, max_desc_token=20, do_norm=True, smooth_traj=smooth)
    samples = [ds[i] for i in range(5)]
    bt = col(samples)
    B, Np = 5, cfg.point.point_token_len
    L = Np + 2 + 20 + 1 + 160
    assert bt["tokens"].shape == (B, L) and bt["attention_masks"].shape == (B, L) and bt["attention_masks"].dtype == torch.bool
    assert bt["pcrgbs"].shape == (B, 8192, 6) and bt["trajectories"].shape == (B, 20, 6) and bt["max_abs"].shape == (B, 6)
    tokens = bt["tokens"].cpu().numpy()
    # ---- trajectory tokens == the oracle chain, sample by sample
    for b, s in enumerate(samples):
        t, m = tt.preprocess_traj(s["traj"], 20)
        if smooth:
            t = tt.smoothing_traj(t)
        ids = tt.poses_to_ids(tt.norm_poses(t), 256)
        ids = ids + [0] * (160 - len(ids))
        assert tokens[b, -160:].tolist() == ids, b
        assert np.array_equal(bt["trajectories"][b].cpu().numpy(), np.asarray(t, np.float32))
        assert bt["trajectory_step_masks"][b].cpu().numpy().astype(int).tolist() == m.tolist()
    # ---- layout of dataset.py:150-194
    assert (tokens[:, 0] == 32001).all() and (tokens[:, 1:1 + Np] == 32000).all() and (tokens[:, 1 + Np] == 32002).all()
    d0 = Np + 2
    words = len(DESC2TRAJ["desc"].format(desc=samples[0]["desc"]).split())
    nd = min(words, 20)
    assert (tokens[0, d0:d0 + nd] != 0).all() and (tokens[0, d0 + nd:d0 + 20] == 0).all()            # mid-sequence padding
    assert (tokens[:, d0 + 20] == 13).all() and (tokens[:, d0 + 21] == 32003).all()                    # <sep>, <ts>
    masks = bt["attention_masks"].cpu().numpy()
    assert np.array_equal(masks, tokens != 0)
    P = bt["prompts"].shape[1]
    assert P == d0 + 21 + 1 + 7 and (tokens[:, P - 1] == 32004).all()                                  # through the first <tsep>
    assert torch.equal(bt["prompts"], bt["tokens"][:, :P]) and torch.equal(bt["trajectory_masks"], bt["tokens"][:, -160:] != 0)
    # ---- point clouds: pc_norm'ed (centroid 0, max radius 1), colours untouched
    pc = bt["pcrgbs"]
    assert float(pc[..., :3].mean(dim=1).abs().max()) < 1e-4
    assert float((pc[..., :3].norm(dim=-1).amax(dim=1) - 1.0).abs().max()) < 1e-5
    assert float(pc[..., 3:].min()) >= 0.0 and float(pc[..., 3:].max()) <= 1.0
    # ---- round trip: ids -> poses -> denorm lands within one bin of the resampled trajectory
    poses, cnt = ops.traj_ids_to_poses(bt["tokens"][:, P - 8:].contiguous(), max_steps=20, scale=ops.SCALE_DENORM, pad_repeat=False)
    assert cnt.tolist() == [20] * B
    err = (poses - bt["trajectories"]).abs().cpu().numpy()
    width = np.array([4.0, 4.0, 2.5, 2 * np.pi, 2 * np.pi, 2 * np.pi]) / 255.0
    assert (err <= width * 1.001).all()
    # ---- the batch drives the model (forward) and the trainer
    from egoscaler_b200 import synth
    from egoscaler_b200.model import TrajPointLLMForCausalLM
    from egoscaler_b200.trainer import Trainer
    m = TrajPointLLMForCausalLM(cfg, "cuda")
    m.load_state_dict(synth.det_state_dict(m.state_dict(), seed=0, round_bf16=True), strict=True)
    m.train()
    loss = Trainer(m, lr=2e-5).train_step(bt)
    assert 9.0 < float(loss) < 12.5


@pytest.mark.gpu
def test_device_collate_standardised_variant(tmp_path):
    """--do_standard: (x - mean) / std then per-sample max-abs scaling; ``denorm`` (dataset.py:146-148) inverts it."""
    from egoscaler_b200 import ops
    args = _make_dataset(str(tmp_path))
    tok = SyntheticTokenizer()
    ds = TrajDataset(args, "val", tok)
    cfg = TrajVLMConfig.tiny()
    mean, std = np.array([0.1, -0.2, 1.2, 0.0, 0.1, -0.1], np.float32), np.array([1.1, 0.9, 0.7, 1.5, 1.6, 1.4], np.float32)
    col = DeviceCollator(cfg, tok, "cuda", do_norm=False, do_standard=True, mean=mean, std=std)
    bt = col([ds[i] for i in range(3)])
    P = bt["prompts"].shape[1]
    raw, cnt = ops.traj_ids_to_poses(bt["tokens"][:, P - 8:].contiguous(), max_steps=20, pad_repeat=False)
    rec = tt.denorm(raw.cpu().numpy(), False, bt["max_abs"].cpu().numpy(), mean, std)
    tol = (2.0 / 255.0) * bt["max_abs"].cpu().numpy()[:, None, :] * std * 1.01
    assert (np.abs(rec - bt["trajectories"].cpu().numpy()) <= tol + 1e-5).all()
    with pytest.raises(AssertionError):
        DeviceCollator(cfg, tok, "cuda", do_norm=True, do_standard=True, mean=mean, std=std)
