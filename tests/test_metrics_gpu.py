"""GPU: on-device trajectory metrics (csrc/metrics.cu) vs outputs of the REAL egoscaler/models/utils/metrics.py
(tests/golden/traj_metrics.npz) and the oracle restatement; plus the decode -> poses -> metrics chain without host loops."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops
from oracle import metrics as om
from oracle import traj_tokenizer as tt
from conftest import GOLDEN


def test_metrics_match_reference_golden():
    g = np.load(os.path.join(GOLDEN, "traj_metrics.npz"))
    gen, gt = torch.from_numpy(g["gen"]).cuda(), torch.from_numpy(g["gt"]).cuda()
    gl, tl = torch.from_numpy(g["gen_len"]).cuda(), torch.from_numpy(g["gt_len"]).cuda()
    out = ops.traj_metrics(gen, gt, gl, tl, cols=(0, 6), rot_col=3).cpu().numpy()
    ref = g["ade_fde_ide_gd_all6"]
    # displacement metrics: fp32 like NumPy on float32 input (summation order differs by an ulp or two)
    np.testing.assert_allclose(out[:, :3], ref[:, :3], rtol=2e-6, atol=1e-7)
    # angular distance: fp64 like scipy; 2*arccos amplifies round-off near dot = 1 (identical rotations) to ~1e-8
    np.testing.assert_allclose(out[:, 3], ref[:, 3], rtol=1e-9, atol=5e-8)
    out3 = ops.traj_metrics(gen, gt, gl, tl, cols=(0, 3), rot_col=-1).cpu().numpy()
    np.testing.assert_allclose(out3[:, :3], g["ade_fde_ide_xyz"], rtol=2e-6, atol=1e-7)
    assert np.isnan(out3[:, 3]).all()


def test_metrics_full_length_and_skipped_rows():
    rng = np.random.default_rng(3)
    B, T = 257, 20
    gen = rng.uniform(-1, 1, (B, T, 6)).astype(np.float32)
    gt = rng.uniform(-1, 1, (B, T, 6)).astype(np.float32)
    gl = np.full(B, T, np.int32)
    gl[5] = 0                                                     # un-parsable generation: train.py:247-248 `continue`
    out = ops.traj_metrics(torch.from_numpy(gen).cuda(), torch.from_numpy(gt).cuda(), torch.from_numpy(gl).cuda()).cpu().numpy()
    assert np.isnan(out[5]).all()
    for b in (0, 1, 100, 256):
        ref = [om.ade(gen[b], gt[b]), om.fde(gen[b], gt[b]), om.ide(gen[b], gt[b]), om.angular_distance(gen[b, :, 3:], gt[b, :, 3:])]
        np.testing.assert_allclose(out[b], ref, rtol=2e-6, atol=5e-8)


def test_ids_to_poses_to_metrics_chain():
    """generated ids -> poses (+count) -> metrics entirely on device == the reference's per-sample host loop
    (train.py:240-260: str_to_float, pad by repeating the last frame, metrics)."""
    rng = np.random.default_rng(9)
    B, steps = 6, 20
    ids = np.zeros((B, 152), np.int64)
    lens = [20, 12, 1, 20, 7, 0]
    for b in range(B):
        seq = [tt.TS_TOKEN]
        for _ in range(lens[b]):
            seq += (rng.integers(0, 256, 6) + tt.TRAJ_TOKEN_BASE).tolist() + [tt.TSEP_TOKEN]
        seq += [tt.TE_TOKEN, 2]
        ids[b, :len(seq)] = seq
    gt = rng.uniform(-1, 1, (B, steps, 6)).astype(np.float32)
    poses, cnt = ops.traj_ids_to_poses(torch.from_numpy(ids).cuda(), max_steps=steps, pad_repeat=False)
    out = ops.traj_metrics(poses, torch.from_numpy(gt).cuda(), cnt).cpu().numpy()
    for b in range(B):
        p = tt.ids_to_poses(ids[b].tolist())
        if p is None or len(p) == 0:
            assert np.isnan(out[b]).all()
            continue
        ref = [om.ade(p, gt[b]), om.fde(p, gt[b]), om.ide(p, gt[b]), om.angular_distance(p[:, 3:], gt[b, :, 3:])]
        np.testing.assert_allclose(out[b], ref, rtol=2e-6, atol=5e-8)
