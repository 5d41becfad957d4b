"""GPU: RGB-D -> coloured point cloud front-end (csrc/rgbd_frontend.cu) vs outputs of the REAL
get_points_colors / pc_norm (tests/golden/rgbd_frontend.npz), the oracle at the Aria frame size, and the chain into the point
encoder's FPS."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops
from oracle import rgbd_frontend as rf
from conftest import GOLDEN

CASES = ["plain", "boxes_thres", "aria_crop", "all_boxed"]


def _run(rgbd, boxes, c, fx, fy, thr, cap=None):
    bx = torch.from_numpy(np.asarray(boxes, np.int32).reshape(-1, 4)).cuda() if boxes is not None and len(boxes) else None
    out, cnt = ops.rgbd_to_points(torch.from_numpy(rgbd).cuda(), bx, c, fx, fy, thr, cap=cap)
    n = int(cnt.item())
    return out.cpu().numpy(), n


@pytest.mark.parametrize("name", CASES)
def test_unprojection_matches_reference_golden(name):
    """Bit-exact: same survivors in the same (row-major) order, fp64 unprojection rounded once to fp32, fp32 colours."""
    g = np.load(os.path.join(GOLDEN, "rgbd_frontend.npz"))
    c, fx, fy, thr = g[f"{name}/cam"]
    out, n = _run(g[f"{name}/rgbd"], g[f"{name}/boxes"], c, fx, fy, None if np.isnan(thr) else float(thr))
    ref_p, ref_c = g[f"{name}/points"], g[f"{name}/colors"]
    assert n == len(ref_p)
    assert np.array_equal(out[:n, :3], ref_p.astype(np.float32))
    assert np.array_equal(out[:n, 3:], ref_c)


@pytest.mark.parametrize("name", CASES[:3])
def test_sample_and_pc_norm_match_reference_golden(name):
    g = np.load(os.path.join(GOLDEN, "rgbd_frontend.npz"))
    c, fx, fy, thr = g[f"{name}/cam"]
    out, cnt = ops.rgbd_to_points(torch.from_numpy(g[f"{name}/rgbd"]).cuda(),
                                  torch.from_numpy(g[f"{name}/boxes"]).cuda() if len(g[f"{name}/boxes"]) else None, c, fx, fy,
                                  None if np.isnan(thr) else float(thr))
    idx = torch.from_numpy(g[f"{name}/sample_idx"]).cuda()
    got = ops.pc_sample_norm(out, idx[None], n=int(cnt.item()))[0].cpu().numpy()
    ref = g[f"{name}/pc_norm"]
    # the reference normalises the fp64 points; the device path stores fp32 points first (rel 6e-8) and normalises in fp64
    np.testing.assert_allclose(got[:, :3], ref[:, :3], rtol=0, atol=2e-6)
    assert np.array_equal(got[:, 3:], ref[:, 3:].astype(np.float32))
    assert abs(float(np.sqrt((got[:, :3].astype(np.float64) ** 2).sum(1)).max()) - 1.0) < 1e-6
    assert float(np.abs(got[:, :3].astype(np.float64).mean(0)).max()) < 1e-6


def test_aria_frame_vs_oracle_and_capacity():
    """Full Aria pinhole frame (1408x1408, configs/camera.py:6-9): bit-exact vs the oracle; cap smaller than the survivor count
    truncates the output but still reports the full count; empty boxes list == None."""
    H = W = 1408
    rgbd = rf.synth_rgbd(H, W, seed=7)
    boxes = [(100, 200, 400, 900), (1000, 0, 1408, 300), (640, 640, 641, 641)]
    ref_p, ref_c = rf.get_points_colors(rgbd, boxes, 703.5, 605.343, 605.343, 5.0)
    out, n = _run(rgbd, boxes, 703.5, 605.343, 605.343, 5.0)
    assert n == len(ref_p) and n > 1_000_000
    assert np.array_equal(out[:n, :3], ref_p.astype(np.float32))
    assert np.array_equal(out[:n, 3:], ref_c)
    out2, n2 = _run(rgbd, boxes, 703.5, 605.343, 605.343, 5.0, cap=4096)
    assert n2 == n and out2.shape == (4096, 6) and np.array_equal(out2, out[:4096])
    o3, n3 = _run(rgbd[:64, :64].copy(), [], 31.5, 50.0, 50.0, None)
    p3, c3 = rf.get_points_colors(rgbd[:64, :64], None, 31.5, 50.0, 50.0, None)
    assert n3 == len(p3) and np.array_equal(o3[:n3, :3], p3.astype(np.float32)) and np.array_equal(o3[:n3, 3:], c3)


def test_frontend_feeds_point_encoder_fps():
    """rgbd -> points -> 8192 sampled + pc_norm -> fps: the output is a valid `pcrgbs` row for the encoder (finite, unit ball)
    and batched pc_sample_norm == per-cloud."""
    rgbd = rf.synth_rgbd(256, 256, seed=9)
    out, cnt = ops.rgbd_to_points(torch.from_numpy(rgbd).cuda(), None, 127.5, 200.0, 200.0, 5.5)
    n = int(cnt.item())
    g = torch.Generator().manual_seed(0)
    idx = torch.stack([torch.randperm(n, generator=g)[:8192] for _ in range(3)]).cuda()
    clouds = out[None].expand(3, -1, -1).contiguous()
    pc = ops.pc_sample_norm(clouds, idx, n=n)
    assert pc.shape == (3, 8192, 6) and bool(torch.isfinite(pc).all())
    for b in range(3):
        one = ops.pc_sample_norm(out, idx[b:b + 1], n=n)[0]
        assert torch.equal(one, pc[b])
        ref = rf.pc_norm(out[:n].cpu().numpy().astype(np.float64)[idx[b].cpu().numpy()])
        np.testing.assert_allclose(pc[b].cpu().numpy(), ref, rtol=0, atol=1e-6)
    centers, cidx = ops.fps(pc, 512, torch.zeros(3, dtype=torch.long, device="cuda"))[:2]
    assert centers.shape[:2] == (3, 512) and bool(torch.isfinite(centers).all())
