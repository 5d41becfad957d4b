"""TEST INFRASTRUCTURE ONLY - CPU restatement (numpy) of the RGB-D -> coloured point cloud front-end.

  get_points_colors   egoscaler/data/tools/pcm_tools.py:68-96 (call sites data/train/7_get_object_trajectory.py:191-200,244-253)
  pc_norm             egoscaler/models/pointllm/pointllm/data/utils.py:146-157

Pinned: oracle/make_golden.py imports both reference functions (open3d stubbed - it is only imported at module scope, the
function itself is numpy) and stores their outputs in tests/golden/rgbd_frontend.npz; tests/test_oracle_golden.py checks this
restatement against them.  Only tests/, smoke() and bench.py's cpu_baseline may import this module.
"""
import numpy as np


def get_points_colors(rgbd, boxes, principal_p, focal_len_x, focal_len_y, d_thres=None):
    """rgbd [H,W,4] float32 (uint8 colours ++ float32 depth concatenate to float32, 7_get_object_trajectory.py:157);
    boxes: iterable of (xmin, ymin, xmax, ymax).  -> points [n,3] float64, colours [n,3] float32, row-major pixel order."""
    rgbd = np.asarray(rgbd)
    H, W = rgbd.shape[:2]
    image, z = rgbd[:, :, :3], rgbd[:, :, 3]
    jj, ii = np.mgrid[0:H, 0:W]                                  # jj = row (y), ii = column (x); int64
    x = (ii - principal_p) / focal_len_x                          # pcm_tools.py:73-75 (float64)
    y = (jj - principal_p) / focal_len_y
    pts = np.stack((x * z, y * z, z), axis=-1).reshape(-1, 3)     # :77
    cols = image.reshape(-1, 3) / 255.0                           # :78 (float32 / python float stays float32)
    valid = np.all(image != 0, axis=2)                            # :79
    for (xmin, ymin, xmax, ymax) in (boxes if boxes is not None else ()):
        valid[ymin:ymax, xmin:xmax] = False                       # :80-85
    if d_thres is not None:
        valid &= z < d_thres                                      # :87-89
    v = valid.ravel()
    return pts[v], cols[v]


def pc_norm(pc):
    """pointllm/data/utils.py:146-157: xyz - centroid, / max L2 norm; other columns untouched."""
    pc = np.asarray(pc)
    xyz = pc[:, :3] - np.mean(pc[:, :3], axis=0)
    m = np.max(np.sqrt(np.sum(xyz ** 2, axis=1)))
    return np.concatenate((xyz / m, pc[:, 3:]), axis=1)


def synth_rgbd(H, W, seed=0, zero_frac=0.05):
    """A deterministic RGB-D frame: uint8 colours (some channels zeroed) ++ float32 depth in (0.2, 6) m."""
    rng = np.random.default_rng(seed)
    img = rng.integers(0, 256, size=(H, W, 3), dtype=np.uint8)
    img[rng.random((H, W, 3)) < zero_frac] = 0
    depth = (0.2 + 5.8 * rng.random((H, W))).astype(np.float32)
    return np.concatenate([img, depth[:, :, None]], axis=-1)     # float32, like the reference's rgbd_seq
