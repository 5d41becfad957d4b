"""NumPy restatement of the trajectory token <-> 6DoF pose path (TEST INFRASTRUCTURE).

Follows ``/root/reference/egoscaler/models/pointllm/utils/utils.py`` (cited per function)
and ``dataset.py:126-148`` (denorm).  Integer/bit-exact work: the CUDA kernels must equal
these functions bit for bit.  Pinned against the imported reference functions by
``oracle/make_golden.py`` -> ``tests/golden/traj_tokenizer.npz``.
"""
from __future__ import annotations

import re
from typing import Optional, Sequence

import numpy as np

# egoscaler/configs/camera.py:6-9 (the reference reads a non-existent ``focal_length``
# attribute at utils/utils.py:10; the value meant is ``focal_len``).
PINHOLE_IMAGE_SIZE = 1408
FOCAL_LEN = 605.343
PRINCIPAL_POINT = 703.5
# egoscaler/configs/dataset.py:1-6
DATASET_RANGE = dict(min_x=-2.0, max_x=2.0, min_y=-2.0, max_y=2.0, min_z=0.0, max_z=2.5)

TRAJ_TOKEN_BASE = 32006     # id of <p0>  (builder.py:33-46: 32000..32002 point, 32003..32005 ts/tsep/te)
TS_TOKEN, TSEP_TOKEN, TE_TOKEN = 32003, 32004, 32005


def bin_edges(num_bins: int = 256) -> np.ndarray:
    """utils/utils.py:14,19: ``np.linspace(-1, 1, num_bins)`` (float64)."""
    return np.linspace(-1, 1, num_bins)


def discretize_action(x, num_bins: int = 256) -> np.ndarray:
    """utils/utils.py:13-16: ``np.digitize(x, bins) - 1``.

    x < -1 -> -1 ; x >= 1 or NaN -> num_bins-1 ; a value on an edge goes to that edge's bin.
    """
    return np.digitize(np.asarray(x), bin_edges(num_bins)) - 1


def token_to_action(bins_idx, num_bins: int = 256) -> np.ndarray:
    """utils/utils.py:18-21: left bin edge (float64)."""
    return bin_edges(num_bins)[np.asarray(bins_idx)]


def rt2_scaler(traj: np.ndarray, maxmin: Sequence[float]) -> np.ndarray:
    """utils/utils.py:23-34 on a float32 [T,6] array (in place, like the reference)."""
    d_max, d_min = maxmin
    traj[:, [3, 4, 5]] = np.pi * traj[:, [3, 4, 5]]
    traj[:, 2] = (1.0 / 2) * traj[:, 2] + (1.0 / 2)
    traj[:, 2] = (d_max - d_min) * traj[:, 2] + d_min
    traj[:, 0] = (PINHOLE_IMAGE_SIZE / 2) * traj[:, 0] + (PINHOLE_IMAGE_SIZE / 2)
    traj[:, 0] = (traj[:, 0] - PRINCIPAL_POINT) * traj[:, 2] / FOCAL_LEN
    traj[:, 1] = (PINHOLE_IMAGE_SIZE / 2) * traj[:, 1] + (PINHOLE_IMAGE_SIZE / 2)
    traj[:, 1] = (traj[:, 1] - PRINCIPAL_POINT) * traj[:, 2] / FOCAL_LEN
    return traj


def simple_scaler(traj: np.ndarray, maxmin: Sequence[float]) -> np.ndarray:
    """utils/utils.py:36-45."""
    d_max, d_min = maxmin
    traj[:, [3, 4, 5]] = np.pi * (2 * (traj[:, [3, 4, 5]] / 100) - 1)
    traj[:, 2] = traj[:, 2] / 100
    traj[:, 2] = traj[:, 2] * (d_max - d_min) + d_min
    traj[:, 0] = (traj[:, 0] - PRINCIPAL_POINT) * traj[:, 2] / FOCAL_LEN
    traj[:, 1] = (traj[:, 1] - PRINCIPAL_POINT) * traj[:, 2] / FOCAL_LEN
    return traj


_RT2_PAT = re.compile(r'<p(\d+)> <p(\d+)> <p(\d+)> <p(\d+)> <p(\d+)> <p(\d+)>')


def str_to_float_rt2(s: str, maxmin: Optional[Sequence[float]], num_bins: int = 256,
                     scale: bool = True) -> Optional[np.ndarray]:
    """utils/utils.py:47-104, ``rt2=True`` full-6DoF branch: split on ``<tsep>``, a segment
    without a match repeats the previous pose (:91-93), float32 cast (:96), rt2 scaler (:98)."""
    traj, last = [], None
    for seg in s.split("<tsep>"):
        m = _RT2_PAT.search(seg)
        if m:
            last = tuple(token_to_action(list(map(int, m.groups())), num_bins))
            traj.append(last)
        elif last is not None:
            traj.append(last)
    if not traj:
        return None
    out = np.array(traj).astype(np.float32)
    return rt2_scaler(out, maxmin) if scale else out


def ids_to_poses(ids: Sequence[int], num_bins: int = 256, base: int = TRAJ_TOKEN_BASE,
                 eos: int = 2) -> Optional[np.ndarray]:
    """Token-id form of the post-processing at train.py:240-246 + utils.py:47-96: cut at EOS,
    keep only <p*>/<tsep> structure, each ``<tsep>``-delimited segment must hold 6 consecutive
    bin tokens (the regex's ``<pA> <pB> ...`` run), otherwise the previous pose is repeated.
    Returns float32 [T,6] un-scaled poses in [-1,1] (left bin edges) or None."""
    ids = list(ids)
    if eos in ids:
        ids = ids[:ids.index(eos)]
    segs, cur = [], []
    for t in ids:
        if t == TSEP_TOKEN:
            segs.append(cur)
            cur = []
        else:
            cur.append(t)
    segs.append(cur)
    traj, last = [], None
    for seg in segs:
        run, found = [], None
        for t in seg:                      # first run of 6 consecutive bin tokens (regex .search)
            if base <= t < base + num_bins:
                run.append(t - base)
                if len(run) == 6:
                    found = run
                    break
            else:
                run = []
        if found is not None:
            last = tuple(token_to_action(found, num_bins))
            traj.append(last)
        elif last is not None:
            traj.append(last)
    if not traj:
        return None
    return np.array(traj).astype(np.float32)


def denorm(traj: np.ndarray, do_norm: bool = True, max_abs: Optional[np.ndarray] = None,
           mean: Optional[np.ndarray] = None, std: Optional[np.ndarray] = None) -> np.ndarray:
    """dataset.py:126-148 on a float32 [B,T,6] array."""
    traj = np.array(traj, copy=True)
    r = DATASET_RANGE
    if do_norm:
        traj[:, :, [0, 1, 2]] = (traj[:, :, [0, 1, 2]] + 1) / 2
        traj[:, :, 0] = traj[:, :, 0] * (r["max_x"] - r["min_x"]) + r["min_x"]
        traj[:, :, 1] = traj[:, :, 1] * (r["max_y"] - r["min_y"]) + r["min_y"]
        traj[:, :, 2] = traj[:, :, 2] * (r["max_z"] - r["min_z"]) + r["min_z"]
        traj[:, :, [3, 4, 5]] *= np.pi
        return traj
    traj = traj * max_abs[:, None, :]
    return traj * std + mean


def pad_repeat_last(gen: np.ndarray, T: int) -> np.ndarray:
    """train.py:251-255: pad a short trajectory by repeating its last frame."""
    if gen.shape[0] < T:
        gen = np.concatenate([gen, np.repeat(gen[-1][None, :], T - gen.shape[0], axis=0)], axis=0)
    return gen


def poses_to_ids(poses: np.ndarray, num_bins: int = 256, base: int = TRAJ_TOKEN_BASE, eos: int = 2) -> list:
    """Forward tokenisation (not shipped by the reference): the sequence whose decode by
    ``str_to_float``/``ids_to_poses`` (utils/utils.py:47-104) gives back ``token_to_action(discretize_action(poses))``:
    ``<ts> (<pA> <pB> <pC> <pD> <pE> <pF> <tsep>) x T <te> eos`` with bins from ``discretize_action`` (:13-16)."""
    ids = [TS_TOKEN]
    for row in np.asarray(poses, np.float32):
        ids += (discretize_action(row, num_bins) + base).tolist() + [TSEP_TOKEN]
    return ids + [TE_TOKEN, eos]


def norm_poses(traj: np.ndarray) -> np.ndarray:
    """Inverse of ``denorm(do_norm=True)`` (dataset.py:139-146): physical [.., 6] -> [-1, 1], float32."""
    t = np.array(traj, np.float64, copy=True)
    r = DATASET_RANGE
    for j, ax in enumerate("xyz"):
        t[..., j] = 2.0 * (t[..., j] - r["min_" + ax]) / (r["max_" + ax] - r["min_" + ax]) - 1.0
    t[..., 3:] = t[..., 3:] / np.pi
    return t.astype(np.float32)


# ---------------------------------------------------------------------------------------------- pre-processing (N3)
def smoothing_traj(traj):
    """egoscaler/models/utils/traj_utils.py:42-96 restated with slices: moving average over columns 0..2 with the reference's
    edge rules (branch order j==0, j==1, j==T-2, j==T-1, general); summation left to right in the input dtype."""
    traj = np.asarray(traj)
    p = traj[:, :3]
    n = p.shape[0]
    q = p.copy()
    if n >= 5:
        q[2:n - 2] = (p[0:n - 4] + p[1:n - 3] + p[2:n - 2] + p[3:n - 1] + p[4:n]) / 5
    if n >= 3:
        q[0] = (3 * p[0] + p[1] + p[2]) / 5
    elif n == 2:
        q[0] = (3 * p[0] + p[1]) / 4
    if n >= 4:
        q[1] = (2 * p[0] + p[1] + p[2] + p[3]) / 5
    elif n == 3:
        q[1] = (2 * p[0] + p[1] + p[2]) / 4
    if n >= 4:                                      # j == n-2 (and j >= 2)
        q[n - 2] = (p[n - 4] + p[n - 3] + p[n - 2] + p[n - 1]) / 4
    if n >= 3:                                      # j == n-1 (and j >= 2)
        q[n - 1] = (p[n - 3] + p[n - 2] + p[n - 1]) / 3
    return np.concatenate([q, traj[:, 3:]], axis=-1)


def preprocess_traj(traj, num_steps):
    """traj_utils.py:3-40: -> (resampled [num_steps, D], padding mask [num_steps] int)."""
    traj = np.asarray(traj)
    T = traj.shape[0]
    if T >= num_steps:
        idx = np.linspace(0, T - 1, num_steps).astype(int)
        return traj[idx], np.ones(num_steps, dtype=int)
    pad = num_steps - T
    return np.vstack([traj, np.repeat(traj[-1:], pad, axis=0)]), np.concatenate([np.ones(T, dtype=int), np.zeros(pad, dtype=int)])
