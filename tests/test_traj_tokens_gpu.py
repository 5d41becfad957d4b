"""GPU: trajectory quantize / dequantize / scalers / ids->poses, bit-exact against the oracle and the goldens
generated from the real reference (utils/utils.py, dataset.py:126-148)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from egoscaler_b200 import ops
from oracle import traj_tokenizer as tt
from conftest import GOLDEN


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(GOLDEN, "traj_tokenizer.npz"))


def _cu(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def test_quantize_golden(g):
    assert np.array_equal(ops.traj_quantize(_cu(g["x32"])).cpu().numpy(), g["q32"])
    assert np.array_equal(ops.traj_quantize(_cu(g["x64"])).cpu().numpy(), g["q64"])
    assert np.array_equal(ops.traj_quantize(_cu(g["x32"]), num_bins=64).cpu().numpy(), g["q32_64bins"])
    assert np.array_equal(ops.traj_quantize(_cu(g["demo_norm"])).cpu().numpy(), g["demo_q"])


def test_quantize_large_random_vs_oracle():
    rng = np.random.default_rng(3)
    n = (1 << 22) + 3                        # ragged tail
    x = rng.uniform(-1.05, 1.05, n).astype(np.float32)
    edges = np.linspace(-1, 1, 256).astype(np.float32)
    x[:256] = edges
    x[256:512] = np.nextafter(edges, np.float32(2))
    x[512:768] = np.nextafter(edges, np.float32(-2))
    x[768:772] = [np.nan, np.inf, -np.inf, -0.0]
    q = ops.traj_quantize(_cu(x), id_offset=32006).cpu().numpy()
    assert np.array_equal(q, tt.discretize_action(x) + 32006)
    assert ops.traj_quantize(torch.empty(0, device="cuda")).numel() == 0


def test_dequantize_and_roundtrip(g):
    ids = torch.arange(256 * 6, device="cuda").reshape(-1, 6) % 256
    d = ops.traj_dequantize(ids).cpu().numpy()
    assert np.array_equal(d, g["deq256"].astype(np.float32)[ids.cpu().numpy()])
    # round trip through float32 edges: the reference is NOT idempotent there (float32(edge) can fall below the
    # float64 edge) -- the kernel must reproduce exactly what np.digitize does with those float32 values ...
    rt = ops.traj_quantize(ops.traj_dequantize(ids)).cpu().numpy()
    assert np.array_equal(rt, tt.discretize_action(d))
    # ... while float64 edges round-trip exactly
    e64 = torch.from_numpy(g["deq256"]).cuda()
    assert torch.equal(ops.traj_quantize(e64), torch.arange(256, device="cuda"))
    d64 = ops.traj_dequantize(torch.arange(60, device="cuda").reshape(-1, 6), num_bins=64).cpu().numpy()
    assert np.array_equal(d64.ravel(), g["deq64"].astype(np.float32)[:60])
    bad = ops.traj_dequantize(torch.full((1, 6), 999, device="cuda"))
    assert torch.isnan(bad).all()


def test_scalers_golden(g):
    dmax, dmin = g["s2f_maxmin"].tolist()
    out = ops.traj_scale_(_cu(g["scaler_in"]).clone(), ops.SCALE_RT2, dmax, dmin).cpu().numpy()
    assert np.array_equal(out, g["rt2_out"])
    out = ops.traj_scale_(_cu(g["simple_in"]).clone(), ops.SCALE_SIMPLE, dmax, dmin).cpu().numpy()
    assert np.array_equal(out, g["simple_out"])
    out = ops.traj_scale_(_cu(g["denorm_in"]).clone(), ops.SCALE_DENORM).cpu().numpy()
    assert np.array_equal(out, g["denorm_norm_out"])


def _ids_from_string(s):
    """Token-id form of a reference-format string (only <p*>, <tsep>; everything else -> a text token)."""
    import re
    ids = []
    for tok in re.findall(r"<p\d+>|<tsep>|<ts>|<te>|[^<\s]+|<", s):
        if tok == "<tsep>":
            ids.append(tt.TSEP_TOKEN)
        elif tok == "<ts>":
            ids.append(tt.TS_TOKEN)
        elif tok == "<te>":
            ids.append(tt.TE_TOKEN)
        elif tok.startswith("<p"):
            ids.append(tt.TRAJ_TOKEN_BASE + int(tok[2:-1]))
        else:
            ids.append(100)
    return ids


def test_ids_to_poses_golden_strings(g):
    dmax, dmin = g["s2f_maxmin"].tolist()
    strs = g["s2f_strings"].tolist()
    rows = [_ids_from_string(s) + [2] for s in strs]
    T = max(len(r) for r in rows)
    ids = torch.tensor([r + [0] * (T - len(r)) for r in rows], device="cuda")
    poses, cnt = ops.traj_ids_to_poses(ids, max_steps=32, scale=ops.SCALE_RT2, d_max=dmax, d_min=dmin, pad_repeat=False)
    for i in range(len(strs)):
        ref = g[f"s2f_{i}"]
        assert int(cnt[i]) == ref.shape[0]
        assert np.array_equal(poses[i, :ref.shape[0]].cpu().numpy(), ref)
        o = tt.ids_to_poses(rows[i])
        assert (o is None and ref.shape[0] == 0) or o.shape[0] == ref.shape[0]


def test_ids_to_poses_random_vs_oracle():
    rng = np.random.default_rng(11)
    B, T = 64, 170
    ids = rng.integers(32000, 32262, (B, T))
    ids[rng.random((B, T)) < 0.15] = tt.TSEP_TOKEN
    ids[rng.random((B, T)) < 0.01] = 2
    ids[0] = 5                                     # nothing parsable
    poses, cnt = ops.traj_ids_to_poses(_cu(ids), max_steps=40, pad_repeat=True)
    poses, cnt = poses.cpu().numpy(), cnt.cpu().numpy()
    for b in range(B):
        ref = tt.ids_to_poses(ids[b].tolist())
        if ref is None:
            assert cnt[b] == 0
            continue
        n = min(ref.shape[0], 40)
        assert cnt[b] == n
        assert np.array_equal(poses[b, :n], ref[:n])
        assert np.array_equal(poses[b], tt.pad_repeat_last(ref[:n], 40))


def test_poses_to_ids_roundtrip_and_oracle():
    """Data-side tokenisation: ids == oracle layout built from the pinned discretize_action; decoding them with
    the (reference-pinned) parser gives token_to_action(discretize_action(poses)) -- the quantiser round trip -- bit-exactly;
    ragged lengths, exact bin edges, out-of-range values and the physical -> normalised mapping (inverse of denorm)."""
    import numpy as np
    import torch
    from egoscaler_b200 import ops
    from oracle import traj_tokenizer as tt
    rng = np.random.default_rng(4)
    B, T = 37, 20
    poses = rng.uniform(-1.02, 1.02, (B, T, 6)).astype(np.float32)
    poses[0, 0] = np.linspace(-1, 1, 256)[[0, 1, 127, 128, 254, 255]].astype(np.float32)     # exact edges
    poses[1, 0] = [np.nan, np.inf, -np.inf, -1.0, 1.0, 0.0]
    lens = rng.integers(0, T + 1, B).astype(np.int32)
    lens[:3] = [T, T, 0]
    ids, olen = ops.traj_poses_to_ids(torch.from_numpy(poses).cuda(), torch.from_numpy(lens).cuda())
    ids, olen = ids.cpu().numpy(), olen.cpu().numpy()
    assert ids.shape == (B, 3 + 7 * T)
    for b in range(B):
        ref = tt.poses_to_ids(poses[b, :lens[b]])
        assert olen[b] == len(ref)
        assert np.array_equal(ids[b, :len(ref)], np.array(ref)) and (ids[b, len(ref):] == 0).all()
    # round trip through the parser kernel and through the oracle parser
    back, cnt = ops.traj_ids_to_poses(torch.from_numpy(ids).cuda(), max_steps=T, pad_repeat=False)
    back, cnt = back.cpu().numpy(), cnt.cpu().numpy()
    for b in range(B):
        if b == 1:
            continue                                               # NaN / inf row: bins -1 / 255 are checked above via the oracle ids
        n = int(lens[b])
        # the reference parser splits on <tsep>: the trailing "<te>" segment has no pose and repeats the previous one
        # (utils/utils.py:91-93), so a well-formed sequence of n poses parses to n + 1 (capped by max_steps)
        ref_parse = tt.ids_to_poses(ids[b].tolist())
        in_range = n > 0 and bool((tt.discretize_action(poses[b, :n]) >= 0).all())
        if in_range:
            assert cnt[b] == min(n + 1, T)
        if ref_parse is None:                                  # x < -1 quantises to bin -1 = id of <te>: such a pose does not parse
            assert cnt[b] == 0
        else:
            assert cnt[b] == min(len(ref_parse), T) and np.array_equal(back[b, :cnt[b]], ref_parse[:cnt[b]])
        q = tt.token_to_action(tt.discretize_action(poses[b, :n]), 256).astype(np.float32)
        ok = (tt.discretize_action(poses[b, :n]) >= 0).all(axis=1) if n else np.zeros(0, bool)
        if n and ok.all():
            assert np.array_equal(back[b, :n], q)
    # physical units -> normalised -> ids  ==  oracle normaliser + tokeniser
    phys = np.stack([rng.uniform(-2, 2, (B, T)), rng.uniform(-2, 2, (B, T)), rng.uniform(0, 2.5, (B, T)),
                     rng.uniform(-3.1, 3.1, (B, T)), rng.uniform(-3.1, 3.1, (B, T)), rng.uniform(-3.1, 3.1, (B, T))], -1).astype(np.float32)
    r = tt.DATASET_RANGE
    ranges = torch.tensor([r["min_x"], r["max_x"], r["min_y"], r["max_y"], r["min_z"], r["max_z"]], dtype=torch.float64).cuda()
    ids2, _ = ops.traj_poses_to_ids(torch.from_numpy(phys).cuda(), None, ranges=ranges)
    ids2 = ids2.cpu().numpy()
    nrm = tt.norm_poses(phys)
    mism = 0
    for b in range(B):
        ref = np.array(tt.poses_to_ids(nrm[b]))
        mism += int((ids2[b, :len(ref)] != ref).sum())
    assert mism == 0
    # and denorm(dequantised) lands within one bin of the physical input
    back2, _ = ops.traj_ids_to_poses(torch.from_numpy(ids2).cuda(), max_steps=T, scale=ops.SCALE_DENORM, pad_repeat=False)
    err = np.abs(back2.cpu().numpy() - phys)
    assert err[..., :2].max() <= 4.0 / 255 + 1e-5 and err[..., 2].max() <= 2.5 / 255 + 1e-5 and err[..., 3:].max() <= 2 * np.pi / 255 + 1e-5


def test_traj_preprocess_matches_reference_golden():
    """csrc/traj_tokens.cu traj_preprocess vs outputs of the real traj_utils.py (smoothing_traj then preprocess_traj) for
    every edge length: bit-exact, incl. the padding mask."""
    import os
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "traj_preprocess.npz"))
    traj = torch.from_numpy(g["traj"]).cuda()
    lens = torch.from_numpy(g["lens"]).cuda()
    for ns in (20, 7):
        for smooth in (0, 1):
            out, mask = ops.traj_preprocess(traj, lens, ns, smooth=bool(smooth))
            assert np.array_equal(out.cpu().numpy(), g[f"out/{ns}/{smooth}"]), (ns, smooth)
            assert np.array_equal(mask.cpu().numpy(), g[f"mask/{ns}/{smooth}"])
    # lens=None: every row uses all T steps
    out, mask = ops.traj_preprocess(traj[:3], None, 20, smooth=True)
    from oracle import traj_tokenizer as ot
    for b in range(3):
        r, m = ot.preprocess_traj(ot.smoothing_traj(g["traj"][b]), 20)
        assert np.array_equal(out[b].cpu().numpy(), r) and np.array_equal(mask[b].cpu().numpy(), m)
