import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


GOLDEN = os.path.join(ROOT, "tests", "golden")


def apply_trained_delta(sd, g):
    """State of tests/golden/traj_vlm_hd128_trained.npz: det_state_dict weights with the over-fitted rows
    (oracle/make_golden.py golden_trained) written over them.  ``sd``: name -> tensor, modified in place."""
    import torch
    lo = int(g["rows_lo"])
    for key, name in (("embed_rows", "model.embed_tokens.weight"), ("lm_head_rows", "lm_head.weight")):
        rows = torch.from_numpy(g[key])
        sd[name] = sd[name].clone()
        sd[name][lo:lo + rows.shape[0]] = rows.to(sd[name].dtype)
    for k in g.files:
        if k.startswith("norm/"):
            sd[k[5:]] = torch.from_numpy(g[k]).to(sd[k[5:]].dtype)
    return sd
