import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _ensure_built():
    """Host-side libraries (oracle restatement, host layer, hostsim) are built on demand; the CUDA
    library is only ever built by __graft_entry__.build()."""
    import __graft_entry__ as g

    g.build_host_only()


@pytest.fixture(scope="session", autouse=True)
def built():
    _ensure_built()


@pytest.fixture(scope="session")
def gpu():
    import torch

    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from visionaray_b200 import api

    api.init(0)
    return api


def tri(v1, v2, v3, prim_id, geom_id=0):
    """64-byte basic_triangle<3,float> record as 16 float32 words."""
    t = np.zeros(16, dtype=np.float32)
    ids = t.view(np.uint32)
    ids[0], ids[1] = geom_id, prim_id
    t[4:7] = v1
    t[8:11] = np.asarray(v2, dtype=np.float32) - np.asarray(v1, dtype=np.float32)
    t[12:15] = np.asarray(v3, dtype=np.float32) - np.asarray(v1, dtype=np.float32)
    return t


def classify_closest(oracle_hits, got_hits, rel_tie=1e-6, rel_tuv=1e-5):
    """Parity rule of BASELINE.json's north_star for closest-hit records:
    hit flags equal; prim ids bit-exact except t-ties within rel_tie; t/u/v within rel_tuv.
    Returns dict(n, id_mismatch, ties, bad) -- `bad` must be 0."""
    a, b = oracle_hits, got_hits
    n = a.shape[0]
    flag_bad = int((a["hit"] != b["hit"]).sum())
    both = (a["hit"] != 0) & (b["hit"] != 0)
    id_diff = both & ((a["prim_id"] != b["prim_id"]) | (a["geom_id"] != b["geom_id"]))
    ta, tb = a["t"].astype(np.float64), b["t"].astype(np.float64)
    tie = id_diff & (np.abs(ta - tb) <= rel_tie * np.maximum(np.abs(ta), 1e-30))
    same = both & ~id_diff
    def rel_bad(x, y, tol):
        x = x.astype(np.float64); y = y.astype(np.float64)
        return np.abs(x - y) > tol * np.maximum(np.maximum(np.abs(x), np.abs(y)), 1e-30)
    tuv_bad = same & (rel_bad(a["t"], b["t"], rel_tuv) | rel_bad(a["u"], b["u"], rel_tuv) | rel_bad(a["v"], b["v"], rel_tuv))
    bad = flag_bad + int((id_diff & ~tie).sum()) + int(tuv_bad.sum())
    return dict(n=n, flag_bad=flag_bad, id_mismatch=int(id_diff.sum()), ties=int(tie.sum()), tuv_bad=int(tuv_bad.sum()), bad=bad)
