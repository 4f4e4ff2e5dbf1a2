"""multi_hit<N>() -- SURVEY section 8f row 2 (detail/traverse_linear.inl:325-380, detail/multi_hit.h,
detail/bvh/intersect.inl:35-134 with MultiHit).  CPU tier: the oracle restatement is pinned
bit-for-bit against the unmodified reference headers (where oracle/_ref exists) and against golden
vectors generated from them; the product's two traversals (csrc/multi_hit.cuh, instantiated on the
host by tests/hostsim) are checked against the oracle.  GPU tier: the same through the C ABI."""
import os

import numpy as np
import pytest

from hostsim_bindings import HostWide, multi_hit_exact
from oracle import bindings as O
from visionaray_b200 import scenes as S
from visionaray_b200._native import NODE_DTYPE

GOLD = os.path.join(os.path.dirname(__file__), "golden")
FLT_MAX = np.float32(3.4028234663852886e38)


def dense_soup(n=4000, seed=31, size=0.25):
    """Large overlapping triangles: tens of intersections per ray, so all N records fill up."""
    tris = S.gen_triangles(5, n, 0, seed, size)
    nodes, idx = S.build_bvh(0, tris)
    return S.HostScene(0, tris, nodes, idx, None, "dense_soup")


def check_records(want, got, rel_tie=1e-6):
    """North-star rule applied per record: same number of hits per ray, t's equal as sorted sequences;
    a record whose ids differ must be a t-tie (same t within rel_tie); same id -> t/u/v bit-identical.
    Returns the number of tie-swapped records."""
    assert want.shape == got.shape
    assert ((want["hit"] != 0).sum(axis=1) == (got["hit"] != 0).sum(axis=1)).all()
    assert (np.diff(got["t"].astype(np.float64), axis=1) >= 0).all(), "records sorted by t"
    assert (got["t"][got["hit"] == 0] == FLT_MAX).all()
    hit = want["hit"] != 0
    same = hit & (want["prim_id"] == got["prim_id"])
    for k in ("t", "u", "v", "geom_id"):
        assert (want[k][same] == got[k][same]).all(), k
    diff = hit & ~same
    ta, tb = want["t"][diff].astype(np.float64), got["t"][diff].astype(np.float64)
    assert (np.abs(ta - tb) <= rel_tie * np.abs(ta)).all(), "id mismatch that is not a t-tie"
    return int(diff.sum())


SCENES = [("cfg1_sphere", 0.05), ("cfg2_grid", 0.01), ("cfg5_soup", 0.0005), ("spheres", 0.03)]


def get_scene(name, scale):
    return dense_soup() if name == "dense_soup" else S.make_scene(name, scale)


@pytest.mark.parametrize("name,scale", SCENES + [("dense_soup", 1.0)])
def test_oracle_multi_hit_is_the_reference(name, scale):
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here): pinned by the golden vectors instead")
    sc = get_scene(name, scale)
    lo, hi = sc.bounds()
    rays = S.random_rays(6000, 61, float(lo.min()) - 0.2, float(hi.max()) + 0.2)
    for n in (1, 2, 5, 16):
        a, b = O.oracle_multi_hit(sc, rays, n), O.ref_multi_hit(sc, rays, n)
        assert a.tobytes() == b.tobytes(), (name, n)
    if name == "dense_soup":
        assert (a["hit"].sum(axis=1) == 16).mean() > 0.1          # the bound really is the 16th record


def test_oracle_multi_hit_over_two_bvhs_is_the_reference():
    if not O.ref_available():
        pytest.skip("oracle/_ref not built")
    a, b = dense_soup(1500, 32), dense_soup(1500, 33)
    rays = S.random_rays(4000, 62, -0.1, 1.1)
    for n in (4, 8):
        assert O.oracle_multi_hit([a, b], rays, n).tobytes() == O.ref_multi_hit([a, b], rays, n).tobytes()


def test_multi_hit_golden_vectors():
    g = np.load(os.path.join(GOLD, "golden_queries.npz"))
    m = np.load(os.path.join(GOLD, "golden_multi_hit.npz"))
    for key, kind in (("tri", 0), ("sph", 1)):
        sc = S.HostScene(kind, g[f"{key}_prims"], g[f"{key}_nodes"].view(NODE_DTYPE).reshape(-1), g[f"{key}_indices"], None, key)
        rays = g[f"{key}_rays"][:600]
        for n in (4, 16):
            want = m[f"{key}_multi{n}"].view(O.HIT_DTYPE).reshape(600, n)
            assert O.oracle_multi_hit(sc, rays, n).tobytes() == want.tobytes()
            # first record == closest hit of the older fixture
            c = g[f"{key}_closest"].view(O.HIT_DTYPE).reshape(-1)[:600]
            assert (want[:, 0]["t"] == c["t"]).all() and (want[:, 0]["hit"] == c["hit"]).all()


@pytest.mark.parametrize("name,scale", SCENES + [("dense_soup", 1.0)])
def test_product_multi_hit_on_the_host(name, scale):
    sc = get_scene(name, scale)
    lo, hi = sc.bounds()
    rays = S.random_rays(3000, 63, float(lo.min()) - 0.2, float(hi.max()) + 0.2)
    w = HostWide(sc)
    for n in (1, 4, 16):
        want = O.oracle_multi_hit(sc, rays, n)
        assert multi_hit_exact(sc, rays, n).tobytes() == want.tobytes()       # reference order: bit-identical
        check_records(want, w.multi_hit(rays, n))
    w.close()


# ---- GPU tier ------------------------------------------------------------------------------------

@pytest.mark.gpu
@pytest.mark.parametrize("name,scale", [("cfg1_sphere", 0.2), ("cfg2_grid", 0.05), ("cfg5_soup", 0.002), ("spheres", 0.2), ("dense_soup", 1.0)])
def test_gpu_multi_hit_parity(gpu, name, scale):
    sc = get_scene(name, scale)
    dev = gpu.Scene(sc)
    lo, hi = sc.bounds()
    rays = S.random_rays(60_000, 64, float(lo.min()) - 0.2, float(hi.max()) + 0.2)
    for n in (1, 3, 16):
        want = O.oracle_multi_hit(sc, rays, n)
        assert gpu.multi_hit(dev, rays, n, traverse=gpu.TRAVERSE_EXACT).tobytes() == want.tobytes()
        check_records(want, gpu.multi_hit(dev, rays, n))
    # record 0 of multi_hit is closest_hit
    c = gpu.closest_hit(dev, rays)
    m = gpu.multi_hit(dev, rays, 4)
    assert (c["hit"] == m[:, 0]["hit"]).all() and (c["t"] == m[:, 0]["t"]).all()
    dev.close()


@pytest.mark.gpu
def test_gpu_multi_hit_golden_two_bvhs_device_buffers_and_errors(gpu):
    import torch

    g = np.load(os.path.join(GOLD, "golden_queries.npz"))
    m = np.load(os.path.join(GOLD, "golden_multi_hit.npz"))
    sc = S.HostScene(0, g["tri_prims"], g["tri_nodes"].view(NODE_DTYPE).reshape(-1), g["tri_indices"], None, "tri")
    dev = gpu.Scene(sc)
    rays = g["tri_rays"][:600]
    for n in (4, 16):
        want = m[f"tri_multi{n}"].view(O.HIT_DTYPE).reshape(600, n)
        assert gpu.multi_hit(dev, rays, n, traverse=gpu.TRAVERSE_EXACT).tobytes() == want.tobytes()
        check_records(want, gpu.multi_hit(dev, rays, n))
    # device buffers == host buffers
    d = gpu.multi_hit(dev, torch.from_numpy(rays).cuda(), 16)
    gpu.sync()
    assert d.cpu().numpy().tobytes() == gpu.multi_hit(dev, rays, 16).tobytes()
    # a range of two bvhs: merged by t
    a, b = dense_soup(1500, 32), dense_soup(1500, 33)
    da, db = gpu.Scene(a), gpu.Scene(b)
    rr = S.random_rays(20_000, 65, -0.1, 1.1)
    want = O.oracle_multi_hit([a, b], rr, 8)
    check_records(want, gpu.multi_hit([da, db], rr, 8, traverse=gpu.TRAVERSE_EXACT))
    check_records(want, gpu.multi_hit([da, db], rr, 8))
    # the GPU-built tree returns the same records
    built = gpu.Scene.build(0, a.prims)
    check_records(O.oracle_multi_hit(a, rr, 16), gpu.multi_hit(built, rr, 16))
    for bad in (0, 17):
        with pytest.raises(RuntimeError):
            gpu.multi_hit(dev, rays, bad)
    for s in (dev, da, db, built):
        s.close()
