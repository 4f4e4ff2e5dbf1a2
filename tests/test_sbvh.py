"""Trees built WITH spatial splits (build<index_bvh<P>>(first, last, true),
detail/bvh/sah.h:430-615): a primitive is referenced from several leaves, so indices[] is longer than
the primitive list and a ray can meet the same primitive twice.  The fixture is such a tree from the
reference's own builder with the reference's hit records on it (tests/golden/make_golden.py sbvh).

closest_hit / any_hit: duplicates are exact t-ties of the SAME primitive, so the records are those of
any other tree (bit-identical on the reference-order walk).  multi_hit: the reference reports the
references it happens to visit (a primitive may appear twice); the reference-order walk reproduces
that bit for bit, the BVH8 walk reports every primitive once."""
import os
import sys

import numpy as np
import pytest

from conftest import classify_closest
from hostsim_bindings import HostWide, multi_hit_exact, query_exact
from oracle import bindings as O
from visionaray_b200 import scenes as S
from visionaray_b200._native import NODE_DTYPE

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as MG  # noqa: E402  (the scene / ray generators of the fixture)

FLT_MAX = np.float32(3.4028234663852886e38)


@pytest.fixture(scope="module")
def sbvh():
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_sbvh.npz"))
    tris = MG.sbvh_scene_prims()
    sc = S.HostScene(0, tris, g["nodes"].view(NODE_DTYPE).reshape(-1), g["indices"], S.face_normals(tris), "sbvh")
    rays = MG.sbvh_rays()
    gold = {k: g[k].view(O.HIT_DTYPE) for k in ("closest", "any")}
    gold["multi16"] = g["multi16"].view(O.HIT_DTYPE).reshape(-1, 16)
    return sc, rays, gold


def unique_records(rec):
    """Drop repeated (prim_id, t) records of each ray, keep order, pad with misses."""
    out = np.zeros_like(rec)
    out["t"] = FLT_MAX
    for i in range(rec.shape[0]):
        seen, k = set(), 0
        for r in rec[i]:
            if not r["hit"]:
                break
            key = (int(r["prim_id"]), float(r["t"]))
            if key in seen:
                continue
            seen.add(key)
            out[i, k] = r
            k += 1
    return out


def test_fixture_is_a_split_tree(sbvh):
    sc, rays, gold = sbvh
    assert sc.indices.shape[0] > 2 * sc.n_prims and np.unique(sc.indices).shape[0] == sc.n_prims
    m = gold["multi16"]
    dup = sum(len({(int(r["prim_id"])) for r in row[row["hit"] != 0]}) < int((row["hit"] != 0).sum()) for row in m)
    assert dup > 10          # the reference does report duplicates on this tree


def test_oracle_on_split_tree(sbvh):
    sc, rays, gold = sbvh
    assert O.oracle_query(sc, rays, False).tobytes() == gold["closest"].tobytes()
    assert O.oracle_query(sc, rays, True, 0.5).tobytes() == gold["any"].tobytes()
    assert O.oracle_multi_hit(sc, rays[:1500], 16).tobytes() == gold["multi16"].tobytes()


def test_host_instantiation_on_split_tree(sbvh):
    """csrc/bvh2_exact.cuh, bvh8*.cuh, multi_hit.cuh compiled for the host (tests/hostsim)."""
    sc, rays, gold = sbvh
    assert query_exact(sc, rays).tobytes() == gold["closest"].tobytes()
    assert multi_hit_exact(sc, rays[:1500], 16).tobytes() == gold["multi16"].tobytes()
    w = HostWide(sc)
    assert w.stats()["wide_prims"] >= sc.indices.shape[0]
    assert classify_closest(gold["closest"], w.query(rays))["bad"] == 0
    assert (w.query(rays, any_hit=True, max_t=0.5)["hit"] == gold["any"]["hit"]).all()
    check_unique_multi(gold["multi16"], w.multi_hit(rays[:1500], 16))
    w.close()


def test_host_builder_with_spatial_splits_reproduces_the_fixture(sbvh):
    """build<index_bvh<P>>(prims, n, true) of include/visionaray_b200/bvh.h == the reference builder's
    tree, byte for byte (nodes and indices)."""
    sc, rays, gold = sbvh
    nodes, idx = S.build_bvh(0, sc.prims, spatial_splits=True)
    assert nodes.tobytes() == sc.nodes.tobytes() and idx.tobytes() == sc.indices.tobytes()


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref is only built where /root/reference exists")
@pytest.mark.parametrize("kind,gen", [
    (0, lambda: S.gen_triangles(5, 4000, 0, 31, 0.1)),
    (0, lambda: S.gen_triangles(5, 1500, 0, 33, 1.0)),
    (0, lambda: S.gen_triangles(3, 300, 0, 3)),
    (0, lambda: S.gen_triangles(2, 60, 0, 2, 0.15)),
    (1, lambda: S.gen_spheres(3000, 7, 0.02, 0.2)),
    (1, lambda: S.gen_spheres(2000, 9, 0.005, 0.5)),
])
def test_host_builder_with_spatial_splits_matches_reference_builder(kind, gen):
    prims = gen()
    nodes, idx = S.build_bvh(kind, prims, spatial_splits=True)
    rn, ri = O.ref_build(kind, prims, True)
    assert nodes.tobytes() == rn.tobytes() and idx.tobytes() == ri.tobytes()


def check_unique_multi(ref, got):
    want = unique_records(ref)
    full = (ref["hit"] != 0).all(axis=1)       # the reference's window was full: a duplicate may have pushed a hit out
    assert full.sum() < ref.shape[0] // 4
    ok = ~full
    assert ((want["hit"] != 0).sum(axis=1)[ok] == (got["hit"] != 0).sum(axis=1)[ok]).all()
    assert (want["t"][ok] == got["t"][ok]).all()
    # same primitives per ray (order among exact ties is free)
    assert (np.sort(want["prim_id"][ok], axis=1) == np.sort(got["prim_id"][ok], axis=1)).all()
    for row in got:
        h = row[row["hit"] != 0]
        assert len({(int(r["prim_id"]), float(r["t"])) for r in h}) == len(h)


@pytest.mark.gpu
def test_gpu_split_tree(gpu, sbvh):
    sc, rays, gold = sbvh
    dev = gpu.Scene(sc)
    try:
        assert dev.stats()["wide_prims"] >= sc.indices.shape[0]
        assert gpu.closest_hit(dev, rays, traverse=gpu.TRAVERSE_EXACT).tobytes() == gold["closest"].tobytes()
        assert gpu.any_hit(dev, rays, max_t=0.5, traverse=gpu.TRAVERSE_EXACT).tobytes() == gold["any"].tobytes()
        assert classify_closest(gold["closest"], gpu.closest_hit(dev, rays))["bad"] == 0
        assert (gpu.any_hit(dev, rays, max_t=0.5)["hit"] == gold["any"]["hit"]).all()
        assert gpu.multi_hit(dev, rays[:1500], 16, traverse=gpu.TRAVERSE_EXACT).tobytes() == gold["multi16"].tobytes()
        check_unique_multi(gold["multi16"], gpu.multi_hit(dev, rays[:1500], 16))
        # a frame over the split tree: reference-order walk reproduces the oracle pixel for pixel
        fp = S.frame_params(1, 128, 96, (15.0, 0.5, 2.5), (15.0, 0.5, 0.5), num_bounces=3)
        mats, light = S.default_material(0), S.point_light((15.0, 0.5, 2.5))
        want = O.oracle_frame(sc, fp, mats, light)
        img = np.zeros((96, 128, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, mats, light, img, traverse=gpu.TRAVERSE_EXACT)
        assert float((np.abs(img - want).max(axis=2) > 1 / 255).mean()) == 0.0
        gpu.frame(dev, fp, None, mats, light, img, traverse=gpu.TRAVERSE_WIDE)
        assert float((np.abs(img - want).max(axis=2) > 1 / 255).mean()) <= 2e-3
    finally:
        dev.close()
