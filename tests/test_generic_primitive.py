"""generic_primitive<basic_triangle, basic_sphere> -- SURVEY section 8f row 2, second half: one tree over
triangles AND spheres, leaves mix both (generic_primitive.h, detail/generic_primitive.inl, variant.h).
The fixture (tests/golden/make_golden.py generic) holds records laid out by the reference's own
constructors, the reference builder's tree and the reference's closest_hit / any_hit / multi_hit<8>
records.  CPU: record layout, host builder, oracle, CPU instantiation of the device headers.
GPU: the same through the C ABI, plus the GPU builder."""
import os

import numpy as np
import pytest

from conftest import classify_closest
from hostsim_bindings import HostWide, multi_hit_exact, query_exact
from oracle import bindings as O
from visionaray_b200 import scenes as S
from visionaray_b200._native import NODE_DTYPE

TRI_WORDS = [0, 1, 4, 5, 6, 8, 9, 10, 12, 13, 14, 16]     # ids, v1, e1, e2, which
SPH_WORDS = [0, 1, 4, 5, 6, 8, 16]                          # ids, center, radius, which


@pytest.fixture(scope="module")
def gen():
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_generic.npz"))
    sc = S.make_scene("generic", 0.05)
    ref = S.HostScene(2, sc.prims, g["nodes"].view(NODE_DTYPE).reshape(-1), g["indices"], None, "generic")
    gold = {k: g[k].view(O.HIT_DTYPE) for k in ("closest", "any")}
    gold["multi8"] = g["multi8"].view(O.HIT_DTYPE).reshape(-1, 8)
    return ref, g["rays"], gold, g["ref_records"]


def test_record_layout_is_the_reference_variant(gen):
    """80 bytes: the alternative in place, its 1-based index at +64 -- as the reference's constructors lay it out."""
    sc, rays, gold, rec = gen
    mine = sc.prims.view(np.uint32)
    assert mine.shape == rec.shape == (sc.n_prims, 20)
    tri = mine[:, 16] == 1
    assert tri.sum() == 3000 and (mine[~tri, 16] == 2).all()
    assert (mine[tri][:, TRI_WORDS] == rec[tri][:, TRI_WORDS]).all()
    assert (mine[~tri][:, SPH_WORDS] == rec[~tri][:, SPH_WORDS]).all()


def test_host_builder_reproduces_the_reference_tree(gen):
    sc, rays, gold, rec = gen
    nodes, idx = S.build_bvh(2, sc.prims)
    assert nodes.tobytes() == sc.nodes.tobytes() and idx.tobytes() == sc.indices.tobytes()
    if O.ref_available():
        n2, i2 = S.build_bvh(2, sc.prims, spatial_splits=True)
        r2, j2 = O.ref_build(2, sc.prims, True)
        assert n2.tobytes() == r2.tobytes() and i2.tobytes() == j2.tobytes() and i2.shape[0] > sc.n_prims


def test_oracle_on_mixed_leaves(gen):
    sc, rays, gold, rec = gen
    want = gold["closest"]
    assert O.oracle_query(sc, rays, False).tobytes() == want.tobytes()
    assert O.oracle_query(sc, rays, True, 0.3).tobytes() == gold["any"].tobytes()
    assert O.oracle_multi_hit(sc, rays[:1500], 8).tobytes() == gold["multi8"].tobytes()
    which = sc.prims.view(np.uint32)[:, 16]
    hit = want["hit"] != 0
    kinds = which[want["prim_id"][hit]]
    assert (kinds == 1).sum() > 200 and (kinds == 2).sum() > 200          # both kinds are hit
    assert (want["u"][hit][kinds == 2] == 0).all()                        # spheres carry no barycentrics


def test_host_instantiation_on_mixed_leaves(gen):
    sc, rays, gold, rec = gen
    assert query_exact(sc, rays).tobytes() == gold["closest"].tobytes()
    assert query_exact(sc, rays, any_hit=True, max_t=0.3).tobytes() == gold["any"].tobytes()
    assert multi_hit_exact(sc, rays[:1500], 8).tobytes() == gold["multi8"].tobytes()
    w = HostWide(sc)
    assert classify_closest(gold["closest"], w.query(rays))["bad"] == 0
    assert (w.query(rays, any_hit=True, max_t=0.3)["hit"] == gold["any"]["hit"]).all()
    m = w.multi_hit(rays[:1500], 8)
    assert (m["t"] == gold["multi8"]["t"]).all() and (m["hit"] == gold["multi8"]["hit"]).all()
    w.close()


def frame_case(kernel, nb, w=96, h=64):
    sc = S.make_scene("generic", 0.05)
    sc.prims.view(np.uint32)[:, 0] = (np.arange(sc.n_prims) // 3) % 4
    fp = S.frame_params(kernel, w, h, (0.5, 0.6, 2.6), (0.5, 0.5, 0.5), n_materials=4, num_bounces=4, ambient=(0.3, 0.3, 0.3, 1.0))
    sd = S.SurfaceData(nb, S.generic_normals(sc.prims, nb == 1))
    return sc, fp, sd, S.generic_materials(), S.point_light((1.5, 2.0, 2.5))


FRAME_CASES = [(0, 0), (0, 1), (1, 0), (1, 1)]        # (kernel, normal binding)


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref is only built where /root/reference exists")
@pytest.mark.parametrize("kernel,nb", FRAME_CASES)
def test_frames_oracle_matches_reference_headers(kernel, nb):
    """get_surface() over a variant: triangles follow the normal binding, spheres take get_normal(hr, sphere)
    (detail/generic_primitive_get_normal.h)."""
    sc, fp, sd, mats, light = frame_case(kernel, nb)
    got = O.oracle_frame_surface(sc, fp, sd, mats, light)
    want = O.ref_frame_generic(sc, fp, sd, mats, light)
    assert np.abs(got - want).max() <= 1e-6 and want[..., :3].std() > 0.1


@pytest.mark.parametrize("kernel,nb", FRAME_CASES)
def test_frames_oracle_matches_fixture(kernel, nb):
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_generic.npz"))
    sc, fp, sd, mats, light = frame_case(kernel, nb, 40, 28)
    assert np.abs(O.oracle_frame_surface(sc, fp, sd, mats, light) - g[f"frame_k{kernel}_nb{nb}"]).max() <= 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("kernel,nb", FRAME_CASES)
def test_gpu_frames_over_mixed_leaves(gpu, kernel, nb):
    sc, fp, sd, mats, light = frame_case(kernel, nb, 160, 112)
    want = O.oracle_frame_surface(sc, fp, sd, mats, light)
    dev = gpu.Scene(sc)
    try:
        gpu.set_surface(dev, sd)
        for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
            img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
            gpu.frame(dev, fp, None, mats, light, img, traverse=trav)
            bad = float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())
            assert bad <= (0.0 if trav == gpu.TRAVERSE_EXACT else 1e-3), (kernel, nb, trav, bad)
    finally:
        dev.close()


@pytest.mark.gpu
def test_gpu_mixed_leaves(gpu, gen):
    sc, rays, gold, rec = gen
    dev = gpu.Scene(sc)
    try:
        assert gpu.closest_hit(dev, rays, traverse=gpu.TRAVERSE_EXACT).tobytes() == gold["closest"].tobytes()
        assert gpu.any_hit(dev, rays, max_t=0.3, traverse=gpu.TRAVERSE_EXACT).tobytes() == gold["any"].tobytes()
        assert gpu.multi_hit(dev, rays[:1500], 8, traverse=gpu.TRAVERSE_EXACT).tobytes() == gold["multi8"].tobytes()
        assert classify_closest(gold["closest"], gpu.closest_hit(dev, rays))["bad"] == 0
        assert (gpu.any_hit(dev, rays, max_t=0.3)["hit"] == gold["any"]["hit"]).all()
        m = gpu.multi_hit(dev, rays[:1500], 8)
        assert (m["t"] == gold["multi8"]["t"]).all() and (m["hit"] == gold["multi8"]["hit"]).all()
    finally:
        dev.close()


@pytest.mark.gpu
def test_gpu_builder_and_full_size_mixed_scene(gpu):
    """200 k triangles + 100 k spheres, 2 M rays: the tree built on the GPU from the 80-byte records gives the
    reference tree's closest hits (ids exact up to t-ties), wide == reference-order walk on both trees."""
    sc = S.make_scene("generic", 200_000 / 60_000)
    lo, hi = sc.bounds()
    rays = S.random_rays(1 << 21, 19, float(lo.min()), float(hi.max()))
    dev = gpu.Scene(sc)
    built = gpu.Scene.build(2, sc.prims)
    try:
        exact = gpu.closest_hit(dev, rays, traverse=gpu.TRAVERSE_EXACT)
        sample = slice(0, 20000)
        assert exact[sample].tobytes() == O.oracle_query(sc, rays[sample], False).tobytes()
        assert classify_closest(exact, gpu.closest_hit(dev, rays))["bad"] == 0
        assert classify_closest(exact, gpu.closest_hit(built, rays))["bad"] == 0
        assert exact["hit"].mean() > 0.05
    finally:
        dev.close(); built.close()
