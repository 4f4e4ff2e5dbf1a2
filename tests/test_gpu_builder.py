"""GPU tier: the BVH built ON the GPU (vsb_scene_build, csrc/build_ploc.cuh) -- SURVEY section 8f
row 1, replacing build<index_bvh<P>>() (bvh.h:455 -> detail/bvh/build.inl:28-140, sah.h:150-762).

Parity: any valid BVH returns the reference's closest hits up to t-ties, so the hits of the
GPU-built scene are compared with the oracle on the REFERENCE builder's tree; the downloaded tree
itself is checked as an index_bvh, is bit-identical to the serial instantiation of the same
element functions (tests/hostsim), and the reference-order traversal over it is bit-identical to
the oracle's traversal of the same downloaded arrays."""
import dataclasses

import numpy as np
import pytest

from conftest import classify_closest, tri
from hostsim_bindings import build_ploc
from oracle import bindings as O
from test_builder_hostsim import check_bvh2, prim_bounds
from test_hostsim import scene_rays
from visionaray_b200 import scenes as S

pytestmark = pytest.mark.gpu

FLT_MAX = 3.4028234663852886e38
SCENES = [("cfg1_sphere", 0.2), ("cfg2_grid", 0.05), ("cfg3_cubes", 0.004), ("cfg5_soup", 0.002), ("spheres", 0.2)]


@pytest.fixture(scope="module", autouse=True)
def ploc_builder(gpu):
    """This file tests the Morton + PLOC builder; the default of vsb_scene_build is build_sah.cuh (test_gpu_sah_builder.py)."""
    gpu.tune("build_algorithm", 0)
    yield
    gpu.tune("build_algorithm", 1)


@pytest.fixture(scope="module")
def built(gpu, ploc_builder):
    cache = {}

    def get(name, scale):
        key = (name, scale)
        if key not in cache:
            sc = S.make_scene(name, scale)
            dev = gpu.Scene.build(sc.prim_kind, sc.prims, sc.normals)
            nodes, idx = dev.download_bvh()
            cache[key] = (sc, dev, dataclasses.replace(sc, nodes=nodes, indices=idx))
        return cache[key]

    yield get
    for _, dev, _ in cache.values():
        dev.close()


@pytest.mark.parametrize("name,scale", SCENES)
def test_gpu_tree_is_valid_and_equals_the_serial_build(gpu, built, name, scale):
    sc, dev, mine = built(name, scale)
    depth = check_bvh2(mine, mine.nodes, mine.indices)
    st = dev.stats()
    assert st["bvh2_depth"] == depth and st["bvh2_nodes"] == mine.nodes.shape[0]
    nodes, idx, info = build_ploc(sc.prim_kind, sc.prims)
    assert nodes.tobytes() == mine.nodes.tobytes(), "GPU tree differs from the serial form of the same algorithm"
    assert (idx == mine.indices).all()
    assert gpu.last_build_ms()["iterations"] >= 1


@pytest.mark.parametrize("name,scale", SCENES)
def test_hits_on_gpu_built_scene_match_reference_tree(gpu, built, name, scale):
    sc, dev, mine = built(name, scale)
    rays = scene_rays(sc, 200_000, 52)
    want = O.oracle_query(sc, rays, False)                 # the reference's traversal over the reference builder's tree
    got = gpu.closest_hit(dev, rays)                       # BVH8 over the GPU-built tree
    r = classify_closest(want, got)
    assert r["bad"] == 0, r
    same = (want["hit"] != 0) & (want["prim_id"] == got["prim_id"])
    for k in ("t", "u", "v"):
        assert (want[k][same] == got[k][same]).all(), k
    # list index follows THIS tree's indices[]
    assert (mine.indices[got["prim_list_index"][got["hit"] != 0]] == got["prim_id"][got["hit"] != 0]).all() or sc.prim_kind == 1
    for mt in (FLT_MAX, 0.3):
        a = O.oracle_query(sc, rays, True, mt)
        b = gpu.any_hit(dev, rays, max_t=mt)
        assert (a["hit"] == b["hit"]).all()
    # reference-order walk over the new tree == the oracle over the same arrays, bit for bit
    sub = rays[:50_000]
    for any_hit, fn in ((False, gpu.closest_hit), (True, gpu.any_hit)):
        a = O.oracle_query(mine, sub, any_hit, 0.7)
        b = fn(dev, sub, max_t=0.7, traverse=gpu.TRAVERSE_EXACT)
        assert a.tobytes() == b.tobytes()


def test_frames_on_gpu_built_scene(gpu, built):
    sc, dev, mine = built("cfg2_grid", 0.05)
    eye = (0.3, 1.2, 2.2)
    for kernel, ps, mat in ((0, 0, 0), (1, 0, 0), (2, 0, 0)):
        fp = S.frame_params(kernel, 256, 192, eye, (0, 0, 0), pixel_sampler=ps, num_bounces=4)
        mats, lights = S.default_material(mat), S.point_light(eye)
        want = O.oracle_frame(sc, fp, mats, lights)        # oracle on the reference tree
        img = np.zeros((192, 256, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, mats, lights, img)
        frac_bad = float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())
        assert frac_bad <= (2e-4 if kernel < 2 else 2e-3), (kernel, frac_bad)


def test_builder_degenerate_inputs(gpu):
    one = np.stack([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), 7)])
    dev = gpu.Scene.build(0, one)
    nodes, idx = dev.download_bvh()
    assert nodes.shape[0] == 1 and nodes[0]["num_prims"] == 1 and idx.tolist() == [0]
    rays = np.zeros((2, 8), dtype=np.float32)
    rays[:, 0:3] = (0.2, 0.2, 1.0)
    rays[0, 4:7] = (0.0, 0.0, -1.0)
    rays[1, 4:7] = (0.0, 0.0, 1.0)
    h = gpu.closest_hit(dev, rays)
    assert h["hit"].tolist() == [1, 0] and h["prim_id"][0] == 7 and h["t"][0] == 1.0
    dev.close()
    same = np.stack([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), i) for i in range(1000)])   # all codes and all areas tie
    dev = gpu.Scene.build(0, same)
    nodes, idx = dev.download_bvh()
    check_bvh2(S.HostScene(0, same, nodes, idx), nodes, idx)
    h = gpu.closest_hit(dev, rays)
    assert h["hit"].tolist() == [1, 0] and h["t"][0] == 1.0
    dev.close()
    with pytest.raises(RuntimeError):
        gpu.Scene.build(0, np.zeros((0, 16), dtype=np.float32))
    with pytest.raises(RuntimeError):
        gpu.Scene.build(5, one)


def test_builder_knobs_keep_validity_and_hits(gpu):
    sc = S.make_scene("cfg3_cubes", 0.002)
    rays = scene_rays(sc, 50_000, 53)
    want = O.oracle_query(sc, rays, False)
    try:
        for radius, cp in ((1, 100), (16, 30), (32, 300)):
            gpu.tune("build_radius", radius)
            gpu.tune("build_cost_prim_x100", cp)
            dev = gpu.Scene.build(0, sc.prims)
            nodes, idx = dev.download_bvh()
            check_bvh2(dataclasses.replace(sc, nodes=nodes, indices=idx), nodes, idx)
            assert classify_closest(want, gpu.closest_hit(dev, rays))["bad"] == 0
            dev.close()
    finally:
        gpu.tune("build_radius", 8)
        gpu.tune("build_cost_prim_x100", 100)


def test_builder_device_primitives_and_full_size(gpu):
    """cfg2's 1.0 M-triangle mesh, primitives already on the device: vectorised validity of the
    downloaded tree and closest hits against the scene built from the reference tree."""
    import torch

    sc = S.make_scene("cfg2_grid", 1.0)
    d_prims = torch.from_numpy(sc.prims).cuda()
    dev = gpu.Scene.build(0, d_prims)
    nodes, idx = dev.download_bvh()
    assert np.array_equal(np.sort(idx), np.arange(sc.n_prims, dtype=np.uint32))
    leaf = nodes["num_prims"] > 0
    assert nodes.shape[0] == 2 * int(leaf.sum()) - 1 and int(nodes["num_prims"].sum()) == sc.n_prims
    assert nodes["num_prims"].max() <= 4
    inner = np.nonzero(~leaf)[0]
    c0 = nodes["first"][inner].astype(np.int64)
    refs = np.concatenate([c0, c0 + 1])
    assert np.array_equal(np.sort(refs), np.arange(1, nodes.shape[0]))           # every non-root node has one parent
    for c in (c0, c0 + 1):
        assert (nodes["bbox_min"][c] >= nodes["bbox_min"][inner]).all() and (nodes["bbox_max"][c] <= nodes["bbox_max"][inner]).all()
    plo, phi = prim_bounds(sc)
    li = np.nonzero(leaf)[0]
    for k in range(4):
        sel = li[nodes["num_prims"][li] > k]
        p = idx[nodes["first"][sel] + k]
        assert (plo[p] >= nodes["bbox_min"][sel]).all() and (phi[p] <= nodes["bbox_max"][sel]).all()
    ref = gpu.Scene(sc)
    rays = torch.from_numpy(scene_rays(sc, 2_000_000, 54)).cuda()
    ha, hb = gpu.closest_hit(ref, rays), gpu.closest_hit(dev, rays)
    gpu.sync()                                             # the library's own stream (stream=None)
    a = ha.cpu().numpy().view(np.uint8).reshape(-1).view(O.HIT_DTYPE)
    b = hb.cpu().numpy().view(np.uint8).reshape(-1).view(O.HIT_DTYPE)
    assert classify_closest(a, b)["bad"] == 0
    assert (a["t"] == b["t"]).mean() > 0.9999
    ms = gpu.last_build_ms()
    print("build 1.0M tris on the GPU:", ms, dev.stats())
    ref.close()
    dev.close()
