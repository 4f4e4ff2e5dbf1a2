"""GPU tier: vsb_scene_build with build_algorithm = 1 (csrc/build_sah.cuh) builds the REFERENCE builder's tree --
build<index_bvh<P>>(first, last) (bvh.h:455 -> detail/bvh/build.inl:28-140 -> binned_sah_builder, detail/bvh/sah.h:150-762)
-- breadth-first on the GPU: the downloaded nodes and indices equal the host builder's (include/visionaray_b200/bvh.h,
itself byte-identical to the reference's: tests/test_host_bvh.py) node for node: child / primitive ranges and indices bit
for bit, bounds equal as floats (the one caveat is the sign of a zero, see build_sah.cuh)."""
import numpy as np
import pytest

from conftest import tri
from visionaray_b200 import scenes as S

pytestmark = pytest.mark.gpu

SCENES = [("cfg1_sphere", 0.2), ("cfg2_grid", 0.05), ("cfg3_cubes", 0.004), ("cfg5_soup", 0.002), ("spheres", 0.2),
          ("cfg2_grid", 0.5), ("cfg5_soup", 0.02)]


@pytest.fixture()
def sah(gpu):
    gpu.tune("build_algorithm", 1)
    yield gpu
    gpu.tune("build_algorithm", 1)                     # the default


def assert_same_tree(nodes, idx, want_nodes, want_idx):
    assert nodes.shape == want_nodes.shape, (nodes.shape, want_nodes.shape)
    for k in ("first", "num_prims"):
        bad = np.flatnonzero(nodes[k] != want_nodes[k])
        assert bad.size == 0, (k, bad[:5], nodes[bad[:5]], want_nodes[bad[:5]])
    for k in ("bbox_min", "bbox_max"):
        assert np.array_equal(nodes[k], want_nodes[k]), k              # as floats: -0.0 == +0.0
    assert np.array_equal(idx, want_idx)


@pytest.mark.parametrize("name,scale", SCENES)
def test_gpu_builds_the_reference_tree(sah, name, scale):
    sc = S.make_scene(name, scale)                                     # sc.nodes / sc.indices: the host builder's tree
    dev = sah.Scene.build(sc.prim_kind, sc.prims)
    nodes, idx = dev.download_bvh()
    assert_same_tree(nodes, idx, sc.nodes, sc.indices)
    ms = sah.last_build_ms()
    assert ms["iterations"] >= 1
    # the hit records are then the ones of a host-built scene, bit for bit (same tree -> same BVH8 up to allocation order)
    ref = sah.Scene(sc)
    rays = S.random_rays(100_000, 5, -1.5, 1.5)
    assert sah.closest_hit(dev, rays).tobytes() == sah.closest_hit(ref, rays).tobytes()
    assert sah.closest_hit(dev, rays, traverse=sah.TRAVERSE_EXACT).tobytes() == sah.closest_hit(ref, rays, traverse=sah.TRAVERSE_EXACT).tobytes()
    dev.close(); ref.close()


def test_generic_primitives_and_device_input(sah):
    import torch

    sc = S.make_scene("generic", 0.05)
    dev = sah.Scene.build(sc.prim_kind, torch.from_numpy(sc.prims).cuda())
    nodes, idx = dev.download_bvh()
    assert_same_tree(nodes, idx, sc.nodes, sc.indices)
    dev.close()


def test_degenerate_inputs(sah):
    # up to 4 primitives: one leaf; identical primitives: no centroid extent, one big leaf; two clusters of duplicates
    one = np.array([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), 0)], dtype=np.float32)
    for prims in (one, np.repeat(one, 3, axis=0), np.repeat(one, 37, axis=0)):
        prims = prims.copy()
        prims.view(np.uint32)[:, 1] = np.arange(prims.shape[0])
        want_nodes, want_idx = S.build_bvh(0, prims)
        dev = sah.Scene.build(0, prims)
        nodes, idx = dev.download_bvh()
        assert_same_tree(nodes, idx, want_nodes, want_idx)
        dev.close()
    a = np.repeat(one, 9, axis=0)
    b = np.repeat(np.array([tri((5, 0, 0), (6, 0, 0), (5, 1, 0), 0)], dtype=np.float32), 11, axis=0)
    prims = np.concatenate([a, b, a]).copy()
    prims.view(np.uint32)[:, 1] = np.arange(prims.shape[0])
    want_nodes, want_idx = S.build_bvh(0, prims)
    dev = sah.Scene.build(0, prims)
    nodes, idx = dev.download_bvh()
    assert_same_tree(nodes, idx, want_nodes, want_idx)
    dev.close()
    # a non-power-of-two soup whose size is not a multiple of the chunk, with a skewed distribution (deep tree)
    rng = np.random.default_rng(3)
    n = 12_345
    c = (rng.random((n, 3)) ** 6).astype(np.float32)
    prims = np.array([tri(tuple(p), tuple(p + np.float32([0.01, 0, 0])), tuple(p + np.float32([0, 0.01, 0])), i) for i, p in enumerate(c)], dtype=np.float32)
    want_nodes, want_idx = S.build_bvh(0, prims)
    dev = sah.Scene.build(0, prims)
    nodes, idx = dev.download_bvh()
    assert_same_tree(nodes, idx, want_nodes, want_idx)
    dev.close()
