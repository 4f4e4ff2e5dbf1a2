"""CPU tier: the algorithm of the GPU's binned-SAH builder (csrc/build_sah.cuh), restated in numpy (tests/sah_levelwise.py),
builds the host builder's tree -- the reference's, byte for byte (tests/test_host_bvh.py) -- when the nodes are processed
level by level instead of recursively: same nodes at the same indices, same bounds, same indices[].  Checks the three
claims in the header of sah_levelwise.py without a GPU; the CUDA implementation itself is compared with the same host
trees in tests/test_gpu_sah_builder.py."""
import numpy as np
import pytest

import sah_levelwise as L
from conftest import tri
from test_builder_hostsim import prim_bounds
from visionaray_b200 import scenes as S


def assert_same(nodes, idx, want_nodes, want_idx):
    assert nodes.shape == want_nodes.shape
    assert np.array_equal(nodes["first"], want_nodes["first"]) and np.array_equal(nodes["num_prims"], want_nodes["num_prims"])
    assert np.array_equal(nodes["bbox_min"], want_nodes["bbox_min"]) and np.array_equal(nodes["bbox_max"], want_nodes["bbox_max"])
    assert np.array_equal(idx, want_idx)


@pytest.mark.parametrize("name,scale", [("cfg1_sphere", 0.05), ("cfg2_grid", 0.004), ("cfg3_cubes", 0.0008), ("cfg5_soup", 0.0002),
                                        ("spheres", 0.05)])
def test_level_by_level_equals_the_recursive_builder(name, scale):
    sc = S.make_scene(name, scale)
    lo, hi = prim_bounds(sc)
    nodes, idx, levels = L.build(lo, hi)
    assert_same(nodes, idx, sc.nodes, sc.indices)
    assert levels >= 3


def test_hoare_closed_form_equals_the_sequential_partition():
    """std::partition of libstdc++ for bidirectional iterators, statement for statement, against the closed form."""
    rng = np.random.default_rng(11)
    for m in (0, 1, 2, 5, 17, 64, 257):
        for p in (0.0, 0.2, 0.5, 0.9, 1.0):
            pred = rng.random(m) < p
            a = list(range(m))
            first, last = 0, m
            while True:
                while True:
                    if first == last:
                        break
                    if pred[a[first]]:
                        first += 1
                    else:
                        break
                if first == last:
                    break
                last -= 1
                while True:
                    if first == last:
                        break
                    if not pred[a[last]]:
                        last -= 1
                    else:
                        break
                if first == last:
                    break
                a[first], a[last] = a[last], a[first]
                first += 1
            perm, t = L.hoare_partition(pred)
            assert t == first == int(pred.sum())
            assert perm.tolist() == a


def test_degenerate_inputs():
    one = np.array([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), 0)], dtype=np.float32)
    a = np.repeat(one, 9, axis=0)
    b = np.repeat(np.array([tri((5, 0, 0), (6, 0, 0), (5, 1, 0), 0)], dtype=np.float32), 11, axis=0)
    for prims in (one, np.repeat(one, 3, axis=0), np.repeat(one, 37, axis=0), np.concatenate([a, b, a])):
        prims = prims.copy()
        prims.view(np.uint32)[:, 1] = np.arange(prims.shape[0])
        want_nodes, want_idx = S.build_bvh(0, prims)
        sc = S.HostScene(0, prims, want_nodes, want_idx, None, "degenerate")
        nodes, idx, _ = L.build(*prim_bounds(sc))
        assert_same(nodes, idx, want_nodes, want_idx)
