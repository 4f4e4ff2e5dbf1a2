"""CPU tier for the GPU BVH builder (csrc/build_ploc.cuh; SURVEY section 8f row 1, replaces
build<index_bvh<P>>(), bvh.h:455 + detail/bvh/{build.inl,sah.h}).  tests/hostsim runs the same
element functions serially, so the tree checked here is the tree the kernels build.

* the output is a valid index_bvh in the reference's layout (bvh.h:52-119,189-236): children
  adjacent, every node reachable once, leaf ranges partition indices[], indices[] a permutation,
  boxes contain what is below them, leaves <= 4 primitives;
* the reference's traversal (the oracle) over this tree returns the same closest hits as over the
  reference builder's tree (ids exact except t-ties), and the same occlusion answers;
* the product's own traversals (wide BVH8 re-encode, exact BVH2 walk) agree with the oracle on it.
"""
import dataclasses

import numpy as np
import pytest

from conftest import classify_closest
from hostsim_bindings import HostWide, build_ploc, query_exact
from oracle import bindings as O
from visionaray_b200 import scenes as S

SCENES = [("cfg1_sphere", 0.05), ("cfg2_grid", 0.01), ("cfg3_cubes", 0.001), ("cfg5_soup", 0.0005), ("spheres", 0.03)]


def prim_bounds(sc):
    if sc.prim_kind == 0:
        v1 = sc.prims[:, 4:7]
        v2 = sc.prims[:, 4:7] + sc.prims[:, 8:11]
        v3 = sc.prims[:, 4:7] + sc.prims[:, 12:15]
        return np.minimum(np.minimum(v1, v2), v3), np.maximum(np.maximum(v1, v2), v3)
    c, r = sc.prims[:, 4:7], sc.prims[:, 8:9]
    return c - r, c + r


def check_bvh2(sc, nodes, indices, max_leaf=4):
    """Validity of a reference-layout tree; returns its depth (root = 1)."""
    n = nodes.shape[0]
    assert sorted(indices.tolist()) == list(range(sc.n_prims)), "indices[] is a permutation of the primitives"
    plo, phi = prim_bounds(sc)
    seen = np.zeros(n, dtype=np.int64)
    covered = np.zeros(indices.shape[0], dtype=np.int64)
    depth = 0
    stack = [(0, 1)]
    while stack:
        i, d = stack.pop()
        seen[i] += 1
        depth = max(depth, d)
        nd = nodes[i]
        lo, hi = nd["bbox_min"], nd["bbox_max"]
        if nd["num_prims"] == 0:
            c = int(nd["first"])
            assert 1 <= c and c + 1 < n
            for k in (c, c + 1):
                assert (nodes[k]["bbox_min"] >= lo).all() and (nodes[k]["bbox_max"] <= hi).all(), "child box outside parent"
                stack.append((k, d + 1))
        else:
            a, cnt = int(nd["first"]), int(nd["num_prims"])
            assert cnt <= max_leaf and a + cnt <= indices.shape[0]
            covered[a:a + cnt] += 1
            p = indices[a:a + cnt]
            assert (plo[p] >= lo).all() and (phi[p] <= hi).all(), "primitive outside its leaf box"
            # tight: the leaf box IS the union of its primitives
            assert (plo[p].min(axis=0) == lo).all() and (phi[p].max(axis=0) == hi).all()
    assert (seen == 1).all(), "every node reachable exactly once"
    assert (covered == 1).all(), "leaf ranges partition indices[]"
    assert n == 2 * int((nodes["num_prims"] > 0).sum()) - 1
    return depth


@pytest.fixture(scope="module", params=SCENES, ids=[s[0] for s in SCENES])
def built(request):
    name, scale = request.param
    sc = S.make_scene(name, scale)
    nodes, idx, info = build_ploc(sc.prim_kind, sc.prims)
    return sc, dataclasses.replace(sc, nodes=nodes, indices=idx), info


def test_ploc_tree_is_a_valid_index_bvh(built):
    sc, mine, info = built
    depth = check_bvh2(mine, mine.nodes, mine.indices)
    assert depth == info["depth"]
    assert (mine.nodes[0]["bbox_min"] == sc.nodes[0]["bbox_min"]).all()      # same root bounds as the reference tree
    assert (mine.nodes[0]["bbox_max"] == sc.nodes[0]["bbox_max"]).all()


def test_ploc_tree_gives_the_reference_hits(built):
    sc, mine, _ = built
    lo, hi = sc.bounds()
    rays = S.random_rays(4000, 11, float(lo.min()) - 0.2, float(hi.max()) + 0.2)
    want = O.oracle_query(sc, rays, False)
    got = O.oracle_query(mine, rays, False)        # the reference's traversal over the new tree
    r = classify_closest(want, got)
    assert r["bad"] == 0, r
    assert ((want["t"] == got["t"]) | (want["prim_id"] != got["prim_id"]))[want["hit"] != 0].all()
    for mt in (0.3, 3.4028234663852886e38):
        a = O.oracle_query(sc, rays, True, mt)
        b = O.oracle_query(mine, rays, True, mt)
        assert (a["hit"] == b["hit"]).all()


def test_product_traversals_on_ploc_tree(built):
    sc, mine, _ = built
    lo, hi = sc.bounds()
    rays = S.random_rays(3000, 12, float(lo.min()) - 0.2, float(hi.max()) + 0.2)
    want = O.oracle_query(mine, rays, False)
    exact = query_exact(mine, rays, False)
    assert exact.tobytes() == want.tobytes()        # reference visit order on the new tree: bit-identical
    w = HostWide(mine)
    got = w.query(rays, False)
    assert classify_closest(want, got)["bad"] == 0
    assert (w.query(rays, True, 0.25)["hit"] == O.oracle_query(mine, rays, True, 0.25)["hit"]).all()
    w.close()


def test_ploc_degenerate_inputs():
    # one primitive, two primitives, many identical primitives (all Morton codes equal, all merge areas tie)
    from conftest import tri

    one = np.stack([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), 0)])
    nodes, idx, info = build_ploc(0, one)
    assert nodes.shape[0] == 1 and nodes[0]["num_prims"] == 1 and idx.tolist() == [0] and info["depth"] == 1
    two = np.stack([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), 0), tri((5, 5, 5), (6, 5, 5), (5, 6, 5), 1)])
    nodes, idx, info = build_ploc(0, two)
    sc = S.HostScene(0, two, nodes, idx)
    check_bvh2(sc, nodes, idx)
    same = np.stack([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), i) for i in range(301)])
    nodes, idx, info = build_ploc(0, same)
    assert info["iterations"] <= 12 and info["depth"] <= 12      # ties pair up (0,1)(2,3)..: log2 n passes, not n
    sc = S.HostScene(0, same, nodes, idx)
    check_bvh2(sc, nodes, idx)
    rays = np.zeros((1, 8), dtype=np.float32)
    rays[0, 0:3] = (0.2, 0.2, 1.0)
    rays[0, 4:7] = (0.0, 0.0, -1.0)
    h = O.oracle_query(sc, rays, False)
    assert h["hit"][0] == 1 and h["t"][0] == 1.0


def test_ploc_radius_and_costs_keep_validity():
    sc = S.make_scene("cfg3_cubes", 0.0005)
    for radius, cp in ((1, 1.0), (4, 0.3), (32, 3.0)):
        nodes, idx, _ = build_ploc(0, sc.prims, radius, 1.0, cp)
        check_bvh2(dataclasses.replace(sc, nodes=nodes, indices=idx), nodes, idx)
