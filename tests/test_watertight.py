"""VSB_TRAVERSE_WATERTIGHT (csrc/watertight.cuh, north_star's "Woop / Moeller-Trumbore watertight triangle test"): the hit /
miss decision of Woop, Benthin, Wald (JCGT 2013), t / u / v in the reference's Moeller-Trumbore form.  The reference's
own test (math/intersect.h:122-179) lets rays through on shared edges and stays the default; this mode is opt-in.

Stress mesh: a closed height field over [0, 1]^2 whose vertices are dyadic (multiples of 1/32 in x / z, 1/2048 in y), so
edge vectors are exact differences and the (v1, e1, e2) records reconstruct every shared vertex to the same floats.
Rays start above the mesh and aim at points on shared edges and at shared vertices (up to the rounding of the direction):
every such ray hits the surface, on one side of the edge or the other.  CPU tier: the host instantiation of the same code (tests/hostsim); GPU tier: through the C ABI."""
import numpy as np
import pytest

from conftest import tri
from hostsim_bindings import HostWide, query_watertight
from visionaray_b200 import scenes as S

N = 32


def _height(i, j):
    # multiples of 1/2048, at most 36/2048 over a 1/32 grid: slopes below 30 degrees, so the near-vertical rays below
    # see every triangle from above (no silhouette edges, where a grazing ray may legitimately miss both neighbours)
    return ((i * 7 + j * 13 + (i * j) % 5) % 37) / 2048.0


def dyadic_mesh():
    tris = []
    pid = 0
    for j in range(N):
        for i in range(N):
            p = lambda a, b: (a / N, _height(a, b), b / N)
            a, b, c, d = p(i, j), p(i + 1, j), p(i + 1, j + 1), p(i, j + 1)
            tris.append(tri(a, b, c, pid)); pid += 1
            tris.append(tri(a, c, d, pid)); pid += 1
    prims = np.array(tris, dtype=np.float32)
    nodes, idx = S.build_bvh(0, prims)
    return S.HostScene(0, prims, nodes, idx, S.face_normals(prims), "dyadic_heightfield")


def edge_rays(seed=5, n=400_000):
    """Rays from dyadic origins above the mesh through points ON shared edges / AT shared vertices (interior ones)."""
    rng = np.random.default_rng(seed)
    i = rng.integers(1, N - 1, n); j = rng.integers(1, N - 1, n)
    kind = rng.integers(0, 4, n)                       # 0 vertex, 1 x-edge midpoint, 2 z-edge midpoint, 3 diagonal midpoint
    tx = (i + np.where((kind == 1) | (kind == 3), 0.5, 0.0)) / N
    tz = (j + np.where((kind == 2) | (kind == 3), 0.5, 0.0)) / N
    h = np.vectorize(_height)
    ty = np.where(kind == 0, h(i, j),
         np.where(kind == 1, (h(i, j) + h(i + 1, j)) / 2,
         np.where(kind == 2, (h(i, j) + h(i, j + 1)) / 2, (h(i, j) + h(i + 1, j + 1)) / 2)))
    # origins within 0.25 of the target horizontally and >= 1 above it (at most ~20 degrees off the vertical), at arbitrary
    # (not dyadic) positions: the direction float32(target - origin) aims at the edge point up to rounding, so in exact
    # arithmetic the ray passes a hair to one side of the edge or the other -- and hits the triangle on that side
    ox = (tx + rng.uniform(-0.25, 0.25, n)).astype(np.float32)
    oz = (tz + rng.uniform(-0.25, 0.25, n)).astype(np.float32)
    oy = (1.0 + rng.uniform(0.0, 1.0, n)).astype(np.float32)
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, 0], rays[:, 1], rays[:, 2] = ox, oy, oz
    rays[:, 4] = (tx - ox.astype(np.float64)).astype(np.float32)
    rays[:, 5] = (ty - oy.astype(np.float64)).astype(np.float32)
    rays[:, 6] = (tz - oz.astype(np.float64)).astype(np.float32)
    return rays


def test_host_instantiation_is_watertight_where_moeller_trumbore_leaks():
    sc = dyadic_mesh()
    rays = edge_rays()
    wide = HostWide(sc)
    mt = wide.query(rays)
    wt = query_watertight(wide, rays)
    leaks_mt, leaks_wt = int((mt["hit"] == 0).sum()), int((wt["hit"] == 0).sum())
    assert leaks_wt == 0, leaks_wt                     # every ray aims at the surface
    assert leaks_mt > 0                                # the reference's test lets some of them through (why the mode exists)
    # where both accept the same primitive, the records are bit-identical (t / u / v in the reference's form)
    same = (mt["hit"] != 0) & (mt["prim_id"] == wt["prim_id"])
    assert same.sum() > 0.3 * rays.shape[0] and mt[same].tobytes() == wt[same].tobytes()     # edge rays often tie between neighbours
    # a hit never gets farther: the watertight decision accepts a superset of what Moeller-Trumbore accepts (up to ulps of t)
    both = (mt["hit"] != 0)
    assert (wt["t"][both] <= mt["t"][both] * (1 + 1e-6)).all()
    occ = query_watertight(wide, rays, any_hit=True, max_t=10.0)
    assert (occ["hit"] != 0).all()
    wide.close()


def test_host_generic_rays_change_almost_nothing():
    sc = S.make_scene("cfg2_grid", 0.02)
    rays = S.random_rays(100_000, 9, -1.2, 1.2)
    wide = HostWide(sc)
    mt, wt = wide.query(rays), query_watertight(wide, rays)
    changed = int((mt["hit"] != wt["hit"]).sum())
    assert changed <= 10, changed                      # random rays practically never meet an edge exactly
    same = (mt["hit"] != 0) & (wt["hit"] != 0) & (mt["prim_id"] == wt["prim_id"])
    assert mt[same].tobytes() == wt[same].tobytes()
    wide.close()


@pytest.mark.gpu
def test_gpu_watertight_mode_through_the_c_abi(gpu):
    import torch

    sc = dyadic_mesh()
    dev = gpu.Scene(sc)
    rays = edge_rays()
    wide = HostWide(sc)
    want = query_watertight(wide, rays)
    got = gpu.closest_hit(dev, rays, traverse=gpu.TRAVERSE_WATERTIGHT)
    assert got.tobytes() == want.tobytes()             # GPU == host instantiation of the same code
    assert int((got["hit"] == 0).sum()) == 0
    mt = gpu.closest_hit(dev, rays)
    leaks = int((mt["hit"] == 0).sum())
    assert leaks > 0
    d_rays = torch.from_numpy(rays).cuda()
    d_hits = gpu.closest_hit(dev, d_rays, traverse=gpu.TRAVERSE_WATERTIGHT)
    torch.cuda.synchronize()
    assert d_hits.cpu().numpy().view(want.dtype).reshape(-1).tobytes() == want.tobytes()
    occ = gpu.any_hit(dev, rays, max_t=10.0, traverse=gpu.TRAVERSE_WATERTIGHT)
    assert (occ["hit"] != 0).all()
    per_ray = np.full(rays.shape[0], 10.0, dtype=np.float32); per_ray[::2] = 1e-3
    occ2 = gpu.any_hit(dev, rays, max_t_per_ray=per_ray, traverse=gpu.TRAVERSE_WATERTIGHT)
    assert (occ2["hit"][::2] == 0).all() and (occ2["hit"][1::2] != 0).all()
    wide.close(); dev.close()
    # the BASELINE cfg2 view: primary rays of a 1024^2 frame on the 1 M-triangle grid -- how many rays change hit / miss
    sc2 = S.make_scene("cfg2_grid", 1.0)
    dev2 = gpu.Scene(sc2)
    fp = S.frame_params(0, 1024, 1024, (0.3, 1.2, 2.2), (0, 0, 0))
    pr = S.primary_rays(fp)
    a, b = gpu.closest_hit(dev2, pr), gpu.closest_hit(dev2, pr, traverse=gpu.TRAVERSE_WATERTIGHT)
    changed = int(((a["hit"] != 0) != (b["hit"] != 0)).sum())
    moved = int(((a["hit"] != 0) & (b["hit"] != 0) & (a["prim_id"] != b["prim_id"])).sum())
    print(f"cfg2 primary rays: {changed} of {pr.shape[0]} change hit/miss, {moved} change primitive; stress list: {leaks} of {rays.shape[0]} leak with the reference's test")
    assert changed <= 1e-4 * pr.shape[0] and moved <= 1e-3 * pr.shape[0]
    sphere = S.make_scene("spheres", 0.01)
    ds = gpu.Scene(sphere)
    with pytest.raises(RuntimeError):
        gpu.closest_hit(ds, rays[:10], traverse=gpu.TRAVERSE_WATERTIGHT)      # triangle scenes only
    ds.close(); dev2.close()
