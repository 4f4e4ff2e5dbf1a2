"""CPU tier: the oracle restatement against (1) the reference's own known-answer tests,
(2) the reference headers compiled here (oracle/_ref), (3) committed golden vectors generated
from the reference, plus the host builder against the reference builder."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import tri
from oracle import bindings as O
from visionaray_b200 import scenes as S
from visionaray_b200._native import NODE_DTYPE

FLT_MAX = 3.4028234663852886e38
GOLD = os.path.join(os.path.dirname(__file__), "golden")


def two_triangles():
    """test/unittests/get_normal.cpp:17-35"""
    t = np.stack([
        tri((-1, -1, 1), (1, -1, 1), (1, 1, 1), 0, 0),
        tri((1, -1, -1), (-1, -1, -1), (-1, 1, -1), 1, 1),
    ])
    nodes, idx = S.build_bvh(0, t)
    return S.HostScene(0, t, nodes, idx, S.face_normals(t), "two_triangles")


def test_known_answer_get_normal_bvh():
    """GetNormal.BVH (test/unittests/get_normal.cpp:36-46): hit, t == 1.0f, prim_id == 0."""
    sc = two_triangles()
    ray = np.array([[0.5, -0.5, 2.0, 0, 0.0, 0.0, -1.0, 0]], dtype=np.float32)
    h = O.oracle_query(sc, ray, False)
    assert h["hit"][0] == 1 and h["t"][0] == np.float32(1.0) and h["prim_id"][0] == 0
    # get_normal(hr, bvh) == normalize(cross(e1, e2)) (get_normal.cpp:48-54)
    n = sc.normals[h["prim_id"][0]][:3]
    e1, e2 = sc.prims[0][8:11], sc.prims[0][12:15]
    want = np.cross(e1, e2); want = want / np.linalg.norm(want)
    assert np.allclose(n, want, rtol=1e-6)


def test_known_answer_build_sizes():
    """BVH.BuildIndexBvh (test/unittests/bvh/build.cpp:69-116): nodes non-empty, prims kept."""
    t = np.stack([tri((0, 0, 0), (1, 0, 0), (0, 1, 0), i) for i in range(3)])
    nodes, idx = S.build_bvh(0, t)
    assert nodes.shape[0] > 0 and idx.shape[0] == 3
    sph = np.zeros((5, 12), dtype=np.float32)
    for i, r in enumerate([1e-9, 1e-3, 1.0, 1e3, 1e5]):
        sph[i, 4:7] = (i, 0, 0); sph[i, 8] = r; sph[i].view(np.uint32)[1] = i
    nodes, idx = S.build_bvh(1, sph)
    assert nodes.shape[0] > 0 and sorted(idx.tolist()) == list(range(5))


def test_layouts_match_reference():
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    lay = (C.c_int * 13)()
    O.ref_lib().ref_layout(lay)
    assert list(lay) == [16, 64, 48, 32, 32, 32, 16, 32, 48, 4, 0, 16, 32]


SCENES = [("cfg1_sphere", 0.05), ("cfg2_grid", 0.01), ("cfg3_cubes", 0.001), ("cfg5_soup", 0.0005), ("spheres", 0.03)]


@pytest.mark.parametrize("name,scale", SCENES)
def test_oracle_bit_identical_to_reference(name, scale):
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    sc = S.make_scene(name, scale)
    rn, ri = O.ref_build(sc.prim_kind, sc.prims)
    assert rn.tobytes() == sc.nodes.tobytes() and ri.tobytes() == sc.indices.tobytes(), "host builder != reference builder"
    lo, hi = sc.bounds()
    rays = S.random_rays(6000, 11, float(lo.min()) - 0.2, float(hi.max()) + 0.2)
    per_ray = np.linspace(0.05, 2.0, rays.shape[0]).astype(np.float32)
    for any_hit in (False, True):
        for mt, pr in ((FLT_MAX, None), (0.3, None), (FLT_MAX, per_ray)):
            a = O.oracle_query(sc, rays, any_hit, mt, pr)
            b = O.ref_query(sc, rays, any_hit, mt, pr)
            assert a.tobytes() == b.tobytes()


@pytest.mark.parametrize("kernel,ps,mat", [(0, 0, 0), (1, 0, 0), (2, 0, 0), (3, 2, 1), (3, 1, 0)])
def test_oracle_frames_match_reference(kernel, ps, mat):
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    sc = S.make_scene("cfg2_grid", 0.01)
    eye = (0.3, 1.2, 2.2)
    fp = S.frame_params(kernel, 64, 48, eye, (0, 0, 0), pixel_sampler=ps, num_bounces=4, ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0))
    mats, lights = S.default_material(mat), S.point_light(eye)
    a = O.oracle_frame(sc, fp, mats, lights)
    b, _, basis = O.ref_frame(sc, fp, mats, lights, mode=0)
    assert np.abs(a - b).max() <= 1e-6
    if kernel in (0, 1):   # RNG-free kernels: the reference's tiled_sched<ray8> is a second oracle
        c, _, _ = O.ref_frame(sc, fp, mats, lights, mode=1)
        assert np.abs(a - c).max() <= 1.0 / 255.0


@pytest.mark.parametrize("kernel,ps", [(0, 3), (1, 4), (0, 5), (2, 4), (3, 3)])
def test_oracle_ssaa_matches_reference(kernel, ps):
    """ssaa_type<2|4|8> (detail/sched_common.h:222-295,640-698): the restatement against the reference's
    own make_primary_rays(ssaa_type<N>) driven per pixel, and for the RNG-free kernels against the
    reference's tiled_sched<ray8> with the ssaa pixel sampler."""
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    sc = S.make_scene("cfg2_grid", 0.01)
    eye = (0.3, 1.2, 2.2)
    fp = S.frame_params(kernel, 48, 40, eye, (0, 0, 0), pixel_sampler=ps, num_bounces=4, ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0))
    mats, lights = S.default_material(1 if kernel == 3 else 0), S.point_light(eye)
    a = O.oracle_frame(sc, fp, mats, lights)
    b, _, _ = O.ref_frame(sc, fp, mats, lights, mode=0)
    assert np.abs(a - b).max() <= 1e-6
    if kernel in (0, 1):
        c, _, _ = O.ref_frame(sc, fp, mats, lights, mode=1)
        assert np.abs(a - c).max() <= 1.0 / 255.0
    # supersampling really happened: differs from the uniform frame on edges
    u = O.oracle_frame(sc, S.frame_params(kernel, 48, 40, eye, (0, 0, 0), pixel_sampler=0, num_bounces=4,
                                          ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0)), mats, lights)
    assert np.abs(a - u).max() > 1e-3


@pytest.mark.parametrize("kernel,ps", [(0, 0), (1, 0), (2, 0), (3, 2), (1, 4), (0, 1), (3, 3)])
def test_oracle_matrix_camera_and_depth_match_reference(kernel, ps):
    """sched_params with view / projection matrices (scheduler.h:60-96): primary rays from the
    inverse matrices (detail/sched_common.h:152-178), depth = hit ? depth_transform(isect_pos) : 1
    (:337-354, :425), stored / blended / SSAA-accumulated with the colour -- restatement against
    the reference headers, including the reference's own inverse() of the camera matrices."""
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    sc = S.make_scene("cfg2_grid", 0.01)
    eye = (0.3, 1.2, 2.2)
    amb = (1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0)
    mats, lights = S.default_material(1 if kernel == 3 else 0), S.point_light(eye)
    view, proj = O.ref_camera_matrices(S.frame_params(kernel, 48, 40, eye, (0, 0, 0)), 0.1, 10.0)
    a = d = b = e = None
    for frame_num in ((1, 2, 3) if ps == 2 else (1,)):
        fp = S.frame_params(kernel, 48, 40, eye, (0, 0, 0), pixel_sampler=ps, num_bounces=4, ambient=amb, frame_num=frame_num)
        a, d = O.oracle_frame_matrix(sc, fp, view, proj, mats, lights, rgba=a, depth=d)
        b, e, inv = O.ref_frame_matrix(sc, fp, view, proj, mats, lights, rgba=b, depth=e)
    assert O.oracle_mat4_inverse(view).tobytes() == inv[:16].tobytes()
    assert O.oracle_mat4_inverse(proj).tobytes() == inv[16:].tobytes()
    assert np.abs(a - b).max() <= 1e-6
    assert np.abs(d - e).max() <= 1e-6
    assert d.min() > 0.5 and d.max() == 1.0 and (d < 1.0).mean() > 0.2      # geometry in front of the far plane, sky at 1
    # the matrix camera sees what the look_at / perspective camera sees (same view, other arithmetic)
    if ps == 0:
        c = O.oracle_frame(sc, S.frame_params(kernel, 48, 40, eye, (0, 0, 0), pixel_sampler=0, num_bounces=4, ambient=amb), mats, lights)
        if kernel in (0, 1):
            assert np.mean(np.abs(a - c).max(axis=2) <= 2.0 / 255.0) > 0.97
    # no depth target: same colours
    a2, none = O.oracle_frame_matrix(sc, S.frame_params(kernel, 48, 40, eye, (0, 0, 0), pixel_sampler=ps, num_bounces=4, ambient=amb),
                                     view, proj, mats, lights, want_depth=False)
    assert none is None
    if ps != 2:
        assert a2.tobytes() == a.tobytes()


def test_host_camera_matrices_match_reference_camera():
    """visionaray_b200.scenes.camera_matrices restates camera::look_at / perspective
    (detail/camera.inl:10-65) bit for bit."""
    if not O.ref_available():
        pytest.skip("oracle/_ref not built (no /root/reference here)")
    for eye, center, up, zn, zf in (((0.3, 1.2, 2.2), (0, 0, 0), (0, 1, 0), 0.1, 10.0),
                                    ((-2.0, 0.4, 0.1), (0.2, 0.1, -0.3), (0, 0, 1), 0.001, 1000.0)):
        fp = S.frame_params(0, 64, 48, eye, center, up=up, fovy_deg=37.0)
        v, p = S.camera_matrices(fp, zn, zf)
        rv, rp = O.ref_camera_matrices(fp, zn, zf)
        assert v.tobytes() == rv.tobytes() and p.tobytes() == rp.tobytes()


def test_golden_vectors():
    """Fixtures generated from the reference itself by tests/golden/make_golden.py."""
    path = os.path.join(GOLD, "golden_queries.npz")
    g = np.load(path)
    for key in ("tri", "sph"):
        prims = g[f"{key}_prims"]; nodes = g[f"{key}_nodes"].view(NODE_DTYPE).reshape(-1); idx = g[f"{key}_indices"]
        sc = S.HostScene(0 if key == "tri" else 1, prims, nodes, idx, None, key)
        rays = g[f"{key}_rays"]
        for tag, any_hit, mt in (("closest", False, FLT_MAX), ("any", True, 0.4)):
            want = g[f"{key}_{tag}"].view(O.HIT_DTYPE).reshape(-1)
            got = O.oracle_query(sc, rays, any_hit, mt)
            assert got.tobytes() == want.tobytes(), f"{key} {tag}"
        # the host builder reproduces the reference's tree for the fixture's primitives
        n2, i2 = S.build_bvh(sc.prim_kind, prims)
        assert n2.tobytes() == nodes.tobytes() and i2.tobytes() == idx.tobytes()
    fr = np.load(os.path.join(GOLD, "golden_frames.npz"))
    sc = S.make_scene("cfg1_sphere", 0.01)
    eye = (0.0, 0.0, 3.0)
    for kernel, ps, mat in ((0, 0, 0), (1, 0, 0), (2, 0, 0), (3, 2, 1)):
        fp = S.frame_params(kernel, 32, 24, eye, (0, 0, 0), pixel_sampler=ps, ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0))
        img = O.oracle_frame(sc, fp, S.default_material(mat), S.point_light(eye))
        assert np.abs(img - fr[f"k{kernel}"]).max() <= 1e-6
    if "k1_ssaa4" in fr:
        fp = S.frame_params(1, 32, 24, eye, (0, 0, 0), pixel_sampler=4)
        img = O.oracle_frame(sc, fp, S.default_material(0), S.point_light(eye))
        assert np.abs(img - fr["k1_ssaa4"]).max() <= 1.0 / 255.0     # generated by tiled_sched<simd::ray8> (packets)
    fp = S.frame_params(1, 32, 24, eye, (0, 0, 0), pixel_sampler=3)
    view, proj = S.camera_matrices(fp, 0.1, 10.0)
    assert view.tobytes() == fr["matrix_view"].tobytes() and proj.tobytes() == fr["matrix_proj"].tobytes()
    img, z = O.oracle_frame_matrix(sc, fp, view, proj, S.default_material(0), S.point_light(eye))
    assert np.abs(img - fr["k1_matrix_ssaa2"]).max() <= 1e-6 and np.abs(z - fr["k1_matrix_ssaa2_depth"]).max() <= 1e-6


def test_brute_force_agrees_with_bvh():
    sc = S.make_scene("cfg5_soup", 0.0001)
    rays = S.random_rays(500, 5)
    a = O.oracle_query(sc, rays, False)
    b = O.oracle_brute(sc, rays, False)
    assert (a["hit"] == b["hit"]).all()
    assert np.allclose(a["t"], b["t"], rtol=1e-6)


def test_edge_cases():
    sc = two_triangles()
    rays = np.array([
        [0.5, -0.5, 2.0, 0, 0, 0, 1.0, 0],      # pointing away: miss
        [0.5, -0.5, 0.0, 0, 0, 0, -1.0, 0],     # starts between the triangles: hits the back one
        [5.0, 5.0, 2.0, 0, 0, 0, -1.0, 0],      # outside the root box: miss
        [0.5, -0.5, 1.0, 0, 0, 0, -1.0, 0],     # origin on the first triangle: t == 0 is a hit (t >= 0)
    ], dtype=np.float32)
    h = O.oracle_query(sc, rays, False)
    assert h["hit"].tolist() == [0, 1, 0, 1]
    assert h["prim_id"][1] == 1 and h["t"][3] == 0.0 and h["prim_id"][3] == 0
    assert h["t"][0] == np.float32(FLT_MAX)
    # max_t cuts the hit (strict '<', update_if.h:75-79)
    assert O.oracle_query(sc, rays[1:2], False, 1.0)["hit"][0] == 0
    assert O.oracle_query(sc, rays[1:2], False, np.float32(1.0000001))["hit"][0] == 1
    # empty ray list
    assert O.oracle_query(sc, np.zeros((0, 8), dtype=np.float32)).shape[0] == 0
