"""get_surface() with normal / colour bindings, tex coords + textures and the generic material list
(kernels.h:194-509, get_surface.h:330-530, texture/detail/sampler2d.h, generic_material.h).

CPU: the oracle restatement against fixtures generated from the reference headers
(tests/golden/golden_surface.npz, make_golden.py) and, where oracle/_ref exists, against the
reference itself on more cases.  GPU (-m gpu): vsb_scene_set_surface + vsb_frame against the oracle."""
import os

import numpy as np
import pytest

from oracle import bindings as O
from visionaray_b200 import scenes as S

import surface_cases as SC

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_surface.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLDEN)


@pytest.mark.parametrize("name", list(SC.CASES))
def test_oracle_matches_reference_fixture(golden, name):
    sc, fp, sd, mats, light = SC.case(name)
    got = O.oracle_frame_surface(sc, fp, sd, mats, light)
    want = golden[name]
    # same libm on both sides in the build container; 1e-6 leaves room for another libm's powf / sinf
    assert np.abs(got - want).max() <= 1e-6, name
    assert want[..., :3].std() > 0.01          # the fixture is not a constant image


def test_tex2d_known_answers(golden):
    """tex2D on vector<4, unorm<8>> texels: nearest / linear x wrap / mirror / clamp, bit for bit
    (the filtered value is truncated to an integer before normalisation)."""
    for key, tex, st in SC.tex_samples():
        got = np.stack([O.oracle_tex2d(tex, float(s), float(t)) for s, t in st])
        assert got.tobytes() == golden["tex_" + key].tobytes(), key


def test_surface_changes_the_image():
    """Each ingredient is live: bindings, colours, textures and material kinds all move pixels."""
    sc, fp, sd, mats, light = SC.case("whitted_vertex_cvert_tex")
    base = O.oracle_frame_surface(sc, fp, sd, mats, light)
    flat = S.make_surface(sc, 0, 0, False, n_geoms=SC.N_GEOMS)
    assert np.abs(O.oracle_frame_surface(sc, fp, flat, mats, light) - base).max() > 0.02
    nocol = S.make_surface(sc, 1, 0, True, n_geoms=SC.N_GEOMS)
    assert np.abs(O.oracle_frame_surface(sc, fp, nocol, mats, light) - base).max() > 0.02
    plastic_only = np.repeat(S.default_material(0), SC.N_GEOMS)
    assert np.abs(O.oracle_frame_surface(sc, fp, sd, plastic_only, light) - base).max() > 0.02
    # per-face binding with the scene's own normals and no colours == the plain vso_frame path
    a = O.oracle_frame_surface(sc, fp, flat, plastic_only, light)
    b = O.oracle_frame(sc, fp, plastic_only, light)
    assert a.tobytes() == b.tobytes()


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref is only built where /root/reference exists")
@pytest.mark.parametrize("kernel", [0, 1, 3])
@pytest.mark.parametrize("nb", [0, 1])
@pytest.mark.parametrize("cb,tex", [(0, False), (0, True), (1, False), (2, True)])
def test_oracle_matches_reference_headers(kernel, nb, cb, tex):
    sc = SC.scene("grid")
    fp = S.frame_params(kernel, 72, 48, (0.9, 0.75, 1.3), (0, 0, 0), pixel_sampler=2 if kernel == 3 else 0,
                        n_materials=SC.N_GEOMS, num_bounces=5, ambient=(1, 1, 1, 1) if kernel == 3 else (0.2, 0.2, 0.2, 1))
    sd = S.make_surface(sc, nb, cb, tex, n_geoms=SC.N_GEOMS, seed=40 + kernel, tex_filter=nb, tex_address=(nb, 2 - cb))
    mats, light = S.generic_materials(), S.point_light((0.4, 1.2, 0.6))
    got = O.oracle_frame_surface(sc, fp, sd, mats, light)
    want = O.ref_frame_surface(sc, fp, sd, mats, light)
    assert np.abs(got - want).max() <= 1e-6


def prim_normal_scene(name):
    if name == "spheres":
        sc = S.make_scene("spheres", 0.02)
        eye, light = (0.0, 0.3, 3.2), S.point_light((1.0, 2.0, 3.0))
    else:
        sc = S.make_scene("cfg2_grid", 0.004)
        eye, light = (0.9, 0.75, 1.3), S.point_light((0.4, 1.2, 0.6))
    sc.prims.view(np.uint32)[:, 0] = (np.arange(sc.n_prims) // 3) % SC.N_GEOMS
    return sc, eye, light


PRIM_NORMAL_CASES = [("cfg2_grid", 0), ("cfg2_grid", 1), ("cfg2_grid", 3), ("spheres", 0), ("spheres", 1)]


def prim_normal_frame(name, kernel, w, h):
    sc, eye, light = prim_normal_scene(name)
    fp = S.frame_params(kernel, w, h, eye, (0, 0, 0), pixel_sampler=2 if kernel == 3 else 0, n_materials=SC.N_GEOMS,
                        num_bounces=4, ambient=(1, 1, 1, 1) if kernel == 3 else (0.3, 0.3, 0.3, 1.0))
    return sc, fp, S.SurfaceData(2, None), S.generic_materials(), light


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref is only built where /root/reference exists")
@pytest.mark.parametrize("name,kernel", PRIM_NORMAL_CASES)
def test_normals_from_the_primitive_match_reference_headers(name, kernel):
    """kernel_params without normals (default make_kernel_params overload): get_normal(hr, triangle | sphere)."""
    sc, fp, sd, mats, light = prim_normal_frame(name, kernel, 96, 64)
    got = O.oracle_frame_surface(sc, fp, sd, mats, light)
    want = O.ref_frame_prim_normals(sc, fp, mats, light)
    assert np.abs(got - want).max() <= 1e-6 and want[..., :3].std() > 0.05


@pytest.mark.parametrize("name,kernel", PRIM_NORMAL_CASES)
def test_normals_from_the_primitive_fixture(golden, name, kernel):
    sc, fp, sd, mats, light = prim_normal_frame(name, kernel, 40, 28)
    got = O.oracle_frame_surface(sc, fp, sd, mats, light)
    assert np.abs(got - golden[f"primnormal_{name}_k{kernel}"]).max() <= 1e-6


# ---- GPU ------------------------------------------------------------------------------------------

def _frac_bad(img, want):
    return float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(SC.CASES))
def test_gpu_frames_match_oracle(gpu, name):
    sc, fp, sd, mats, light = SC.case(name, 160, 112)
    want = O.oracle_frame_surface(sc, fp, sd, mats, light)
    dev = gpu.Scene(sc)
    try:
        gpu.set_surface(dev, sd)
        for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
            img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
            gpu.frame(dev, fp, None, mats, light, img, traverse=trav)
            bad = _frac_bad(img, want)
            # RNG-free kernels on the reference-order walk reproduce every pixel; the wide walk may take
            # the other triangle of an exact t-tie on a shared edge, the path tracer may fork on an ulp
            limit = 0.0 if trav == gpu.TRAVERSE_EXACT else (1e-3 if fp.kernel in (0, 1) else 3e-3)
            assert bad <= limit, (name, trav, bad, float(np.abs(img - want).max()))
        # detaching restores the per-face path
        gpu.set_surface(dev, None)
        plastic_only = np.repeat(S.default_material(0), SC.N_GEOMS)
        img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, plastic_only, light, img, traverse=gpu.TRAVERSE_EXACT)
        assert _frac_bad(img, O.oracle_frame(sc, fp, plastic_only, light)) <= (0.0 if fp.kernel in (0, 1) else 3e-3)
    finally:
        dev.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,ps", [("whitted_vertex_cvert_tex", 4), ("simple_face_tex_linear", 3), ("path_vertex_tex", 5)])
def test_gpu_surface_with_ssaa(gpu, name, ps):
    """Surface description AND an SSAA pixel sampler: the instantiation that carries both extensions."""
    sc, fp, sd, mats, light = SC.case(name, 128, 96)
    fp.pixel_sampler = ps
    want = O.oracle_frame_surface(sc, fp, sd, mats, light)
    dev = gpu.Scene(sc)
    try:
        gpu.set_surface(dev, sd)
        img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, mats, light, img, traverse=gpu.TRAVERSE_EXACT)
        assert _frac_bad(img, want) <= (0.0 if fp.kernel in (0, 1) else 4e-3)
    finally:
        dev.close()


@pytest.mark.gpu
@pytest.mark.parametrize("name,kernel", PRIM_NORMAL_CASES)
def test_gpu_normals_from_the_primitive(gpu, name, kernel):
    """Triangle and SPHERE scenes through vsb_frame with kernel_params that carry no normals."""
    sc, fp, sd, mats, light = prim_normal_frame(name, kernel, 160, 112)
    want = O.oracle_frame_surface(sc, fp, sd, mats, light)
    dev = gpu.Scene(sc)
    try:
        gpu.set_surface(dev, sd)
        for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
            img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
            gpu.frame(dev, fp, None, mats, light, img, traverse=trav)
            limit = 0.0 if trav == gpu.TRAVERSE_EXACT else (1e-3 if kernel in (0, 1) else 4e-3)
            assert _frac_bad(img, want) <= limit, (name, kernel, trav)
        if name == "spheres":
            fp.kernel = 3          # the reference's path tracer has no defined sphere normal (isect_pos unset)
            with pytest.raises(RuntimeError):
                gpu.frame(dev, fp, None, mats, light, img)
    finally:
        dev.close()


@pytest.mark.gpu
def test_gpu_border_addressing_clamps(gpu):
    """tex_address_mode Border falls through to Clamp in the reference (filter/common.h:41-46)."""
    sc, fp, sd, mats, light = SC.case("simple_face_tex_linear", 128, 96)
    for t in sd.textures:
        if t is not None:
            t.address = (3, 3)
    want = O.oracle_frame_surface(sc, fp, sd, mats, light)
    clamp = S.make_surface(sc, 0, 0, True, n_geoms=SC.N_GEOMS, tex_filter=1, tex_address=(2, 2))
    assert want.tobytes() == O.oracle_frame_surface(sc, fp, clamp, mats, light).tobytes()
    dev = gpu.Scene(sc)
    try:
        gpu.set_surface(dev, sd)
        img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, mats, light, img, traverse=gpu.TRAVERSE_EXACT)
        assert _frac_bad(img, want) == 0.0
    finally:
        dev.close()


@pytest.mark.gpu
def test_gpu_ids_are_checked_against_array_sizes(gpu):
    """prim_id indexes normals / colours / tex coords, geom_id the materials: the reference trusts them, the
    library computes the scene's largest ids once and refuses arrays they would leave."""
    sc, fp, sd, mats, light = SC.case("simple_vertex", 64, 48)
    img = np.zeros((48, 64, 4), dtype=np.float32)
    dev = gpu.Scene(sc)
    try:
        with pytest.raises(RuntimeError):
            gpu.frame(dev, S.frame_params(0, 64, 48, (0.3, 0.5, 3.0), (0, 0, 0), n_materials=2), None, mats[:2], light, img)   # geom ids reach 3
    finally:
        dev.close()
    bad = S.HostScene(sc.prim_kind, sc.prims.copy(), sc.nodes, sc.indices, None, "bad")
    bad.prims.view(np.uint32)[5, 1] = sc.n_prims + 7          # a prim_id past the per-primitive arrays
    dev = gpu.Scene(bad)
    try:
        with pytest.raises(RuntimeError):
            gpu.set_surface(dev, sd)
        with pytest.raises(RuntimeError):
            dev.set_normals(sc.normals)
        assert gpu.closest_hit(dev, S.random_rays(100, 1, -1.5, 1.5)).shape[0] == 100     # ray queries do not care
    finally:
        dev.close()


@pytest.mark.gpu
def test_gpu_generic_materials_without_surface(gpu):
    """A generic_material list with the plain per-face normals (no vsb_scene_set_surface): device
    materials need the generic_materials flag, host materials are detected."""
    import torch

    sc, fp, sd, mats, light = SC.case("whitted_face_generic", 128, 96)
    want = O.oracle_frame(sc, fp, mats, light)
    dev = gpu.Scene(sc)
    try:
        img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
        gpu.frame(dev, fp, None, mats, light, img, traverse=gpu.TRAVERSE_EXACT)
        assert _frac_bad(img, want) == 0.0
        d_m = torch.from_numpy(mats.view(np.uint8)).cuda()
        d_l = torch.from_numpy(light.view(np.uint8)).cuda()
        d_img = torch.zeros((fp.height, fp.width, 4), dtype=torch.float32, device="cuda")
        gpu.frame(dev, fp, None, d_m, d_l, d_img, traverse=gpu.TRAVERSE_EXACT, generic_materials=True)
        gpu.sync()
        assert _frac_bad(d_img.cpu().numpy(), want) == 0.0
    finally:
        dev.close()


@pytest.mark.gpu
def test_gpu_surface_full_size_properties(gpu):
    """cfg2-sized scene (1.0 M triangles, 2048 x 2048): wide == reference-order walk except tie pixels,
    and the per-face surface with no colours reproduces the plain frame bit for bit."""
    sc = S.make_scene("cfg2_grid", 1.0)
    sc.prims.view(np.uint32)[:, 0] = (np.arange(sc.n_prims) // 3) % SC.N_GEOMS
    fp = S.frame_params(1, 2048, 2048, (0.9, 0.75, 1.3), (0, 0, 0), n_materials=SC.N_GEOMS, num_bounces=4,
                        ambient=(0.3, 0.3, 0.3, 1.0))
    mats, light = S.generic_materials(), S.point_light((0.4, 1.2, 0.6))
    dev = gpu.Scene(sc)
    try:
        sd = S.make_surface(sc, 1, 2, True, n_geoms=SC.N_GEOMS)
        gpu.set_surface(dev, sd)
        a = np.zeros((2048, 2048, 4), dtype=np.float32); b = np.zeros_like(a)
        gpu.frame(dev, fp, None, mats, light, a, traverse=gpu.TRAVERSE_WIDE)
        gpu.frame(dev, fp, None, mats, light, b, traverse=gpu.TRAVERSE_EXACT)
        assert _frac_bad(a, b) <= 1e-3
        assert np.isfinite(a).all() and a[..., :3].std() > 0.01
        flat = S.make_surface(sc, 0, 0, False, n_geoms=SC.N_GEOMS)
        gpu.set_surface(dev, flat)
        gpu.frame(dev, fp, None, mats, light, a, traverse=gpu.TRAVERSE_EXACT)
        gpu.set_surface(dev, None)
        gpu.frame(dev, fp, None, mats, light, b, traverse=gpu.TRAVERSE_EXACT)
        assert a.tobytes() == b.tobytes()
    finally:
        dev.close()
