"""Shared definition of the get_surface() parity cases: used by
tests/golden/make_golden.py (fixtures from the reference) and tests/test_surface.py."""
import numpy as np

from visionaray_b200 import scenes as S

N_GEOMS = 4

# name: (scene, kernel, normal_binding, color_binding, textured, tex_filter)
CASES = {
    "simple_vertex":          ("sphere", 0, 1, 0, False, 0),
    "simple_face_tex_linear": ("grid",   0, 0, 0, True,  1),
    "simple_vertex_cface":    ("sphere", 0, 1, 1, False, 0),
    "whitted_vertex_cvert_tex": ("grid", 1, 1, 2, True,  1),
    "whitted_face_generic":   ("grid",   1, 0, 0, False, 0),
    "whitted_vertex_tex_nearest": ("sphere", 1, 1, 0, True, 0),
    "path_vertex_tex":        ("grid",   3, 1, 0, True,  1),
    "whitted_face_tex_bspline": ("grid", 1, 0, 0, True,  2),
    "simple_vertex_tex_cardinal": ("sphere", 0, 1, 1, True, 4),
    "path_face_cvert":        ("sphere", 3, 0, 2, False, 0),
}

_scenes = {}


def scene(name):
    """cfg1 sphere / cfg2 terrain at oracle size; geom_id = prim index mod 4 (one material of each kind)."""
    if name not in _scenes:
        sc = S.make_scene("cfg1_sphere" if name == "sphere" else "cfg2_grid", 0.01 if name == "sphere" else 0.004)
        sc.prims.view(np.uint32)[:, 0] = (np.arange(sc.n_prims) // 3) % N_GEOMS
        _scenes[name] = sc
    return _scenes[name]


def case(name, width=40, height=28):
    scn, kernel, nb, cb, tex, filt = CASES[name]
    sc = scene(scn)
    eye, center = ((0.3, 0.5, 3.0), (0.0, 0.0, 0.0)) if scn == "sphere" else ((0.9, 0.75, 1.3), (0.0, 0.0, 0.0))
    fp = S.frame_params(kernel, width, height, eye, center, pixel_sampler=2 if kernel == 3 else 0, frame_num=1,
                        n_materials=N_GEOMS, num_bounces=4, epsilon=1e-3,
                        ambient=(1, 1, 1, 1) if kernel == 3 else (0.3, 0.3, 0.3, 1.0))
    sd = S.make_surface(sc, nb, cb, tex, n_geoms=N_GEOMS, tex_filter=filt, tex_address=(0, 1) if filt else (2, 0))
    light = S.point_light((1.0, 2.0, 3.0) if scn == "sphere" else (0.4, 1.2, 0.6))
    return sc, fp, sd, S.generic_materials(), light


def tex_samples():
    """(texture, coords) pairs for the tex2D known-answer vectors: every filter x address mode."""
    rng = np.random.default_rng(17)
    out = []
    for filt in (0, 1, 2, 4):
        for am in ((0, 0), (1, 1), (2, 2), (0, 2), (1, 0), (3, 3)):      # 3 = Border: clamps, like the reference
            tex = S.checker_texture(13, 7, 5, filt, am)
            st = rng.uniform(-3.0, 3.0, (64, 2)).astype(np.float32)
            out.append((f"f{filt}_a{am[0]}{am[1]}", tex, st))
    return out
