"""Generates tests/golden/*.npz FROM THE REFERENCE ITSELF (oracle/_ref/libvsnray_ref.so, i.e. the
unmodified headers under /root/reference/include compiled by oracle/Makefile).  Run in the
build container only:   python tests/golden/make_golden.py
The fixtures pin the oracle restatement where /root/reference is absent (the GPU box).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import bindings as O  # noqa: E402
from visionaray_b200 import scenes as S  # noqa: E402
from visionaray_b200._native import NODE_DTYPE  # noqa: E402

FLT_MAX = 3.4028234663852886e38
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    assert O.ref_available(), "oracle/_ref missing: run `make -C oracle ref` where /root/reference exists"
    out = {}
    tris = S.gen_triangles(5, 600, 0, 21, 0.08)                 # small soup
    sph = S.gen_spheres(300, 22, 0.02, 0.12)
    for key, kind, prims, lo, hi in (("tri", 0, tris, -0.2, 1.2), ("sph", 1, sph, -1.3, 1.3)):
        nodes, idx = O.ref_build(kind, prims)                    # the REFERENCE builder
        sc = S.HostScene(kind, prims, nodes, idx, None, key)
        rays = S.random_rays(2000, 23, lo, hi)
        out[f"{key}_prims"] = prims
        out[f"{key}_nodes"] = nodes.view(np.uint8)
        out[f"{key}_indices"] = idx
        out[f"{key}_rays"] = rays
        out[f"{key}_closest"] = O.ref_query(sc, rays, False, FLT_MAX).view(np.uint8)
        out[f"{key}_any"] = O.ref_query(sc, rays, True, 0.4).view(np.uint8)
    np.savez_compressed(os.path.join(HERE, "golden_queries.npz"), **out)

    sc = S.make_scene("cfg1_sphere", 0.01)
    eye = (0.0, 0.0, 3.0)
    frames = {}
    for kernel, ps, mat in ((0, 0, 0), (1, 0, 0), (2, 0, 0), (3, 2, 1)):
        fp = S.frame_params(kernel, 32, 24, eye, (0, 0, 0), pixel_sampler=ps, ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0))
        img, _, _ = O.ref_frame(sc, fp, S.default_material(mat), S.point_light(eye), mode=0)
        frames[f"k{kernel}"] = img
    # ssaa_type<4> through the reference's own tiled_sched<simd::ray8> (RNG-free kernel)
    fp = S.frame_params(1, 32, 24, eye, (0, 0, 0), pixel_sampler=4)
    frames["k1_ssaa4"], _, _ = O.ref_frame(sc, fp, S.default_material(0), S.point_light(eye), mode=1)
    # sched_params from view / projection matrices + depth buffer (scalar driver over the reference's
    # make_primary_ray_impl(matrices) / depth_transform); the matrices are the reference camera's
    fp = S.frame_params(1, 32, 24, eye, (0, 0, 0), pixel_sampler=3)
    view, proj = O.ref_camera_matrices(fp, 0.1, 10.0)
    frames["matrix_view"], frames["matrix_proj"] = view, proj
    frames["k1_matrix_ssaa2"], frames["k1_matrix_ssaa2_depth"], _ = O.ref_frame_matrix(sc, fp, view, proj, S.default_material(0), S.point_light(eye))
    np.savez_compressed(os.path.join(HERE, "golden_frames.npz"), **frames)
    print("wrote golden_queries.npz, golden_frames.npz")
    multi_hit_vectors()
    surface_vectors()
    sbvh_vectors()
    format_vectors()
    generic_vectors()


def multi_hit_vectors():
    """multi_hit<4> / multi_hit<16> of the reference on the scenes and rays of golden_queries.npz."""
    g = np.load(os.path.join(HERE, "golden_queries.npz"))
    out = {}
    for key, kind in (("tri", 0), ("sph", 1)):
        sc = S.HostScene(kind, g[f"{key}_prims"], g[f"{key}_nodes"].view(NODE_DTYPE).reshape(-1), g[f"{key}_indices"], None, key)
        rays = g[f"{key}_rays"][:600]
        for n in (4, 16):
            out[f"{key}_multi{n}"] = O.ref_multi_hit(sc, rays, n).view(np.uint8)
    np.savez_compressed(os.path.join(HERE, "golden_multi_hit.npz"), **out)
    print("wrote golden_multi_hit.npz")


def surface_vectors():
    """Frames of the reference kernels with kernel_params that carry a surface description
    (normal / colour bindings, tex coords + textures, generic_material list) and tex2D samples,
    through the unmodified headers: make_kernel_params / get_surface / tex2D."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import surface_cases as SC
    out = {}
    for name in SC.CASES:
        sc, fp, sd, mats, light = SC.case(name)
        out[name] = O.ref_frame_surface(sc, fp, sd, mats, light)
    import test_surface as TS
    for name, kernel in TS.PRIM_NORMAL_CASES:            # kernel_params without normals, triangle and sphere scenes
        sc, fp, sd, mats, light = TS.prim_normal_frame(name, kernel, 40, 28)
        out[f"primnormal_{name}_k{kernel}"] = O.ref_frame_prim_normals(sc, fp, mats, light)
    import math
    import test_lights as TL
    for kernel in (0, 1):                                # spot_light list
        sc, fp, mats, lights = TL.case(kernel, 40, 28)
        out[f"spot_k{kernel}"] = O.ref_frame_spot(sc, fp, mats, lights, [math.radians(a) for a in TL.ANGLES])
    for key, tex, st in SC.tex_samples():
        out["tex_" + key] = np.stack([O.ref_tex2d(tex, float(s), float(t)) for s, t in st])
    np.savez_compressed(os.path.join(HERE, "golden_surface.npz"), **out)
    print("wrote golden_surface.npz")


FORMAT_CASES = [(0, 0, 0), (1, 0, 0), (1, 4, 0), (3, 2, 1), (1, 5, 0)]     # (kernel, pixel sampler, material kind)


def format_case(kernel, ps, mat, width=40, height=28):
    sc = S.make_scene("cfg2_grid", 0.004)
    fp = S.frame_params(kernel, width, height, (0.9, 0.75, 1.3), (0, 0, 0), pixel_sampler=ps, num_bounces=4,
                        ambient=(1, 1, 1, 1) if kernel == 3 else (0.1, 0.1, 0.1, 1.0), bg=(0.1, 0.4, 1.0, 0.8))
    return sc, fp, S.default_material(mat), S.point_light((0.4, 1.2, 0.6))


def format_vectors():
    """Frames into PF_RGBA8 / PF_RGB32F / PF_RGB8 targets through the reference's own pixel_access::store /
    blend and colour conversion (3 accumulated frames for jittered_blend)."""
    out = {}
    for fmt in ("PF_RGBA8", "PF_RGB32F", "PF_RGB8"):
        for kernel, ps, mat in FORMAT_CASES:
            sc, fp, mats, light = format_case(kernel, ps, mat)
            buf = O.format_buffer(fmt, fp.width, fp.height)
            for fn in ((1, 2, 3) if ps == 2 else (1,)):
                fp.frame_num = fn
                O.ref_frame_format(sc, fp, mats, light, fmt, buf)
            out[f"{fmt}_k{kernel}_ps{ps}"] = buf
    np.savez_compressed(os.path.join(HERE, "golden_formats.npz"), **out)
    print("wrote golden_formats.npz")


def generic_vectors():
    """index_bvh<generic_primitive<basic_triangle, basic_sphere>>: records laid out by the reference's own
    constructors, tree by its builder, closest_hit / any_hit / multi_hit<8> records by its traversal."""
    sc = S.make_scene("generic", 0.05)
    w = sc.prims.view(np.uint32)
    tags = (w[:, 16] == 2).astype(np.int32)
    rec = O.ref_make_generic(tags, np.ascontiguousarray(sc.prims[:, :16]), np.ascontiguousarray(sc.prims[:, :12]))
    nodes, idx = O.ref_build(2, sc.prims)
    ref = S.HostScene(2, sc.prims, nodes, idx, None, "generic")
    lo, hi = ref.bounds()
    rays = S.random_rays(6000, 9, float(lo.min()) - 0.1, float(hi.max()) + 0.1)
    out = {"ref_records": rec.view(np.uint32), "nodes": nodes.view(np.uint8), "indices": idx, "rays": rays,
           "closest": O.ref_query(ref, rays, False, FLT_MAX).view(np.uint8),
           "any": O.ref_query(ref, rays, True, 0.3).view(np.uint8),
           "multi8": O.ref_multi_hit(ref, rays[:1500], 8).view(np.uint8)}
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import test_generic_primitive as TG
    for kernel, nb in TG.FRAME_CASES:                  # frames: get_surface() over the variant with a normals array
        fsc, fp, sd, mats, light = TG.frame_case(kernel, nb, 40, 28)
        out[f"frame_k{kernel}_nb{nb}"] = O.ref_frame_generic(fsc, fp, sd, mats, light)
    np.savez_compressed(os.path.join(HERE, "golden_generic.npz"), **out)
    print("wrote golden_generic.npz:", sc.n_prims, "primitives,", int(tags.sum()), "spheres")


def sbvh_scene_prims():
    """Long thin triangles (a soup stretched 30x along x): the case spatial splits exist for."""
    t = S.gen_triangles(5, 3000, 0, 31, 0.05).copy()
    for o in (4, 8, 12):
        t[:, o] *= 30.0
    return t


def sbvh_rays(n=6000):
    rays = S.random_rays(n, 5, -0.2, 1.2)
    rays[:, 0] *= 30.0
    return rays


def sbvh_vectors():
    """A tree of the reference's builder WITH spatial splits (build<index_bvh<P>>(first, last, true),
    detail/bvh/sah.h:430-615: primitives referenced from several leaves, indices > primitives) and the
    reference's closest_hit / any_hit / multi_hit<16> records on it."""
    tris, rays = sbvh_scene_prims(), sbvh_rays()
    nodes, idx = O.ref_build(0, tris, True)
    assert idx.shape[0] > tris.shape[0] * 3 // 2
    sc = S.HostScene(0, tris, nodes, idx, None, "sbvh")
    out = {"nodes": nodes.view(np.uint8), "indices": idx,
           "closest": O.ref_query(sc, rays, False, FLT_MAX).view(np.uint8),
           "any": O.ref_query(sc, rays, True, 0.5).view(np.uint8),
           "multi16": O.ref_multi_hit(sc, rays[:1500], 16).view(np.uint8)}
    np.savez_compressed(os.path.join(HERE, "golden_sbvh.npz"), **out)
    print("wrote golden_sbvh.npz:", nodes.shape[0], "nodes,", idx.shape[0], "indices for", tris.shape[0], "triangles")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "multi_hit":
        multi_hit_vectors()          # adds golden_multi_hit.npz without touching the older fixtures
    elif len(sys.argv) > 1 and sys.argv[1] == "surface":
        surface_vectors()
    elif len(sys.argv) > 1 and sys.argv[1] == "sbvh":
        sbvh_vectors()
    elif len(sys.argv) > 1 and sys.argv[1] == "generic":
        generic_vectors()
    elif len(sys.argv) > 1 and sys.argv[1] == "formats":
        format_vectors()
    else:
        main()
