"""The host C++ adaptor (include/visionaray_b200/scheduler.h): a reference-style program compiled
with g++ against libvsnray_b200.so.  CPU tier: it builds, links, and fails loudly without a GPU.
GPU tier: its hit records and frames match the oracle on the same seeded scene."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import classify_closest
from oracle import bindings as O
from visionaray_b200 import scenes as S

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "cpp", "dropin_example")


def run(outdir):
    env = dict(os.environ, LD_LIBRARY_PATH=os.path.join(ROOT, "visionaray_b200", "lib") + ":" + os.environ.get("LD_LIBRARY_PATH", ""))
    obj = os.path.join(ROOT, "tests", "golden", "obj", "smooth.obj")
    return subprocess.run([EXE, outdir, obj], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)


def test_builds_and_fails_loudly_without_gpu():
    assert os.path.exists(EXE), "tests/cpp/dropin_example not built (__graft_entry__.build())"
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    with tempfile.TemporaryDirectory() as d:
        r = run(d)
    assert r.returncode == 3 and "no CUDA device" in r.stdout, r.stdout


@pytest.mark.gpu
def test_dropin_program_matches_oracle(gpu):
    with tempfile.TemporaryDirectory() as d:
        r = run(d)
        assert r.returncode == 0, r.stdout
        sc = S.make_scene("cfg1_sphere", (50 / 224) ** 2)     # uv_sphere(50, 50)
        assert sc.n_prims == 2 * 50 * 49
        rays = S.random_rays(20000, 3, -1.5, 1.5)
        want = O.oracle_query(sc, rays, False)
        exact = np.fromfile(os.path.join(d, "hits_exact.bin"), dtype=O.HIT_DTYPE)
        wide = np.fromfile(os.path.join(d, "hits_wide.bin"), dtype=O.HIT_DTYPE)
        occ = np.fromfile(os.path.join(d, "hits_any.bin"), dtype=O.HIT_DTYPE)
        assert exact.tobytes() == want.tobytes()
        assert classify_closest(want, wide)["bad"] == 0
        assert (occ["hit"] == O.oracle_query(sc, rays, True, 0.5)["hit"]).all()
        multi = np.fromfile(os.path.join(d, "hits_multi4.bin"), dtype=O.HIT_DTYPE).reshape(-1, 4)     # multi_hit<4>(), reference order
        assert multi.tobytes() == O.oracle_multi_hit(sc, rays, 4).tobytes()
        built = np.fromfile(os.path.join(d, "hits_gpu_built.bin"), dtype=O.HIT_DTYPE)     # build<b200_index_bvh<P>>()
        assert classify_closest(want, built)["bad"] == 0
        eye = (0.0, 0.0, 3.0)
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), num_bounces=4)
        ref = O.oracle_frame(sc, fp, S.default_material(0), S.point_light(eye))
        img = np.fromfile(os.path.join(d, "frame_whitted.bin"), dtype=np.float32).reshape(120, 160, 4)
        assert float((np.abs(img - ref).max(axis=2) > 1 / 255).mean()) <= 2e-4
        acc = np.zeros((120, 160, 4), dtype=np.float32)
        for frame_num in (1, 2, 3):
            fp = S.frame_params(3, 160, 120, eye, (0, 0, 0), pixel_sampler=2, frame_num=frame_num, num_bounces=4, ambient=(1, 1, 1, 1))
            O.oracle_frame(sc, fp, S.default_material(1), S.point_light(eye), rgba=acc)
        pt = np.fromfile(os.path.join(d, "frame_pt.bin"), dtype=np.float32).reshape(120, 160, 4)
        assert float((np.abs(pt - acc).max(axis=2) > 1 / 255).mean()) <= 5e-3
        # matrix sched_params + depth buffer, ssaa_type<4>: the adaptor's camera matrices are the reference camera's
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), pixel_sampler=4, num_bounces=4)
        view, proj = S.camera_matrices(fp, 0.1, 10.0)
        want_c, want_z = O.oracle_frame_matrix(sc, fp, view, proj, S.default_material(0), S.point_light(eye))
        mc = np.fromfile(os.path.join(d, "frame_matrix_color.bin"), dtype=np.float32).reshape(120, 160, 4)
        mz = np.fromfile(os.path.join(d, "frame_matrix_depth.bin"), dtype=np.float32).reshape(120, 160)
        assert float((np.abs(mc - want_c).max(axis=2) > 1 / 255).mean()) <= 2e-4
        assert float(np.abs(mz - want_z).max()) <= 1e-5
        # kernel_params with normals / colours per vertex, tex coords + textures and a generic_material list:
        # the program's own arrays go to the oracle
        from visionaray_b200._native import MATERIAL_DTYPE
        ss = S.make_scene("cfg1_sphere", (50 / 224) ** 2)
        ss.prims.view(np.uint32)[:, 0] = (np.arange(ss.n_prims) // 3) % 4
        texels = np.fromfile(os.path.join(d, "surface_texels.bin"), dtype=np.uint8).reshape(16, 24, 4)
        sd = S.SurfaceData(1, np.fromfile(os.path.join(d, "surface_normals.bin"), dtype=np.float32).reshape(-1, 4), 2,
                           np.fromfile(os.path.join(d, "surface_colors.bin"), dtype=np.float32).reshape(-1, 4),
                           np.fromfile(os.path.join(d, "surface_tex_coords.bin"), dtype=np.float32).reshape(-1, 2),
                           [S.Texture(texels, 1, (0, 1)), S.Texture(texels, 0, (2, 2)), None, None])
        mats = np.fromfile(os.path.join(d, "surface_materials.bin"), dtype=MATERIAL_DTYPE)
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), num_bounces=4, n_materials=4, ambient=(0.3, 0.3, 0.3, 1.0))
        want_s = O.oracle_frame_surface(ss, fp, sd, mats, S.point_light(eye))
        got_s = np.fromfile(os.path.join(d, "surface_whitted.bin"), dtype=np.float32).reshape(120, 160, 4)
        assert float((np.abs(got_s - want_s).max(axis=2) > 1 / 255).mean()) == 0.0      # reference-order walk
        # load_obj -> build -> make_kernel_params(normals_per_vertex_binding, ...) -> frame(): the oracle is fed the arrays
        # the same loader returns through the host library (tests/test_obj_loader.py pins them)
        m = S.load_obj(os.path.join(ROOT, "tests", "golden", "obj", "smooth.obj"))
        osc, osd = m.host_scene(), m.surface()
        ofp = S.frame_params(1, 224, 160, (1.6, 2.1, 2.9), (0.0, 0.6, 0.0), num_bounces=3, n_materials=len(m.materials),
                             ambient=(0.2, 0.2, 0.2, 1.0))
        want_o = O.oracle_frame_surface(osc, ofp, osd, m.materials, S.point_light((2.5, 4.0, 3.0)))
        got_o = np.fromfile(os.path.join(d, "frame_obj.bin"), dtype=np.float32).reshape(160, 224, 4)
        assert float(np.abs(got_o - want_o).max()) <= 1 / 255                             # reference-order walk: every pixel
        # PF_RGBA8 host target and PF_RGB8 device target (SSAA 2x), reference-order walk: byte for byte
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), num_bounces=4)
        want8 = O.oracle_frame_format(sc, fp, S.default_material(0), S.point_light(eye), "PF_RGBA8")
        got8 = np.fromfile(os.path.join(d, "frame_whitted_rgba8.bin"), dtype=np.uint8).reshape(120, 160, 4)
        assert int(np.abs(got8.astype(np.int32) - want8.astype(np.int32)).max()) <= 1 and float((got8 != want8).mean()) <= 1e-3
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), pixel_sampler=3, num_bounces=4)
        want3 = O.oracle_frame_format(sc, fp, S.default_material(0), S.point_light(eye), "PF_RGB8")
        got3 = np.fromfile(os.path.join(d, "frame_whitted_ssaa2_rgb8.bin"), dtype=np.uint8).reshape(120, 160, 3)
        assert int(np.abs(got3.astype(np.int32) - want3.astype(np.int32)).max()) <= 1 and float((got3 != want3).mean()) <= 1e-3
        # sphere bvh, kernel_params without normals: the program's spheres through the oracle
        balls = np.fromfile(os.path.join(d, "spheres.bin"), dtype=np.float32).reshape(-1, 12)
        nodes, idx = S.build_bvh(1, balls)
        ssc = S.HostScene(1, balls, nodes, idx, None, "balls")
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), num_bounces=4)
        want_b = O.oracle_frame_surface(ssc, fp, S.SurfaceData(2, None), S.default_material(0), S.point_light(eye))
        got_b = np.fromfile(os.path.join(d, "frame_spheres.bin"), dtype=np.float32).reshape(120, 160, 4)
        assert float((np.abs(got_b - want_b).max(axis=2) > 1 / 255).mean()) == 0.0 and want_b[..., 0].std() > 0.05
        # a spot_light list: the program's flattened light through the oracle
        from visionaray_b200._native import LIGHT_DTYPE
        spot = np.fromfile(os.path.join(d, "spot_light.bin"), dtype=LIGHT_DTYPE)
        assert spot["kind"][0] == 1
        fp = S.frame_params(1, 160, 120, eye, (0, 0, 0), num_bounces=4)
        want_sp = O.oracle_frame(sc, fp, S.default_material(0), spot)
        got_sp = np.fromfile(os.path.join(d, "frame_spot.bin"), dtype=np.float32).reshape(120, 160, 4)
        assert float((np.abs(got_sp - want_sp).max(axis=2) > 1 / 255).mean()) <= 1e-3 and want_sp[..., 0].std() > 0.02
