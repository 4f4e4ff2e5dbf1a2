"""Frame rate of the built-in kernels with and without a surface description on the
cfg2 scene and view: plain (plastic, per-face normals, lean instantiation), the same inputs through the SURF
extended instantiation (generic material list), and the full surface (normals + colours per vertex,
tex coords + RGBA8 textures, generic_material<emissive, matte, mirror, plastic>).  Device render target,
CUDA events of the library around the launch.  Also: multi_hit<N> and split-tree rates on random rays.
Usage: python tools/surface_bench.py [scale]"""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 1.0
W = H = 4096
eye = (0.3, 1.2, 2.2)
sc = S.make_scene("cfg2_grid", scale)
sc.prims.view(np.uint32)[:, 0] = (np.arange(sc.n_prims) // 3) % 4
dev = api.Scene(sc)
light = torch.from_numpy(S.point_light(eye).view(np.float32).copy()).cuda()
img = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
rays = torch.zeros(1, dtype=torch.int64, device="cuda")


def rate(fp, mats, generic):
    d_m = torch.from_numpy(mats.view(np.float32).copy()).cuda()
    best, n = 1e30, 0
    for _ in range(5):
        rays.zero_()
        api.frame(dev, fp, None, d_m, light, img, ray_count=rays, generic_materials=generic)
        api.sync(); torch.cuda.synchronize()
        best = min(best, api.last_kernel_ms())
        n = int(rays.item())
    return {"ms": round(best, 3), "rays": n, "mrays_s": round(n / best / 1e3, 1)}


out = {"scene": "cfg2_grid", "scale": scale, "triangles": sc.n_prims, "frame": [W, H]}
plastic = np.repeat(S.default_material(0), 4)
for kernel, name in ((1, "whitted"), (3, "pathtracing"), (0, "simple")):
    fp = S.frame_params(kernel, W, H, eye, (0, 0, 0), pixel_sampler=2 if kernel == 3 else 0, n_materials=4,
                        num_bounces=8 if kernel == 3 else 4, ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0))
    r = {}
    api.set_surface(dev, None)
    r["plain"] = rate(fp, plastic, False)
    r["plain_inputs_surf_instantiation"] = rate(fp, plastic, True)
    r["generic_materials"] = rate(fp, S.generic_materials(), True)
    # the face normal at all three vertices: the same rays as `plain`, so the difference is the surface fetch alone
    flat = S.SurfaceData(1, np.ascontiguousarray(np.repeat(sc.normals, 3, axis=0)))
    api.set_surface(dev, flat)
    r["normals_per_vertex_same_rays"] = rate(fp, plastic, False)
    api.set_surface(dev, S.make_surface(sc, 1, 0, False, n_geoms=4))
    r["normals_per_vertex"] = rate(fp, plastic, False)
    api.set_surface(dev, S.make_surface(sc, 1, 2, True, n_geoms=4))
    r["vertex_normals_colors_textures_generic"] = rate(fp, S.generic_materials(), True)
    out[name] = r
api.set_surface(dev, None)

# multi_hit<N> on random rays (one thread per ray, records in the output array)
lo, hi = sc.bounds()
rr = torch.from_numpy(S.random_rays(1 << 22, 77, float(lo.min()), float(hi.max()))).cuda()
mh = {}
for n in (1, 4, 16):
    hits = torch.empty((rr.shape[0], n, 8), dtype=torch.int32, device="cuda")
    for trav, key in ((api.TRAVERSE_WIDE, "wide"), (api.TRAVERSE_EXACT, "exact")):
        best = 1e30
        for _ in range(3):
            api.multi_hit(dev, rr, n, hits=hits, traverse=trav)
            api.sync(); torch.cuda.synchronize()
            best = min(best, api.last_kernel_ms())
        mh[f"n{n}_{key}_mrays_s"] = round(rr.shape[0] / best / 1e3, 1)
h1 = torch.empty((rr.shape[0], 8), dtype=torch.int32, device="cuda")
best = 1e30
for _ in range(3):
    api.closest_hit(dev, rr, hits=h1)
    api.sync(); torch.cuda.synchronize()
    best = min(best, api.last_kernel_ms())
mh["closest_hit_wide_mrays_s"] = round(rr.shape[0] / best / 1e3, 1)
out["multi_hit_random_rays"] = mh
dev.close()
print(json.dumps(out), flush=True)
