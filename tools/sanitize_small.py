"""Small end-to-end run (seconds): GPU build with both builders, ray queries (wide / exact), one frame per kernel.
Meant to run under a memory checker where one is available (compute-sanitizer is closed on this pool)."""
import sys

import numpy as np

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
sc = S.make_scene("cfg1_sphere", 0.1)
rays = S.random_rays(20000, 3, -1.5, 1.5)
for algo in (1, 0):
    api.tune("build_algorithm", algo)
    dev = api.Scene.build(sc.prim_kind, sc.prims, sc.normals)
    a = api.closest_hit(dev, rays)
    b = api.closest_hit(dev, rays, traverse=api.TRAVERSE_EXACT)
    c = api.any_hit(dev, rays, max_t=0.5)
    assert (a["hit"] == b["hit"]).all()
    eye = (0.0, 0.0, 3.0)
    mats, lights = S.default_material(0), S.point_light(eye)
    for kernel in (0, 1, 2, 3):
        fp = S.frame_params(kernel, 96, 64, eye, (0, 0, 0), num_bounces=3, ao_samples=4, ao_radius=0.2)
        img = np.zeros((64, 96, 4), dtype=np.float32)
        api.frame(dev, fp, None, mats, lights, img)
    dev.close()
api.tune("build_algorithm", 1)
print("sanitize_small ok", int(a["hit"].sum()))
