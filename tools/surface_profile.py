"""Two frames of the full-surface Whitted workload (cfg2 scene and view; normals + colours per vertex,
textures, generic materials) and two of the path tracer -- the process ncu attaches to for the profile
of the EXT frame programs.  Usage: python tools/surface_profile.py [kernel]"""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
kernel = int(sys.argv[1]) if len(sys.argv) > 1 else 1
W = H = 4096
eye = (0.3, 1.2, 2.2)
sc = S.make_scene("cfg2_grid", 1.0)
sc.prims.view(np.uint32)[:, 0] = (np.arange(sc.n_prims) // 3) % 4
dev = api.Scene(sc)
api.set_surface(dev, S.make_surface(sc, 1, 2, True, n_geoms=4))
mats = torch.from_numpy(S.generic_materials().view(np.float32).copy()).cuda()
light = torch.from_numpy(S.point_light(eye).view(np.float32).copy()).cuda()
img = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
fp = S.frame_params(kernel, W, H, eye, (0, 0, 0), pixel_sampler=2 if kernel == 3 else 0, n_materials=4,
                    num_bounces=8 if kernel == 3 else 4, ambient=(1, 1, 1, 1) if kernel == 3 else (0, 0, 0, 0))
for _ in range(2):
    api.frame(dev, fp, None, mats, light, img, generic_materials=True)
    api.sync(); torch.cuda.synchronize()
print("ms", api.last_kernel_ms())
dev.close()
