"""End-to-end frame time into HOST render targets of each colour format (cfg2 view, Whitted): the device
trace is the same, the D2H copy shrinks with the pixel size (16 / 4 / 16 / 3 bytes).  Pinned host buffers,
wall clock around vsb_frame (which returns with the pixels in place).  Usage: python tools/format_bench.py"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
W = H = 4096
eye = (0.3, 1.2, 2.2)
sc = S.make_scene("cfg2_grid", 1.0)
dev = api.Scene(sc)
mats, light = S.default_material(0), S.point_light(eye)
fp = S.frame_params(1, W, H, eye, (0, 0, 0), num_bounces=4)
out = {"frame": [W, H], "kernel": "whitted", "triangles": sc.n_prims}
for name, fid, bpp in (("PF_RGBA32F", 0, 16), ("PF_RGBA8", 1, 4), ("PF_RGB32F", 2, 16), ("PF_RGB8", 3, 3)):
    buf = torch.zeros(W * H * bpp, dtype=torch.uint8).pin_memory().numpy()
    rays = np.zeros(1, dtype=np.uint64)
    best = 1e30
    for _ in range(6):
        t0 = time.perf_counter()
        api.frame(dev, fp, None, mats, light, buf, ray_count=rays, color_format=fid)
        best = min(best, time.perf_counter() - t0)
    out[name] = {"ms": round(best * 1e3, 3), "d2h_mb": round(W * H * bpp / 1e6, 1), "mrays_s": round(int(rays[0]) / best / 1e6, 1)}
dev.close()
print(json.dumps(out), flush=True)
