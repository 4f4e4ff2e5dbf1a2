"""Timeline of a banded host-target frame (cfg2): when each band is traced and when it is on the host."""
import sys
import numpy as np, torch
sys.path.insert(0, "/root/repo")
from visionaray_b200 import api, scenes as S
api.init(0)
sc = S.make_scene("cfg2_grid", 1.0); dev = api.Scene(sc)
eye = (0.3, 1.2, 2.2)
fp = S.frame_params(1, 4096, 4096, eye, (0, 0, 0), num_bounces=4)
mats, lights = S.default_material(0), S.point_light(eye)
host = torch.zeros((4096 * 4096, 4), dtype=torch.float32).pin_memory().numpy()
bands = int(sys.argv[1]) if len(sys.argv) > 1 else 16
api.tune("frame_bands", bands)
for _ in range(3):
    api.frame(dev, fp, None, mats, lights, host)
api.tune("band_trace", 1)
api.frame(dev, fp, None, mats, lights, host)
