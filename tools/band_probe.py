import time, numpy as np, torch, sys
sys.path.insert(0, "/root/repo")
from visionaray_b200 import api, scenes as S
api.init(0)
sc = S.make_scene("cfg2_grid", 1.0); dev = api.Scene(sc)
eye = (0.3, 1.2, 2.2)
fp = S.frame_params(1, 4096, 4096, eye, (0, 0, 0), num_bounces=4)
mats, lights = S.default_material(0), S.point_light(eye)
host = torch.zeros((4096 * 4096, 4), dtype=torch.float32).pin_memory().numpy()
d_m = torch.from_numpy(mats.view(np.float32).copy()).cuda(); d_l = torch.from_numpy(lights.view(np.float32).copy()).cuda()
img = torch.zeros((4096, 4096, 4), dtype=torch.float32, device="cuda")
def t_host(b, n=5):
    api.tune("frame_bands", b)
    api.frame(dev, fp, None, mats, lights, host)
    t0 = time.perf_counter()
    for _ in range(n): api.frame(dev, fp, None, mats, lights, host)
    return (time.perf_counter() - t0) / n * 1e3
for b in (1, 2, 3, 4, 8): print("host bands", b, "%.2f ms" % t_host(b))
api.frame(dev, fp, None, d_m, d_l, img); api.sync()
t0 = time.perf_counter()
for _ in range(5): api.frame(dev, fp, None, d_m, d_l, img); api.sync()
print("device frame %.2f ms" % ((time.perf_counter() - t0) / 5 * 1e3))
