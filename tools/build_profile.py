"""Two GPU builds of one scene (the second is the profiled one): python tools/build_profile.py cfg2_grid 1.0 [algorithm]"""
import sys

import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
sc = S.make_scene(sys.argv[1], float(sys.argv[2]), build_tree=False)
if len(sys.argv) > 3:
    api.tune("build_algorithm", int(sys.argv[3]))          # 1 (default): binned SAH breadth-first, 0: Morton + PLOC
d_prims = torch.from_numpy(sc.prims).cuda()
for _ in range(2):
    dev = api.Scene.build(sc.prim_kind, d_prims)
    print(api.last_build_ms(), dev.stats(), flush=True)
    dev.close()
