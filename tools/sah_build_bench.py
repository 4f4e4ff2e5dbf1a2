"""Build time of the two GPU builders (vsb_tune build_algorithm 0 = Morton + PLOC, 1 = the reference's binned SAH
breadth-first) and the traced rate on their trees beside the host-built reference tree.
Usage: python tools/sah_build_bench.py [scene scale]..."""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
jobs = sys.argv[1:] or ["cfg2_grid", "1.0", "cfg5_soup", "0.25"]
for name, scale in zip(jobs[0::2], jobs[1::2]):
    sc = S.make_scene(name, float(scale))
    d_prims = torch.from_numpy(sc.prims).cuda()
    lo, hi = sc.bounds()
    rays = torch.from_numpy(S.random_rays(1 << 22, 77, float(lo.min()), float(hi.max()))).cuda()
    hits = torch.empty((rays.shape[0], 8), dtype=torch.int32, device="cuda")
    out = {"scene": name, "prims": sc.n_prims}

    def rate(dev):
        best = 0.0
        for _ in range(3):
            api.closest_hit(dev, rays, hits=hits)
            api.sync(); torch.cuda.synchronize()
            best = max(best, rays.shape[0] / api.last_kernel_ms() / 1e3)
        return round(best, 1)

    ref = api.Scene(sc)
    out["host_tree"] = {"bvh2_nodes": int(sc.nodes.shape[0]), "mrays": rate(ref)}
    ref.close()
    for algo in (0, 1):
        api.tune("build_algorithm", algo)
        for _ in range(3):
            dev = api.Scene.build(sc.prim_kind, d_prims)
            ms = api.last_build_ms()
            st = dev.stats()
            best = rate(dev)
            dev.close()
        out["ploc" if algo == 0 else "sah"] = {"build": {k: round(v, 2) for k, v in ms.items()}, "bvh2_nodes": st["bvh2_nodes"],
                                                 "wide_nodes": st["wide_nodes"], "mrays": best}
    api.tune("build_algorithm", 0)
    print(json.dumps(out))
