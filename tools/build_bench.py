"""Build time and tree quality of the GPU builder (vsb_scene_build) beside the reference's host
builder, per scene: build ms, node counts, and the traced Mrays/s of the same random rays on both
trees.  Usage: python tools/build_bench.py [scene scale]..."""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
api.tune("build_algorithm", 0)          # this tool sweeps the Morton + PLOC builder's knobs (tools/sah_build_bench.py compares both builders)
jobs = sys.argv[1:] or ["cfg2_grid", "1.0", "cfg5_soup", "0.25"]
for name, scale in zip(jobs[0::2], jobs[1::2]):
    t0 = time.time()
    sc = S.make_scene(name, float(scale))
    host_s = time.time() - t0                      # generator + reference-equivalent host build
    t0 = time.time()
    nodes, idx = S.build_bvh(sc.prim_kind, sc.prims)
    host_build_s = time.time() - t0
    ref = api.Scene(sc)
    d_prims = torch.from_numpy(sc.prims).cuda()
    out = {"scene": name, "prims": sc.n_prims, "host_builder_s": round(host_build_s, 3), "ref_tree": ref.stats()}
    lo, hi = sc.bounds()
    rays = torch.from_numpy(S.random_rays(1 << 22, 77, float(lo.min()), float(hi.max()))).cuda()
    hits = torch.empty((rays.shape[0], 8), dtype=torch.int32, device="cuda")

    def rate(dev):
        best = 0.0
        for _ in range(4):
            api.closest_hit(dev, rays, hits=hits)
            api.sync(); torch.cuda.synchronize()
            best = max(best, rays.shape[0] / api.last_kernel_ms() / 1e3)
        return round(best, 1)

    out["ref_tree_mrays"] = rate(ref)
    for knobs in ({}, {"build_radius": 16}, {"build_cost_prim_x100": 50}, {"build_cost_prim_x100": 30}, {"build_cost_prim_x100": 200}):
        api.tune("build_radius", 8); api.tune("build_cost_prim_x100", 100)
        for k, v in knobs.items():
            api.tune(k, v)
        for _ in range(2):
            dev = api.Scene.build(sc.prim_kind, d_prims)
            ms = api.last_build_ms()
            st = dev.stats()
            r = rate(dev)
            dev.close()
        out.setdefault("gpu", []).append({"knobs": knobs, "build": {k: round(v, 2) for k, v in ms.items()}, "tree": st, "mrays": r})
    ref.close()
    print(json.dumps(out), flush=True)
