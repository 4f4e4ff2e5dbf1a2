"""Traced-rate gap of the GPU builder (no spatial splits) against the host builder WITH spatial splits
(detail/bvh/sah.h:428-615) on the case spatial splits exist for: a soup of long thin triangles (the generator of
tests/golden/make_golden.py:sbvh_scene_prims at N triangles, stretched 30x along x).  Same rays on three trees:
host SAH without splits, host SAH with splits, and the two GPU builders (binned SAH = the host tree; Morton + PLOC).  Usage: python tools/sbvh_gap.py [n_triangles]"""
import json
import sys
import time

import numpy as np
import torch

sys.path.insert(0, ".")
from visionaray_b200 import api, scenes as S  # noqa: E402

api.init(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
tris = S.gen_triangles(5, n, 0, 31, 0.05 * (3000.0 / n) ** (1.0 / 3.0) * 3.0).copy()     # keep the overlap per triangle roughly constant
for o in (4, 8, 12):
    tris[:, o] *= 30.0
rays = S.random_rays(1 << 21, 5, -0.2, 1.2)
rays[:, 0] *= 30.0
d_rays = torch.from_numpy(rays).cuda()
hits = torch.empty((rays.shape[0], 8), dtype=torch.int32, device="cuda")


def rate(dev):
    best = 0.0
    for _ in range(4):
        api.closest_hit(dev, d_rays, hits=hits)
        api.sync(); torch.cuda.synchronize()
        best = max(best, rays.shape[0] / api.last_kernel_ms() / 1e3)
    return round(best, 1), hits.cpu().numpy().copy()


out = {"triangles": n, "rays": rays.shape[0]}
ref_hits = None
for label, split in (("host_sah", False), ("host_sbvh", True)):
    t0 = time.time()
    nodes, idx = S.build_bvh(0, tris, spatial_splits=split)
    dt = time.time() - t0
    dev = api.Scene(S.HostScene(0, tris, nodes, idx, None, label))
    r, h = rate(dev)
    out[label] = {"build_s": round(dt, 3), "nodes": int(nodes.shape[0]), "indices": int(idx.shape[0]), "mrays": r, "tree": dev.stats()}
    if ref_hits is None:
        ref_hits = h
    else:
        out[label]["t_equal_to_host_sah"] = bool(np.array_equal(h[:, 3], ref_hits[:, 3]))
    dev.close()
for label, algo in (("gpu_sah", 1), ("gpu_ploc", 0)):
    api.tune("build_algorithm", algo)
    for _ in range(2):
        dev = api.Scene.build(0, torch.from_numpy(tris).cuda())
        ms = api.last_build_ms()
        r, h = rate(dev)
        st = dev.stats()
        dev.close()
    out[label] = {"build_ms": round(ms["total_ms"], 2), "iterations": ms["iterations"], "mrays": r, "tree": st,
                  "t_equal_to_host_sah": bool(np.array_equal(h[:, 3], ref_hits[:, 3]))}
api.tune("build_algorithm", 1)
out["gpu_vs_sbvh"] = round(out["gpu_ploc"]["mrays"] / out["host_sbvh"]["mrays"], 3)
out["gpu_vs_sah"] = round(out["gpu_ploc"]["mrays"] / out["host_sah"]["mrays"], 3)
out["gpu_sah_vs_sbvh"] = round(out["gpu_sah"]["mrays"] / out["host_sbvh"]["mrays"], 3)
print(json.dumps(out))
