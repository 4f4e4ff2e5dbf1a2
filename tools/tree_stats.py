"""CPU only: per-ray work of closest_hit on the BVH8 made from (a) the reference builder's tree and (b) the GPU
builder's tree (its serial instantiation, bit-identical to vsb_scene_build: tests/test_gpu_builder.py), for the primary
rays of the cfg2 view and for random rays: node visits, leaf groups entered, primitive tests; plus how full the
BVH8's nodes and leaf groups are.  Usage: python tools/tree_stats.py [scene scale [cost_prim]]"""
import json
import sys

import numpy as np

sys.path.insert(0, "."); sys.path.insert(0, "tests")
from hostsim_bindings import HostWide, build_ploc, decode_wide_nodes  # noqa: E402
from visionaray_b200 import scenes as S  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2_grid"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 0.25
cost_prim = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
sc = S.make_scene(name, scale)
fp = S.frame_params(0, 192, 192, (0.3, 1.2, 2.2), (0, 0, 0))
primary = S.primary_rays(fp)
lo, hi = sc.bounds()
rnd = S.random_rays(20_000, 77, float(lo.min()), float(hi.max()))
out = {"scene": name, "prims": sc.n_prims}


def fill(wide):
    d = decode_wide_nodes(wide.nodes())
    return d


for label, (nodes, idx) in (("reference_tree", (sc.nodes, sc.indices)), ("gpu_tree", build_ploc(sc.prim_kind, sc.prims, 8, 1.0, cost_prim)[:2])):
    w = HostWide(S.HostScene(sc.prim_kind, sc.prims, nodes, idx, None, label))
    leaves = int((nodes["num_prims"] > 0).sum())
    out[label] = {"bvh2_nodes": int(nodes.shape[0]), "prims_per_bvh2_leaf": round(sc.n_prims / leaves, 2), **w.stats(),
                  "primary": w.work_counters(primary), "random": w.work_counters(rnd)}
    w.close()
print(json.dumps(out, indent=1))
