"""TEST-ONLY ctypes bindings of tests/hostsim/libvsb_hostsim.so: the product's host/device
headers (csrc/*.cuh) instantiated for the CPU, so the BVH2->BVH8 collapse and both traversal
loops can be checked in the no-GPU tier.  Not part of the product."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
HIT_DTYPE = np.dtype(
    [("hit", "<i4"), ("prim_id", "<u4"), ("geom_id", "<u4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4"),
     ("prim_list_index", "<u4"), ("pad", "<u4")]
)
_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(os.path.join(_HERE, "hostsim", "libvsb_hostsim.so"))
        vp, sz = C.c_void_p, C.c_size_t
        l.hs_build.restype = vp
        l.hs_build.argtypes = [C.c_int, vp, sz, vp, sz, vp, sz]
        l.hs_free.argtypes = [vp]
        l.hs_stats.argtypes = [vp, C.POINTER(C.c_uint64)]
        l.hs_nodes.restype = vp
        l.hs_nodes.argtypes = [vp]
        l.hs_prims.restype = vp
        l.hs_prims.argtypes = [vp]
        l.hs_query_wide.argtypes = [vp, C.c_int, vp, sz, C.c_float, vp, vp]
        l.hs_query_watertight.argtypes = [vp, C.c_int, vp, sz, C.c_float, vp]
        l.hs_count_wide.argtypes = [vp, vp, sz, C.POINTER(C.c_uint64)]
        l.hs_query_exact.argtypes = [C.c_int, C.c_int, vp, vp, vp, vp, sz, C.c_float, vp, vp]
        l.hs_multi_hit_wide.argtypes = [vp, vp, sz, C.c_int, vp]
        l.hs_multi_hit_exact.argtypes = [C.c_int, vp, vp, vp, vp, sz, C.c_int, vp]
        l.hs_build_ploc.restype = sz
        l.hs_build_ploc.argtypes = [C.c_int, vp, sz, C.c_int, C.c_float, C.c_float, vp, sz, vp, vp]
        _lib = l
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


class HostWide:
    def __init__(self, scene):
        self.scene = scene
        self.h = lib().hs_build(scene.prim_kind, _p(scene.prims), scene.prims.shape[0], _p(scene.nodes),
                                scene.nodes.shape[0], _p(scene.indices),
                                scene.indices.shape[0] if scene.indices is not None else 0)

    def stats(self):
        out = (C.c_uint64 * 3)()
        lib().hs_stats(self.h, out)
        return dict(wide_nodes=int(out[0]), wide_prims=int(out[1]), levels=int(out[2]))

    def nodes(self) -> np.ndarray:
        n = self.stats()["wide_nodes"]
        buf = (C.c_uint32 * (20 * n)).from_address(lib().hs_nodes(self.h))
        return np.frombuffer(buf, dtype=np.uint32).reshape(n, 20).copy()

    def prims(self) -> np.ndarray:
        n = self.stats()["wide_prims"]
        words = 12 if self.scene.prim_kind == 0 else 8
        buf = (C.c_uint32 * (words * n)).from_address(lib().hs_prims(self.h))
        return np.frombuffer(buf, dtype=np.uint32).reshape(n, words).copy()

    def query(self, rays, any_hit=False, max_t=3.4028234663852886e38, max_t_per_ray=None):
        hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        lib().hs_query_wide(self.h, int(any_hit), _p(rays), rays.shape[0], max_t, _p(max_t_per_ray), _p(hits))
        return hits

    def work_counters(self, rays):
        """Per-ray averages of closest_hit's work on this BVH8: node visits, leaf groups entered, primitive tests."""
        out = (C.c_uint64 * 4)()
        lib().hs_count_wide(self.h, _p(rays), rays.shape[0], out)
        n = float(rays.shape[0])
        return dict(node_visits=out[0] / n, leaf_groups=out[1] / n, prim_tests=out[2] / n, deepest_stack=int(out[3]))

    def multi_hit(self, rays, max_hits):
        hits = np.zeros((rays.shape[0], max_hits), dtype=HIT_DTYPE)
        lib().hs_multi_hit_wide(self.h, _p(rays), rays.shape[0], max_hits, _p(hits))
        return hits

    def close(self):
        if self.h:
            lib().hs_free(self.h)
            self.h = None

    def __del__(self):
        self.close()


def query_watertight(wide: "HostWide", rays, any_hit=False, max_t=3.4028234663852886e38):
    hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
    lib().hs_query_watertight(wide.h, int(any_hit), _p(rays), rays.shape[0], max_t, _p(hits))
    return hits


def query_exact(scene, rays, any_hit=False, max_t=3.4028234663852886e38, max_t_per_ray=None):
    hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
    lib().hs_query_exact(scene.prim_kind, int(any_hit), _p(scene.prims), _p(scene.nodes), _p(scene.indices), _p(rays),
                         rays.shape[0], max_t, _p(max_t_per_ray), _p(hits))
    return hits


def multi_hit_exact(scene, rays, max_hits):
    hits = np.zeros((rays.shape[0], max_hits), dtype=HIT_DTYPE)
    lib().hs_multi_hit_exact(scene.prim_kind, _p(scene.prims), _p(scene.nodes), _p(scene.indices), _p(rays), rays.shape[0],
                             max_hits, _p(hits))
    return hits


NODE_DTYPE = np.dtype(
    [("bbox_min", "<f4", (3,)), ("first", "<u4"), ("bbox_max", "<f4", (3,)), ("num_prims", "<u4")]
)


def build_ploc(prim_kind, prims, radius=8, cost_node=1.0, cost_prim=1.0):
    """Serial instantiation of the GPU builder (csrc/build_ploc.cuh): (nodes, indices, info)."""
    n = prims.shape[0]
    nodes = np.zeros(max(2 * n - 1, 1), dtype=NODE_DTYPE)
    idx = np.zeros(n, dtype=np.uint32)
    info = np.zeros(4, dtype=np.uint32)
    nn = lib().hs_build_ploc(prim_kind, _p(prims), n, radius, cost_node, cost_prim, _p(nodes), nodes.shape[0], _p(idx), _p(info))
    assert nn > 0
    return nodes[:nn].copy(), idx, dict(depth=int(info[0]), iterations=int(info[1]), leaves=int(info[2]))


def decode_wide_nodes(nodes: np.ndarray):
    """Dequantised child boxes of Node80 records (canonical form p + q * 2^(e-127)).
    Returns dict(lo[n,8,3], hi[n,8,3], meta[n,8], imask[n], child_base[n], prim_base[n])."""
    n = nodes.shape[0]
    p = nodes[:, 0:3].copy().view(np.float32).astype(np.float64)
    e = nodes[:, 3]
    ex = np.stack([(e >> (8 * k)) & 0xFF for k in range(3)], axis=1).astype(np.int64)
    scale = np.where(ex > 0, np.ldexp(1.0, ex - 127), 0.0)
    imask = (e >> 24) & 0xFF
    b = nodes[:, 6:20].copy().view(np.uint8).reshape(n, 14, 4)  # w[6..19] as bytes

    def eight(w0):  # two consecutive words -> 8 bytes
        return np.concatenate([b[:, w0 - 6], b[:, w0 - 5]], axis=1)

    # w[6] = `valid`: nibble s = (1 << count) - 1 of the leaf in slot s; records in slot order from prim_base.
    # Reported as meta = count << 5 | offset (0 for inner children / empty slots).
    valid = nodes[:, 6].astype(np.uint32)
    nib = np.stack([(valid >> (4 * k)) & 0xF for k in range(8)], axis=1)
    assert np.isin(nib, (0, 1, 3, 7, 15)).all() and (nodes[:, 7] == 0).all()
    count = np.array([0, 1, 0, 2, 0, 0, 0, 3, 0, 0, 0, 0, 0, 0, 0, 4], dtype=np.uint32)[nib]
    ofs = np.cumsum(count, axis=1) - count
    meta = np.where(count > 0, (count << 5) | ofs, 0).astype(np.uint32)
    qlo = np.stack([eight(8), eight(10), eight(12)], axis=2).astype(np.float64)
    qhi = np.stack([eight(14), eight(16), eight(18)], axis=2).astype(np.float64)
    lo = p[:, None, :] + qlo * scale[:, None, :]
    hi = p[:, None, :] + qhi * scale[:, None, :]
    return dict(lo=lo, hi=hi, meta=meta, imask=imask, child_base=nodes[:, 4], prim_base=nodes[:, 5])
