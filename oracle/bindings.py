"""TEST INFRASTRUCTURE ONLY -- ctypes bindings of the CPU oracle.

* ``oracle/libvsnray_oracle.so``         plain-C restatement (vsnray_oracle.c), built anywhere.
* ``oracle/_ref/libvsnray_ref.so``       the UNMODIFIED reference headers behind ref_harness.cpp,
  parity build (-ffp-contract=off); prebuilt in the build container, travels to the GPU box.
* ``oracle/_ref/libvsnray_ref_fast.so``  same, -O3 -mavx2 -mfma: the timed CPU baseline.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may import
this module.  The product package (visionaray_b200/) never does.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

HIT_DTYPE = np.dtype(
    [("hit", "<i4"), ("prim_id", "<u4"), ("geom_id", "<u4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4"),
     ("prim_list_index", "<u4"), ("pad", "<u4")]
)


class Bvh(C.Structure):
    _fields_ = [("prim_kind", C.c_int), ("prims", C.c_void_p), ("n_prims", C.c_size_t),
                ("nodes", C.c_void_p), ("n_nodes", C.c_size_t), ("indices", C.c_void_p), ("n_indices", C.c_size_t)]


class Counters(C.Structure):
    _fields_ = [("n_inner", C.c_uint64), ("n_prim", C.c_uint64), ("n_rays", C.c_uint64), ("n_hits", C.c_uint64)]


class Texture(C.Structure):
    _fields_ = [("data", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("filter", C.c_int32),
                ("address", C.c_int32 * 2)]


class Surface(C.Structure):
    _fields_ = [("normal_binding", C.c_int32), ("color_binding", C.c_int32), ("normals", C.c_void_p),
                ("colors", C.c_void_p), ("tex_coords", C.c_void_p), ("textures", C.c_void_p), ("n_textures", C.c_int32)]


def make_texture(t) -> Texture:
    o = Texture()
    if t is not None:
        o.data = t.data.ctypes.data; o.height, o.width = t.data.shape[0], t.data.shape[1]
        o.filter = t.filter; o.address[0], o.address[1] = t.address
    return o


def make_surface(sd):
    """visionaray_b200.scenes.SurfaceData -> (Surface, keepalive)."""
    s = Surface()
    s.normal_binding, s.color_binding = sd.normal_binding, sd.color_binding
    s.normals = sd.normals.ctypes.data if sd.normals is not None else None
    s.colors = sd.colors.ctypes.data if sd.colors is not None else None
    s.tex_coords = sd.tex_coords.ctypes.data if sd.tex_coords is not None else None
    texs = None
    if sd.textures:
        texs = (Texture * len(sd.textures))(*[make_texture(t) for t in sd.textures])
        s.textures = C.cast(texs, C.c_void_p); s.n_textures = len(sd.textures)
    return s, texs


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


_oracle = None
_ref = {}


def oracle_lib():
    global _oracle
    if _oracle is None:
        path = os.path.join(_HERE, "libvsnray_oracle.so")
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle oracle` (or __graft_entry__.build())")
        lib = C.CDLL(path)
        vp, sz = C.c_void_p, C.c_size_t
        lib.vso_query.argtypes = [C.c_int, C.POINTER(Bvh), vp, sz, C.c_float, vp, vp, C.POINTER(Counters), C.c_int]
        lib.vso_query_brute.argtypes = [C.c_int, C.c_int, vp, sz, vp, sz, C.c_float, vp, C.c_int]
        lib.vso_frame.argtypes = [vp, C.POINTER(Bvh), vp, vp, vp, vp, C.POINTER(Counters), C.c_int]
        lib.vso_camera_basis.argtypes = [vp, vp, vp, vp, vp]
        lib.vso_frame_matrix.argtypes = [vp, vp, vp, C.POINTER(Bvh), vp, vp, vp, vp, vp, C.POINTER(Counters), C.c_int]
        lib.vso_mat4_inverse.argtypes = [vp, vp]
        lib.vso_frame_format.argtypes = [vp, C.POINTER(Bvh), vp, vp, vp, vp, C.c_int, C.POINTER(Counters), C.c_int]
        lib.vso_frame_surface.argtypes = [vp, C.POINTER(Bvh), vp, vp, vp, vp, C.POINTER(Counters), C.c_int]
        lib.vso_tex2d.argtypes = [vp, C.c_float, C.c_float, vp]
        lib.vso_multi_hit.argtypes = [C.POINTER(Bvh), C.c_int, vp, sz, C.c_int, vp, C.c_int]
        lib.vso_bvh_max_depth.argtypes = [C.POINTER(Bvh)]
        lib.vso_sampler_init.restype = C.c_uint32
        lib.vso_sampler_init.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32]
        lib.vso_sampler_next.restype = C.c_float
        lib.vso_sampler_next.argtypes = [C.POINTER(C.c_uint32)]
        _oracle = lib
    return _oracle


def ref_available(fast: bool = False) -> bool:
    return os.path.exists(os.path.join(_HERE, "_ref", "libvsnray_ref_fast.so" if fast else "libvsnray_ref.so"))


def ref_lib(fast: bool = False):
    key = bool(fast)
    if key not in _ref:
        path = os.path.join(_HERE, "_ref", "libvsnray_ref_fast.so" if fast else "libvsnray_ref.so")
        lib = C.CDLL(path)
        vp, sz = C.c_void_p, C.c_size_t
        lib.ref_layout.argtypes = [C.POINTER(C.c_int)]
        for f in (lib.ref_build_triangles, lib.ref_build_spheres) + ((lib.ref_build_generic,) if hasattr(lib, "ref_build_generic") else ()):
            f.argtypes = [vp, sz, C.c_int, vp, sz, C.POINTER(sz), vp, sz, C.POINTER(sz)]
        lib.ref_query.argtypes = [C.c_int, C.c_int, vp, sz, vp, sz, vp, sz, vp, sz, C.c_float, vp, vp, C.c_int]
        lib.ref_frame.restype = C.c_double
        lib.ref_frame.argtypes = [vp, vp, sz, vp, sz, vp, sz, vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, C.c_int, vp]
        lib.ref_frame_matrix.argtypes = [vp, vp, vp, vp, sz, vp, sz, vp, sz, vp, vp, vp, vp, vp, C.c_int, vp]
        lib.ref_camera_matrices.argtypes = [vp, C.c_float, C.c_float, vp, vp]
        if hasattr(lib, "ref_frame_format"):
            lib.ref_frame_format.argtypes = [vp, vp, sz, vp, sz, vp, sz, vp, vp, vp, vp, C.c_int, C.c_int]
        if hasattr(lib, "ref_frame_surface"):
            lib.ref_frame_surface.argtypes = [vp, vp, sz, vp, sz, vp, sz, C.c_int, C.c_int, vp, vp, vp, vp, C.c_int, vp, vp,
                                              vp, C.c_int]
            lib.ref_tex2d.argtypes = [vp, C.c_float, C.c_float, vp]
        if hasattr(lib, "ref_multi_hit"):
            lib.ref_multi_hit.argtypes = [C.c_int, C.c_int, vp, sz, vp, sz, vp, sz, vp, sz, vp, sz, vp, sz, vp, sz, vp, C.c_int]
        lib.ref_query_timed.restype = C.c_double
        lib.ref_query_timed.argtypes = [C.c_int, vp, sz, vp, sz, vp, sz, vp, sz, C.c_float, C.c_int, C.c_int, C.c_int,
                                        C.POINTER(C.c_uint64)]
        _ref[key] = lib
    return _ref[key]


def make_bvh(scene) -> Bvh:
    """scene: visionaray_b200.scenes.HostScene (kept alive by the caller)."""
    b = Bvh()
    b.prim_kind = scene.prim_kind
    b.prims = scene.prims.ctypes.data
    b.n_prims = scene.prims.shape[0]
    b.nodes = scene.nodes.ctypes.data
    b.n_nodes = scene.nodes.shape[0]
    b.indices = scene.indices.ctypes.data if scene.indices is not None else None
    b.n_indices = scene.indices.shape[0] if scene.indices is not None else 0
    return b


def oracle_query(scene, rays, any_hit=False, max_t=3.4028234663852886e38, max_t_per_ray=None, nthreads=0,
                 counters=None):
    hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
    b = make_bvh(scene)
    c = counters if counters is not None else Counters()
    oracle_lib().vso_query(int(any_hit), C.byref(b), _p(rays), rays.shape[0], max_t, _p(max_t_per_ray), _p(hits),
                           C.byref(c), nthreads or (os.cpu_count() or 1))
    return hits


def oracle_multi_hit(scenes, rays, n, nthreads=0):
    """multi_hit<n>() over a range of bvhs (one HostScene or a list): (n_rays, n) records sorted by t."""
    if not isinstance(scenes, (list, tuple)):
        scenes = [scenes]
    arr = (Bvh * len(scenes))(*[make_bvh(s) for s in scenes])
    hits = np.zeros((rays.shape[0], n), dtype=HIT_DTYPE)
    oracle_lib().vso_multi_hit(arr, len(scenes), _p(rays), rays.shape[0], n, _p(hits), nthreads or (os.cpu_count() or 1))
    return hits


def ref_multi_hit(scenes, rays, n, nthreads=0):
    """The same through the unmodified reference headers (oracle/_ref); n in {1, 2, 4, 5, 8, 16}, <= 2 bvhs."""
    if not isinstance(scenes, (list, tuple)):
        scenes = [scenes]
    hits = np.zeros((rays.shape[0], n), dtype=HIT_DTYPE)
    a = scenes[0]
    b = scenes[1] if len(scenes) > 1 else None
    args = [a.prim_kind, n, _p(a.prims), a.prims.shape[0], _p(a.nodes), a.nodes.shape[0], _p(a.indices), a.indices.shape[0]]
    if b is not None:
        args += [_p(b.prims), b.prims.shape[0], _p(b.nodes), b.nodes.shape[0], _p(b.indices), b.indices.shape[0]]
    else:
        args += [None, 0, None, 0, None, 0]
    rc = ref_lib().ref_multi_hit(*args, _p(rays), rays.shape[0], _p(hits), nthreads or (os.cpu_count() or 1))
    if rc != 0:
        raise ValueError(f"ref_multi_hit: unsupported n = {n}")
    return hits


def oracle_brute(scene, rays, any_hit=False, max_t=3.4028234663852886e38, nthreads=0):
    hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
    oracle_lib().vso_query_brute(int(any_hit), scene.prim_kind, _p(scene.prims), scene.prims.shape[0], _p(rays),
                                 rays.shape[0], max_t, _p(hits), nthreads or (os.cpu_count() or 1))
    return hits


def oracle_frame(scene, fp, materials, lights, rgba=None, nthreads=0, counters=None):
    w, h = fp.width, fp.height
    if rgba is None:
        rgba = np.zeros((h, w, 4), dtype=np.float32)
    b = make_bvh(scene)
    c = counters if counters is not None else Counters()
    oracle_lib().vso_frame(C.byref(fp), C.byref(b), _p(scene.normals), _p(materials), _p(lights), _p(rgba),
                           C.byref(c), nthreads or (os.cpu_count() or 1))
    return rgba


# colour formats of the render target: name -> (id, bytes per pixel, numpy element type, channels stored)
PIXEL_FORMATS = {"PF_RGBA32F": (0, 16, np.float32, 4), "PF_RGBA8": (1, 4, np.uint8, 4),
                 "PF_RGB32F": (2, 16, np.float32, 4), "PF_RGB8": (3, 3, np.uint8, 3)}


def format_buffer(fmt: str, width: int, height: int) -> np.ndarray:
    """A zeroed colour buffer of that format: (h, w, channels stored); PF_RGB32F pixels are 16 bytes
    (vector<3,float> is 16-byte aligned in the reference), the 4th word is never written."""
    _, _, dt, ch = PIXEL_FORMATS[fmt]
    return np.zeros((height, width, ch), dtype=dt)


def oracle_frame_format(scene, fp, materials, lights, fmt: str, color=None, nthreads=0):
    """vso_frame_format: the frame into a colour buffer of format `fmt` (store / blend / SSAA through it)."""
    if color is None:
        color = format_buffer(fmt, fp.width, fp.height)
    b = make_bvh(scene)
    c = Counters()
    oracle_lib().vso_frame_format(C.byref(fp), C.byref(b), _p(scene.normals), _p(materials), _p(lights), _p(color),
                                  PIXEL_FORMATS[fmt][0], C.byref(c), nthreads or (os.cpu_count() or 1))
    return color


def ref_frame_format(scene, fp, materials, lights, fmt: str, color=None, nthreads=0):
    """The same through the reference's own pixel_access::store / blend and colour conversion."""
    if color is None:
        color = format_buffer(fmt, fp.width, fp.height)
    rc = ref_lib().ref_frame_format(C.byref(fp), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes), scene.nodes.shape[0],
                                    _p(scene.indices), scene.indices.shape[0], _p(scene.normals), _p(materials), _p(lights),
                                    _p(color), PIXEL_FORMATS[fmt][0], nthreads or (os.cpu_count() or 1))
    if rc != 0:
        raise ValueError("ref_frame_format: unsupported kernel / format")
    return color


def oracle_frame_surface(scene, fp, surface, materials, lights, rgba=None, nthreads=0):
    """vso_frame_surface: kernel_params with a surface description + generic materials."""
    if rgba is None:
        rgba = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
    b = make_bvh(scene)
    s, keep = make_surface(surface)
    c = Counters()
    oracle_lib().vso_frame_surface(C.byref(fp), C.byref(b), C.byref(s), _p(materials), _p(lights), _p(rgba), C.byref(c),
                                   nthreads or (os.cpu_count() or 1))
    del keep
    return rgba


def ref_frame_surface(scene, fp, surface, materials, lights, rgba=None, nthreads=0):
    """The same through the unmodified reference headers (make_kernel_params / get_surface / tex2D)."""
    if rgba is None:
        rgba = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
    s, keep = make_surface(surface)
    rc = ref_lib().ref_frame_surface(C.byref(fp), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes), scene.nodes.shape[0],
                                     _p(scene.indices), scene.indices.shape[0], s.normal_binding, s.color_binding,
                                     s.normals, s.colors, s.tex_coords, s.textures, s.n_textures, _p(materials), _p(lights),
                                     _p(rgba), nthreads or (os.cpu_count() or 1))
    del keep
    if rc != 0:
        raise ValueError("ref_frame_surface: unsupported kernel")
    return rgba


def ref_frame_prim_normals(scene, fp, materials, lights, rgba=None, nthreads=0):
    """The reference kernels with kernel_params that carry NO normals (default make_kernel_params overload):
    the normal comes from the primitive.  Triangle and sphere scenes; simple / whitted (+ pathtracing for triangles)."""
    if rgba is None:
        rgba = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
    lib = ref_lib()
    lib.ref_frame_prim_normals.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_void_p,
                                           C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    rc = lib.ref_frame_prim_normals(scene.prim_kind, C.byref(fp), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes),
                                    scene.nodes.shape[0], _p(scene.indices), scene.indices.shape[0], _p(materials), _p(lights),
                                    _p(rgba), nthreads or (os.cpu_count() or 1))
    if rc != 0:
        raise ValueError("ref_frame_prim_normals: unsupported kernel / primitive kind")
    return rgba


def ref_frame_spot(scene, fp, materials, lights, cutoff_angles, rgba=None, nthreads=0):
    """simple / whitted through the reference headers with a spot_light<float> list."""
    if rgba is None:
        rgba = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
    lib = ref_lib()
    vp, sz = C.c_void_p, C.c_size_t
    lib.ref_frame_spot.argtypes = [vp, vp, sz, vp, sz, vp, sz, vp, vp, vp, vp, vp, C.c_int]
    ang = np.ascontiguousarray(cutoff_angles, dtype=np.float32)
    rc = lib.ref_frame_spot(C.byref(fp), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes), scene.nodes.shape[0],
                            _p(scene.indices), scene.indices.shape[0], _p(scene.normals), _p(materials), _p(lights), _p(ang),
                            _p(rgba), nthreads or (os.cpu_count() or 1))
    if rc != 0:
        raise ValueError("ref_frame_spot: unsupported kernel")
    return rgba


def ref_frame_generic(scene, fp, surface, materials, lights, rgba=None, nthreads=0):
    """simple / whitted over an index_bvh<generic_primitive<triangle, sphere>> with a normals array."""
    if rgba is None:
        rgba = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
    lib = ref_lib()
    vp, sz = C.c_void_p, C.c_size_t
    lib.ref_frame_generic.argtypes = [vp, vp, sz, vp, sz, vp, sz, C.c_int, vp, vp, vp, vp, C.c_int]
    rc = lib.ref_frame_generic(C.byref(fp), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes), scene.nodes.shape[0],
                               _p(scene.indices), scene.indices.shape[0], surface.normal_binding, _p(surface.normals),
                               _p(materials), _p(lights), _p(rgba), nthreads or (os.cpu_count() or 1))
    if rc != 0:
        raise ValueError("ref_frame_generic: unsupported kernel")
    return rgba


def oracle_tex2d(tex, s, t):
    o = make_texture(tex)
    out = np.zeros(4, dtype=np.float32)
    oracle_lib().vso_tex2d(C.byref(o), s, t, _p(out))
    return out


def ref_tex2d(tex, s, t):
    o = make_texture(tex)
    out = np.zeros(4, dtype=np.float32)
    ref_lib().ref_tex2d(C.byref(o), s, t, _p(out))
    return out


def oracle_frame_matrix(scene, fp, view, proj, materials, lights, rgba=None, depth=None, want_depth=True, nthreads=0):
    """vso_frame_matrix: sched_params with view / proj matrices (16 floats, column-major); returns (rgba, depth)."""
    w, h = fp.width, fp.height
    if rgba is None:
        rgba = np.zeros((h, w, 4), dtype=np.float32)
    if depth is None and want_depth:
        depth = np.zeros((h, w), dtype=np.float32)
    view = np.ascontiguousarray(view, dtype=np.float32).reshape(16)
    proj = np.ascontiguousarray(proj, dtype=np.float32).reshape(16)
    b = make_bvh(scene)
    c = Counters()
    oracle_lib().vso_frame_matrix(C.byref(fp), _p(view), _p(proj), C.byref(b), _p(scene.normals), _p(materials),
                                  _p(lights), _p(rgba), _p(depth), C.byref(c), nthreads or (os.cpu_count() or 1))
    return rgba, depth


def oracle_mat4_inverse(m):
    m = np.ascontiguousarray(m, dtype=np.float32).reshape(16)
    out = np.zeros(16, dtype=np.float32)
    oracle_lib().vso_mat4_inverse(_p(m), _p(out))
    return out


def oracle_max_depth(scene) -> int:
    b = make_bvh(scene)
    return int(oracle_lib().vso_bvh_max_depth(C.byref(b)))


def ref_build(prim_kind, prims, spatial_splits=False):
    lib = ref_lib()
    n = prims.shape[0]
    nodes = np.zeros(max(2 * n - 1, 1) * (4 if spatial_splits else 1), dtype=np.dtype(
        [("bbox_min", "<f4", (3,)), ("first", "<u4"), ("bbox_max", "<f4", (3,)), ("num_prims", "<u4")]))
    idx = np.zeros(max(n, 1) * (4 if spatial_splits else 1), dtype=np.uint32)
    nn, ni = C.c_size_t(0), C.c_size_t(0)
    f = lib.ref_build_triangles if prim_kind == 0 else (lib.ref_build_spheres if prim_kind == 1 else lib.ref_build_generic)
    rc = f(_p(prims), n, int(spatial_splits), _p(nodes), nodes.shape[0], C.byref(nn), _p(idx), idx.shape[0], C.byref(ni))
    if rc != 0:
        raise RuntimeError("ref_build: capacity")
    return nodes[: nn.value].copy(), idx[: ni.value].copy()


def ref_make_generic(tags, tris, sphs):
    """generic_primitive records built by the reference's own constructors (pins the 80-byte layout)."""
    lib = ref_lib()
    lib.ref_make_generic.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    n = tags.shape[0]
    size = lib.ref_make_generic(_p(tags), _p(tris), _p(sphs), n, None)
    out = np.zeros((n, size // 4), dtype=np.float32)
    lib.ref_make_generic(_p(tags), _p(tris), _p(sphs), n, _p(out))
    return out


def ref_query(scene, rays, any_hit=False, max_t=3.4028234663852886e38, max_t_per_ray=None, nthreads=0):
    hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
    ref_lib().ref_query(scene.prim_kind, int(any_hit), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes),
                        scene.nodes.shape[0], _p(scene.indices), scene.indices.shape[0], _p(rays), rays.shape[0],
                        max_t, _p(max_t_per_ray), _p(hits), nthreads or (os.cpu_count() or 1))
    return hits


def ref_frame(scene, fp, materials, lights, rgba=None, nthreads=0, mode=0, reps=1, warmup=0, fast=False):
    """mode 0: scalar parity driver; 1: tiled_sched<ray8>; 2: tiled_sched<ray>. Returns (rgba, seconds)."""
    w, h = fp.width, fp.height
    if rgba is None:
        rgba = np.zeros((h, w, 4), dtype=np.float32)
    basis = np.zeros(12, dtype=np.float32)
    secs = ref_lib(fast).ref_frame(C.byref(fp), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes),
                                   scene.nodes.shape[0], _p(scene.indices), scene.indices.shape[0], _p(scene.normals),
                                   _p(materials), _p(lights), _p(rgba), nthreads or (os.cpu_count() or 1), mode, reps,
                                   warmup, _p(basis))
    return rgba, secs, basis


def ref_frame_matrix(scene, fp, view, proj, materials, lights, rgba=None, depth=None, want_depth=True, nthreads=0):
    """Scalar parity driver of the reference headers with matrix sched_params; returns (rgba, depth, inverses[32])."""
    w, h = fp.width, fp.height
    if rgba is None:
        rgba = np.zeros((h, w, 4), dtype=np.float32)
    if depth is None and want_depth:
        depth = np.zeros((h, w), dtype=np.float32)
    view = np.ascontiguousarray(view, dtype=np.float32).reshape(16)
    proj = np.ascontiguousarray(proj, dtype=np.float32).reshape(16)
    inv = np.zeros(32, dtype=np.float32)
    ref_lib().ref_frame_matrix(C.byref(fp), _p(view), _p(proj), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes),
                               scene.nodes.shape[0], _p(scene.indices), scene.indices.shape[0], _p(scene.normals),
                               _p(materials), _p(lights), _p(rgba), _p(depth), nthreads or (os.cpu_count() or 1), _p(inv))
    return rgba, depth, inv


def ref_camera_matrices(fp, z_near=0.001, z_far=1000.0):
    """The reference camera's view / projection matrices for fp's look_at + perspective."""
    view = np.zeros(16, dtype=np.float32)
    proj = np.zeros(16, dtype=np.float32)
    ref_lib().ref_camera_matrices(C.byref(fp), z_near, z_far, _p(view), _p(proj))
    return view, proj


def ref_query_timed(scene, rays, any_hit=False, max_t=3.4028234663852886e38, nthreads=0, reps=1, warmup=0):
    nh = C.c_uint64(0)
    secs = ref_lib(True).ref_query_timed(int(any_hit), _p(scene.prims), scene.prims.shape[0], _p(scene.nodes),
                                         scene.nodes.shape[0], _p(scene.indices), scene.indices.shape[0], _p(rays),
                                         rays.shape[0], max_t, nthreads or (os.cpu_count() or 1), reps, warmup,
                                         C.byref(nh))
    return secs, nh.value


# ---- the reference's own CUDA path (oracle/refcuda/ref_cuda_harness.cu -> oracle/_ref/libvsnray_ref_cuda.so) ----
_ref_cuda = None


def ref_cuda_path() -> str:
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libvsnray_ref_cuda.so")


def ref_cuda_available() -> bool:
    return os.path.exists(ref_cuda_path())


def ref_cuda_whitted(scene, fp, light_pos, reps=1, warmup=0, want_image=True):
    """cuda_sched<basic_ray<float>>::frame() with the stock whitted::kernel on the GPU (needs a CUDA device): returns
    (rgba or None, ms per frame, build + upload ms).  The tree is the reference's own build<index_bvh>()."""
    global _ref_cuda
    if _ref_cuda is None:
        lib = C.CDLL(ref_cuda_path())
        lib.ref_cuda_whitted.restype = C.c_float
        lib.ref_cuda_whitted.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_float),
                                         C.c_float, C.POINTER(C.c_float), C.c_uint, C.c_float, C.POINTER(C.c_float), C.c_int, C.c_int,
                                         C.c_void_p, C.POINTER(C.c_float)]
        _ref_cuda = lib
    f3 = lambda v: (C.c_float * 3)(*[float(x) for x in v[:3]])
    img = np.zeros((fp.height, fp.width, 4), dtype=np.float32) if want_image else None
    bms = C.c_float(0)
    bg = (C.c_float * 4)(*[float(fp.bg[i]) for i in range(4)])
    ms = _ref_cuda.ref_cuda_whitted(_p(scene.prims), scene.prims.shape[0], fp.width, fp.height, f3(fp.eye), f3(fp.center), f3(fp.up),
                                    fp.fovy, f3(light_pos), fp.num_bounces, fp.epsilon, bg, reps, warmup,
                                    _p(img) if want_image else None, C.byref(bms))
    if ms < 0:
        raise RuntimeError("ref_cuda_whitted: CUDA error")
    return img, float(ms), float(bms.value)

