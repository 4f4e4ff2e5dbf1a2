"""ctypes bindings of the two in-tree shared libraries.

* ``lib/libvsnray_b200.so``       -- the product: hand-written sm_100a CUDA kernels behind the
  C ABI declared in ``include/vsnray_b200.h``.
* ``lib/libvsnray_b200_host.so``  -- host C++ layer (scene generators, binned-SAH builder).

There is no CPU fallback for the ray-query path: if the CUDA library is missing, loading it
raises, and every compute entry point fails loudly when no CUDA device is present.
Nothing in this package touches ``oracle/``.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_DIR = os.path.join(_HERE, "lib")

HIT_DTYPE = np.dtype(
    [("hit", "<i4"), ("prim_id", "<u4"), ("geom_id", "<u4"), ("t", "<f4"), ("u", "<f4"), ("v", "<f4"),
     ("prim_list_index", "<u4"), ("pad", "<u4")]
)
NODE_DTYPE = np.dtype(
    [("bbox_min", "<f4", (3,)), ("first", "<u4"), ("bbox_max", "<f4", (3,)), ("num_prims", "<u4")]
)
MATERIAL_DTYPE = np.dtype(
    [("ca", "<f4", (3,)), ("ka", "<f4"), ("cd", "<f4", (3,)), ("kd", "<f4"), ("cs", "<f4", (3,)), ("ks", "<f4"),
     ("spec_exp", "<f4"), ("kind", "<i4"), ("pad", "<f4", (2,))]
)
LIGHT_DTYPE = np.dtype(
    [("pos", "<f4", (3,)), ("kl", "<f4"), ("cl", "<f4", (3,)), ("att_const", "<f4"), ("att_lin", "<f4"),
     ("att_quad", "<f4"), ("kind", "<i4"), ("spot_exponent", "<f4"), ("spot_dir", "<f4", (3,)), ("spot_cos_cutoff", "<f4")]
)
assert HIT_DTYPE.itemsize == 32 and NODE_DTYPE.itemsize == 32
assert MATERIAL_DTYPE.itemsize == 64 and LIGHT_DTYPE.itemsize == 64


class FrameParams(C.Structure):
    """vsb_frame_params (include/vsnray_b200.h); same layout as the oracle's vso_frame_params."""

    _fields_ = [
        ("kernel", C.c_int32), ("pixel_sampler", C.c_int32), ("width", C.c_int32), ("height", C.c_int32),
        ("eye", C.c_float * 4), ("center", C.c_float * 4), ("up", C.c_float * 4),
        ("fovy", C.c_float), ("aspect", C.c_float),
        ("frame_num", C.c_uint32), ("seed", C.c_uint32), ("num_bounces", C.c_uint32), ("epsilon", C.c_float),
        ("bg", C.c_float * 4), ("ambient", C.c_float * 4),
        ("ao_samples", C.c_int32), ("ao_radius", C.c_float), ("n_lights", C.c_int32), ("n_materials", C.c_int32),
    ]


class FrameDesc(C.Structure):
    """vsb_frame_desc (include/vsnray_b200.h)."""

    _fields_ = [
        ("p", FrameParams),
        ("cam_eye", C.c_float * 4), ("cam_u", C.c_float * 4), ("cam_v", C.c_float * 4), ("cam_w", C.c_float * 4),
        ("scissor", C.c_int32 * 4),
        ("normals", C.c_void_p), ("materials", C.c_void_p), ("lights", C.c_void_p),
        ("color", C.c_void_p), ("depth", C.c_void_p),
        ("mem", C.c_int32), ("traverse", C.c_int32), ("tile_rank", C.c_int32), ("tile_nranks", C.c_int32),
        ("ray_count", C.c_void_p),
        ("params_mem", C.c_int32), ("tile_target", C.c_int32),
        ("camera_kind", C.c_int32), ("view_matrix", C.c_float * 16), ("proj_matrix", C.c_float * 16),
        ("color_format", C.c_int32), ("generic_materials", C.c_int32), ("spot_lights", C.c_int32),
        ("mirror_color", C.c_void_p),
    ]


class Texture(C.Structure):
    """vsb_texture (include/vsnray_b200.h)."""

    _fields_ = [("data", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32), ("filter", C.c_int32),
                ("address", C.c_int32 * 2)]


class SurfaceDesc(C.Structure):
    """vsb_surface_desc (include/vsnray_b200.h)."""

    _fields_ = [("normal_binding", C.c_int32), ("color_binding", C.c_int32), ("normals", C.c_void_p),
                ("colors", C.c_void_p), ("tex_coords", C.c_void_p), ("textures", C.c_void_p), ("n_textures", C.c_int32)]


# every symbol include/vsnray_b200.h declares (tests check the library exports them all)
ABI_SYMBOLS = [
    "vsb_init", "vsb_shutdown", "vsb_sync", "vsb_strerror", "vsb_version", "vsb_device_sm_count", "vsb_tune",
    "vsb_scene_create", "vsb_scene_build", "vsb_scene_download_bvh", "vsb_last_build_ms", "vsb_scene_destroy", "vsb_scene_set_normals", "vsb_scene_set_surface", "vsb_scene_stats", "vsb_scene_download_wide",
    "vsb_closest_hit", "vsb_any_hit", "vsb_multi_hit", "vsb_closest_hit_static", "vsb_camera_basis", "vsb_frame", "vsb_tiles_of_rank",
    "vsb_deinterleave_tiles", "vsb_device_alloc", "vsb_device_free", "vsb_device_copy", "vsb_ipc_export", "vsb_ipc_open", "vsb_ipc_close", "vsb_last_kernel_ms", "vsb_launch_count",
]

_cuda_lib = None
_host_lib = None


def cuda_lib_path() -> str:
    # VSB_CUDA_LIB: developer hook to load a tuning variant of the same library (tools/build_variant.sh)
    return os.environ.get("VSB_CUDA_LIB") or os.path.join(LIB_DIR, "libvsnray_b200.so")


def host_lib_path() -> str:
    return os.path.join(LIB_DIR, "libvsnray_b200_host.so")


def cuda_lib() -> C.CDLL:
    """Load libvsnray_b200.so (raises if it was not built: there is no fallback)."""
    global _cuda_lib
    if _cuda_lib is None:
        path = cuda_lib_path()
        if not os.path.exists(path):
            raise RuntimeError(
                f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
                "the ray-query path has no CPU fallback"
            )
        lib = C.CDLL(path)
        vp, sz, u32p = C.c_void_p, C.c_size_t, C.POINTER(C.c_uint32)
        lib.vsb_init.argtypes = [C.c_int]
        lib.vsb_strerror.restype = C.c_char_p
        lib.vsb_strerror.argtypes = [C.c_int]
        lib.vsb_device_sm_count.argtypes = [C.POINTER(C.c_int)]
        lib.vsb_tune.argtypes = [C.c_char_p, C.c_int]
        for kv in os.environ.get("VSB_TUNE", "").split(","):      # e.g. VSB_TUNE=frame_service_min=32,phase_min=8
            if "=" in kv:
                k, v = kv.split("=")
                if lib.vsb_tune(k.strip().encode(), int(v)) != 0:
                    raise RuntimeError(f"VSB_TUNE: bad setting {kv!r}")
        lib.vsb_scene_create.argtypes = [C.c_int, vp, sz, vp, sz, vp, sz, C.POINTER(vp)]
        lib.vsb_scene_destroy.argtypes = [vp]
        lib.vsb_scene_build.argtypes = [C.c_int, vp, sz, C.c_int, C.POINTER(vp)]
        lib.vsb_scene_download_bvh.argtypes = [vp, vp, sz, vp, sz]
        lib.vsb_last_build_ms.argtypes = [C.POINTER(C.c_float), C.POINTER(C.c_int)]
        lib.vsb_scene_set_normals.argtypes = [vp, vp, sz, C.c_int]
        lib.vsb_scene_set_surface.argtypes = [vp, C.POINTER(SurfaceDesc), C.c_int]
        lib.vsb_scene_stats.argtypes = [vp, C.POINTER(C.c_uint64), C.c_int]
        lib.vsb_scene_download_wide.argtypes = [vp, vp, sz, vp, sz]
        for f in (lib.vsb_closest_hit, lib.vsb_any_hit):
            f.argtypes = [C.POINTER(vp), C.c_int, vp, sz, C.c_float, vp, vp, C.c_int, C.c_int, vp]
        lib.vsb_multi_hit.argtypes = [C.POINTER(vp), C.c_int, vp, sz, C.c_int, vp, C.c_int, C.c_int, vp]
        lib.vsb_closest_hit_static.argtypes = [vp, vp, sz, C.c_float, vp, vp]
        lib.vsb_camera_basis.argtypes = [C.POINTER(FrameDesc)]
        lib.vsb_frame.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(FrameDesc), vp]
        lib.vsb_tiles_of_rank.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
        lib.vsb_deinterleave_tiles.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
        lib.vsb_device_alloc.argtypes = [sz, C.POINTER(vp)]
        lib.vsb_device_free.argtypes = [vp]
        lib.vsb_device_copy.argtypes = [vp, vp, sz, C.c_int, C.c_int]
        lib.vsb_ipc_export.argtypes = [vp, vp]
        lib.vsb_ipc_open.argtypes = [vp, C.POINTER(vp)]
        lib.vsb_ipc_close.argtypes = [vp]
        lib.vsb_last_kernel_ms.argtypes = [C.POINTER(C.c_float)]
        lib.vsb_launch_count.argtypes = [C.POINTER(C.c_uint64)]
        _cuda_lib = lib
    return _cuda_lib


def host_lib() -> C.CDLL:
    global _host_lib
    if _host_lib is None:
        path = host_lib_path()
        if not os.path.exists(path):
            raise RuntimeError(f"{path} is missing: run __graft_entry__.build()")
        lib = C.CDLL(path)
        vp, sz = C.c_void_p, C.c_size_t
        lib.vsbh_gen_triangles.restype = sz
        lib.vsbh_gen_triangles.argtypes = [C.c_int, C.c_long, C.c_long, C.c_uint64, C.c_float, vp, sz]
        lib.vsbh_gen_spheres.restype = sz
        lib.vsbh_gen_spheres.argtypes = [sz, C.c_uint64, C.c_float, C.c_float, vp]
        lib.vsbh_gen_random_rays.argtypes = [vp, sz, C.c_uint64, C.c_float, C.c_float]
        lib.vsbh_face_normals.argtypes = [vp, sz, vp]
        lib.vsbh_primary_rays.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, vp]
        lib.vsbh_build_bvh.argtypes = [C.c_int, vp, sz, vp, sz, C.POINTER(sz), vp, sz, C.POINTER(sz)]
        lib.vsbh_obj_load.restype = vp
        lib.vsbh_obj_load.argtypes = [C.c_char_p]
        lib.vsbh_obj_free.argtypes = [vp]
        lib.vsbh_obj_counts.argtypes = [vp, C.POINTER(sz)]
        lib.vsbh_obj_copy.argtypes = [vp, vp, vp, vp, vp, vp, vp]
        lib.vsbh_obj_texture.argtypes = [vp, sz, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(vp)]
        lib.vsbh_obj_warning.restype = sz
        lib.vsbh_obj_warning.argtypes = [vp, sz, C.c_char_p, sz]
        lib.vsbh_spot_cos_cutoff.restype = C.c_float
        lib.vsbh_spot_cos_cutoff.argtypes = [C.c_float]
        lib.vsbh_build_bvh_ex.argtypes = [C.c_int, vp, sz, C.c_int, vp, sz, C.POINTER(sz), vp, sz, C.POINTER(sz)]
        _host_lib = lib
    return _host_lib


def ptr(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)
