"""Python mirror of the host adaptor above the C ABI (include/vsnray_b200.h).

Names follow the reference's domain: a `Scene` is what `cuda_index_bvh<P>(index_bvh<P> const&)`
produces (bvh.h:344-350) -- the uploaded tree, re-encoded on the GPU into the wide BVH8;
`closest_hit` / `any_hit` replace the calls of detail/traverse_linear.inl:186-318 over a ray list;
`frame` replaces `cuda_sched<R>::frame()` (detail/cuda_sched.inl:307-320) for the built-in kernels.

Every call goes through libvsnray_b200.so.  There is no CPU fallback; errors raise RuntimeError
(the reference's convention for the path is std::runtime_error at setup).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N

FLT_MAX = 3.4028234663852886e38
TRAVERSE_WIDE, TRAVERSE_EXACT, TRAVERSE_WATERTIGHT = 0, 1, 2
MEM_DEVICE, MEM_HOST = 0, 1
KERNEL_SIMPLE, KERNEL_WHITTED, KERNEL_AO, KERNEL_PATHTRACING = 0, 1, 2, 3
PS_UNIFORM, PS_JITTERED, PS_JITTERED_BLEND = 0, 1, 2
TILE = 64


def _check(rc: int, what: str):
    if rc != 0:
        msg = N.cuda_lib().vsb_strerror(rc).decode()
        raise RuntimeError(f"{what} failed: {msg} ({rc})")


def init(device: int = 0):
    _check(N.cuda_lib().vsb_init(device), "vsb_init")


def shutdown():
    N.cuda_lib().vsb_shutdown()


def sync():
    """Wait for everything the library queued on its own stream."""
    _check(N.cuda_lib().vsb_sync(), "vsb_sync")


def tune(key: str, value: int):
    _check(N.cuda_lib().vsb_tune(key.encode(), int(value)), "vsb_tune")


def sm_count() -> int:
    n = C.c_int(0)
    _check(N.cuda_lib().vsb_device_sm_count(C.byref(n)), "vsb_device_sm_count")
    return n.value


def launch_count() -> int:
    n = C.c_uint64(0)
    N.cuda_lib().vsb_launch_count(C.byref(n))
    return n.value


def last_build_ms() -> dict:
    """Device time of the last Scene.build(): total / sort / clustering / layout + re-encode, iterations."""
    ms = (C.c_float * 4)()
    it = C.c_int(0)
    _check(N.cuda_lib().vsb_last_build_ms(ms, C.byref(it)), "vsb_last_build_ms")
    return dict(total_ms=ms[0], sort_ms=ms[1], cluster_ms=ms[2], layout_encode_ms=ms[3], iterations=it.value)


def last_kernel_ms() -> float:
    ms = C.c_float(0)
    _check(N.cuda_lib().vsb_last_kernel_ms(C.byref(ms)), "vsb_last_kernel_ms")
    return ms.value


class Scene:
    """Device-resident scene: reference-layout BVH2 (exact traversal) + re-encoded BVH8."""

    def __init__(self, host_scene):
        self.host = host_scene
        h = C.c_void_p(0)
        idx = host_scene.indices
        _check(N.cuda_lib().vsb_scene_create(
            host_scene.prim_kind, N.ptr(host_scene.prims), host_scene.prims.shape[0],
            N.ptr(host_scene.nodes), host_scene.nodes.shape[0],
            N.ptr(idx) if idx is not None else None, idx.shape[0] if idx is not None else 0, C.byref(h)),
            "vsb_scene_create")
        self.handle = h
        if host_scene.normals is not None:
            self.set_normals(host_scene.normals)

    @classmethod
    def build(cls, prim_kind: int, prims, normals=None) -> "Scene":
        """The BVH is built on the GPU from the primitives alone (vsb_scene_build): replaces
        build<index_bvh<P>>() (bvh.h:455, detail/bvh/{build.inl,sah.h}) + the upload.
        prims: numpy array (host) or torch CUDA tensor of 64-byte triangles / 48-byte spheres."""
        self = cls.__new__(cls)
        self.host = None
        h = C.c_void_p(0)
        if isinstance(prims, np.ndarray):
            p, n, mem = N.ptr(prims), prims.shape[0], MEM_HOST
        else:
            p, n, mem = C.c_void_p(prims.data_ptr()), prims.shape[0], MEM_DEVICE
        self.handle = None
        _check(N.cuda_lib().vsb_scene_build(prim_kind, p, n, mem, C.byref(h)), "vsb_scene_build")
        self.handle = h
        self.n_prims = n
        if normals is not None:
            self.set_normals(normals)
        return self

    def download_bvh(self):
        """(nodes, indices) of the scene's reference-layout tree (index_bvh<P>::nodes()/indices())."""
        st = self.stats()
        nodes = np.zeros(st["bvh2_nodes"], dtype=N.NODE_DTYPE)
        n_idx = self.n_prims if self.host is None else (self.host.indices.shape[0] if self.host.indices is not None else 0)
        idx = np.zeros(n_idx, dtype=np.uint32)
        _check(N.cuda_lib().vsb_scene_download_bvh(self.handle, N.ptr(nodes), nodes.shape[0],
                                                   N.ptr(idx) if n_idx else None, n_idx), "vsb_scene_download_bvh")
        return nodes, idx

    def set_normals(self, normals):
        """Per-face normals (n_prims x 4 float32) kept on the device with the scene."""
        if isinstance(normals, np.ndarray):
            p, n, mem = N.ptr(normals), normals.shape[0], MEM_HOST
        else:
            p, n, mem = C.c_void_p(normals.data_ptr()), normals.shape[0], MEM_DEVICE
        _check(N.cuda_lib().vsb_scene_set_normals(self.handle, p, n, mem), "vsb_scene_set_normals")

    def stats(self) -> dict:
        out = (C.c_uint64 * 7)()
        _check(N.cuda_lib().vsb_scene_stats(self.handle, out, 7), "vsb_scene_stats")
        keys = ["wide_nodes", "wide_node_bytes", "wide_prims", "wide_prim_bytes", "bvh2_depth", "bvh2_nodes", "wide_levels"]
        return dict(zip(keys, [int(v) for v in out]))

    def download_wide(self):
        st = self.stats()
        nodes = np.zeros((st["wide_nodes"], 20), dtype=np.uint32)
        prims = np.zeros(st["wide_prim_bytes"] // 4, dtype=np.uint32)
        _check(N.cuda_lib().vsb_scene_download_wide(self.handle, N.ptr(nodes), nodes.nbytes, N.ptr(prims), prims.nbytes),
               "vsb_scene_download_wide")
        return nodes, prims

    def close(self):
        if self.handle:
            N.cuda_lib().vsb_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _torch_stream(stream):
    """Device buffers are torch tensors: unless the caller names a stream, the call runs on torch's CURRENT stream,
    so it is ordered behind the kernels that produced its inputs and ahead of whatever consumes (or recycles) its
    outputs.  The legacy default stream has handle 0, which the C ABI reads as "the library's own stream": it is
    passed as cudaStreamLegacy (0x1) instead."""
    if stream is not None:
        return stream
    import torch

    h = torch.cuda.current_stream().cuda_stream
    return h if h else 1


def _host_array(a, what, dtype=None, itemsize=None):
    """numpy buffers cross the ABI as raw pointers: they must be C-contiguous and of the expected record type."""
    if not a.flags["C_CONTIGUOUS"]:
        raise ValueError(f"{what}: array must be C-contiguous")
    if dtype is not None and a.dtype != dtype:
        raise ValueError(f"{what}: dtype {a.dtype} where {dtype} is expected")
    if itemsize is not None and a.nbytes % itemsize:
        raise ValueError(f"{what}: {a.nbytes} bytes is not a whole number of {itemsize}-byte records")
    return a


def _device_tensor(a, what, itemsize=None):
    if not a.is_contiguous():
        raise ValueError(f"{what}: tensor must be contiguous")
    if itemsize is not None and (a.numel() * a.element_size()) % itemsize:
        raise ValueError(f"{what}: not a whole number of {itemsize}-byte records")
    return a


def _scene_array(scenes):
    arr = (C.c_void_p * len(scenes))(*[s.handle for s in scenes])
    return arr


def _query(fn, scenes, rays, max_t, max_t_per_ray, hits, traverse, stream):
    """rays / hits: numpy arrays (host memory, copies inside the call) or torch CUDA tensors."""
    if isinstance(scenes, Scene):
        scenes = [scenes]
    arr = _scene_array(scenes)
    if isinstance(rays, np.ndarray):
        n = rays.shape[0]
        _host_array(rays, "rays", itemsize=32)
        if rays.nbytes != n * 32:
            raise ValueError("rays: 32-byte records (ori, pad, dir, pad) expected")
        if hits is None:
            hits = np.zeros(n, dtype=N.HIT_DTYPE)
        _host_array(hits, "hits", itemsize=32)
        if hits.nbytes < n * 32:
            raise ValueError("hits: one 32-byte record per ray expected")
        if max_t_per_ray is not None:
            _host_array(max_t_per_ray, "max_t_per_ray", dtype=np.float32)
            if max_t_per_ray.size < n:
                raise ValueError("max_t_per_ray: one float per ray expected")
        mt = N.ptr(max_t_per_ray) if max_t_per_ray is not None else None
        _check(fn(arr, len(scenes), N.ptr(rays), n, max_t, mt, N.ptr(hits), MEM_HOST, traverse, stream), fn.__name__)
        return hits
    import torch

    n = rays.shape[0]
    _device_tensor(rays, "rays", itemsize=32)
    if rays.numel() * rays.element_size() != n * 32:
        raise ValueError("rays: 32-byte records (ori, pad, dir, pad) expected")
    if hits is None:
        hits = torch.empty((n, 8), dtype=torch.int32, device=rays.device)
    _device_tensor(hits, "hits", itemsize=32)
    if max_t_per_ray is not None:
        _device_tensor(max_t_per_ray, "max_t_per_ray")
        if max_t_per_ray.dtype != torch.float32 or max_t_per_ray.numel() < n:
            raise ValueError("max_t_per_ray: one float32 per ray expected")
    mt = C.c_void_p(max_t_per_ray.data_ptr()) if max_t_per_ray is not None else None
    _check(fn(arr, len(scenes), C.c_void_p(rays.data_ptr()), n, max_t, mt, C.c_void_p(hits.data_ptr()), MEM_DEVICE,
              traverse, _torch_stream(stream)), fn.__name__)
    return hits


def closest_hit(scenes, rays, max_t=FLT_MAX, max_t_per_ray=None, hits=None, traverse=TRAVERSE_WIDE, stream=None):
    return _query(N.cuda_lib().vsb_closest_hit, scenes, rays, max_t, max_t_per_ray, hits, traverse, stream)


def any_hit(scenes, rays, max_t=FLT_MAX, max_t_per_ray=None, hits=None, traverse=TRAVERSE_WIDE, stream=None):
    return _query(N.cuda_lib().vsb_any_hit, scenes, rays, max_t, max_t_per_ray, hits, traverse, stream)


def multi_hit(scenes, rays, max_hits=16, hits=None, traverse=TRAVERSE_WIDE, stream=None):
    """multi_hit<max_hits>(r, begin, end) over a ray list (detail/traverse_linear.inl:325-380): per ray
    the max_hits closest intersections sorted by t, misses behind.  numpy rays -> (n, max_hits) HIT_DTYPE
    array; torch CUDA rays -> (n, max_hits, 8) int32 tensor."""
    if isinstance(scenes, Scene):
        scenes = [scenes]
    arr = _scene_array(scenes)
    n = rays.shape[0]
    if isinstance(rays, np.ndarray):
        if hits is None:
            hits = np.zeros((n, max_hits), dtype=N.HIT_DTYPE)
        _check(N.cuda_lib().vsb_multi_hit(arr, len(scenes), N.ptr(rays), n, max_hits, N.ptr(hits), MEM_HOST, traverse, stream),
               "vsb_multi_hit")
        return hits
    import torch

    if hits is None:
        hits = torch.empty((n, max_hits, 8), dtype=torch.int32, device=rays.device)
    _device_tensor(rays, "rays", itemsize=32); _device_tensor(hits, "hits", itemsize=32)
    _check(N.cuda_lib().vsb_multi_hit(arr, len(scenes), C.c_void_p(rays.data_ptr()), n, max_hits,
                                      C.c_void_p(hits.data_ptr()), MEM_DEVICE, traverse, _torch_stream(stream)), "vsb_multi_hit")
    return hits


def tiles_of_rank(width, height, rank, nranks) -> int:
    n = C.c_int(0)
    _check(N.cuda_lib().vsb_tiles_of_rank(width, height, rank, nranks, C.byref(n)), "vsb_tiles_of_rank")
    return n.value


class DeviceBuffer:
    """A raw device allocation of the library (cudaMalloc), shareable between the per-GPU processes
    of a node through CUDA IPC: `export()` on the owner, `DeviceBuffer.open(handle)` on the peers."""

    def __init__(self, nbytes: int = 0, ptr: int = 0, peer: bool = False):
        self.nbytes, self.peer = nbytes, peer
        if ptr:
            self.ptr = ptr
        else:
            p = C.c_void_p(0)
            _check(N.cuda_lib().vsb_device_alloc(nbytes, C.byref(p)), "vsb_device_alloc")
            self.ptr = p.value

    def data_ptr(self) -> int:
        return self.ptr

    def export(self) -> bytes:
        h = C.create_string_buffer(64)
        _check(N.cuda_lib().vsb_ipc_export(C.c_void_p(self.ptr), h), "vsb_ipc_export")
        return h.raw

    @staticmethod
    def open(handle: bytes, nbytes: int = 0) -> "DeviceBuffer":
        p = C.c_void_p(0)
        _check(N.cuda_lib().vsb_ipc_open(C.create_string_buffer(handle, 64), C.byref(p)), "vsb_ipc_open")
        return DeviceBuffer(nbytes, p.value, peer=True)

    def download(self, out: np.ndarray):
        _check(N.cuda_lib().vsb_device_copy(N.ptr(out), C.c_void_p(self.ptr), out.nbytes, MEM_HOST, MEM_DEVICE), "vsb_device_copy")
        return out

    def close(self):
        if self.ptr:
            (N.cuda_lib().vsb_ipc_close if self.peer else N.cuda_lib().vsb_device_free)(C.c_void_p(self.ptr))
            self.ptr = 0


def set_surface(scene: "Scene", surface) -> None:
    """vsb_scene_set_surface: attach a scenes.SurfaceData (host arrays) to the device scene -- the
    normals / colours / tex coords / textures of make_kernel_params (kernels.h:395-509); None detaches."""
    if surface is None:
        _check(N.cuda_lib().vsb_scene_set_surface(scene.handle, None, MEM_HOST), "vsb_scene_set_surface")
        return
    d = N.SurfaceDesc()
    d.normal_binding, d.color_binding = surface.normal_binding, surface.color_binding
    d.normals = surface.normals.ctypes.data if surface.normals is not None else None
    d.colors = surface.colors.ctypes.data if surface.colors is not None else None
    d.tex_coords = surface.tex_coords.ctypes.data if surface.tex_coords is not None else None
    texs = None
    if surface.textures:
        texs = (N.Texture * len(surface.textures))()
        for i, t in enumerate(surface.textures):
            if t is None:
                continue
            texs[i].data = t.data.ctypes.data
            texs[i].height, texs[i].width = t.data.shape[0], t.data.shape[1]
            texs[i].filter = t.filter
            texs[i].address[0], texs[i].address[1] = t.address
        d.textures = C.cast(texs, C.c_void_p)
        d.n_textures = len(surface.textures)
    _check(N.cuda_lib().vsb_scene_set_surface(scene.handle, C.byref(d), MEM_HOST), "vsb_scene_set_surface")


def frame(scenes, fp: N.FrameParams, normals, materials, lights, color, traverse=TRAVERSE_WIDE, tile_rank=0,
          tile_nranks=0, ray_count=None, stream=None, tiles_in_place=False, view_matrix=None, proj_matrix=None,
          depth=None, generic_materials=False, color_format=0, spot_lights=False, mirror=None):
    """One frame() call.  `color` (and `ray_count`): numpy array -> host render target (D2H inside
    the call, returns with pixels in place); torch CUDA tensor -> device render target
    (asynchronous on `stream`).  `materials` / `lights` (and `normals`, which may be None to use the
    normals attached to the scene) are host or device buffers independently of the target.
    `view_matrix` / `proj_matrix` (16 floats each, column-major): sched_params built from matrices
    instead of fp's camera; with them `depth` (float32 per pixel, same kind of buffer as `color`)
    receives the PF_DEPTH32F target."""
    if isinstance(scenes, Scene):
        scenes = [scenes]
    d = N.FrameDesc()
    d.p = fp
    _check(N.cuda_lib().vsb_camera_basis(C.byref(d)), "vsb_camera_basis")
    d.scissor[0], d.scissor[1], d.scissor[2], d.scissor[3] = 0, 0, fp.width, fp.height
    host = isinstance(color, np.ndarray)

    def p(a):
        if a is None:
            return None
        if isinstance(a, np.ndarray):
            return _host_array(a, "frame buffer / parameter array").ctypes.data
        if hasattr(a, "is_contiguous"):
            _device_tensor(a, "frame buffer / parameter tensor")
        return a.data_ptr()

    d.normals, d.materials, d.lights, d.color = p(normals), p(materials), p(lights), p(color)
    d.depth = p(depth)
    if depth is not None and isinstance(depth, np.ndarray) != host:
        raise ValueError("color and depth must live in the same memory")
    if (view_matrix is None) != (proj_matrix is None):
        raise ValueError("view_matrix and proj_matrix come together")
    if view_matrix is not None:
        d.camera_kind = 1
        v = np.ascontiguousarray(view_matrix, dtype=np.float32).reshape(16)
        q = np.ascontiguousarray(proj_matrix, dtype=np.float32).reshape(16)
        for i in range(16):
            d.view_matrix[i], d.proj_matrix[i] = float(v[i]), float(q[i])
    d.mem = MEM_HOST if host else MEM_DEVICE
    params = [a for a in (normals, materials, lights) if a is not None]
    kinds = {isinstance(a, np.ndarray) for a in params}
    if len(kinds) > 1:
        raise ValueError("normals / materials / lights must all be host arrays or all device tensors")
    d.params_mem = MEM_HOST if (not kinds or kinds.pop()) else MEM_DEVICE
    d.traverse = traverse
    d.tile_rank, d.tile_nranks = tile_rank, tile_nranks
    d.tile_target = 1 if tiles_in_place else 0
    d.generic_materials = 1 if generic_materials else 0
    d.spot_lights = 1 if spot_lights else 0
    d.color_format = color_format          # PF_RGBA32F 0, PF_RGBA8 1, PF_RGB32F 2 (16-byte pixels), PF_RGB8 3
    d.ray_count = p(ray_count)
    d.mirror_color = p(mirror)             # compact tile buffers only: the whole row-major frame every finished pixel is also stored into
    uses_torch = any(hasattr(a, "is_contiguous") for a in (normals, materials, lights, color, depth, ray_count) if a is not None)
    _check(N.cuda_lib().vsb_frame(_scene_array(scenes), len(scenes), C.byref(d), _torch_stream(stream) if uses_torch else stream), "vsb_frame")
    return color


def deinterleave_tiles(gathered, width, height, nranks, tiles_per_rank_padded, image, stream=None):
    _check(N.cuda_lib().vsb_deinterleave_tiles(C.c_void_p(gathered.data_ptr()), width, height, nranks,
                                               tiles_per_rank_padded, C.c_void_p(image.data_ptr()), _torch_stream(stream)),
           "vsb_deinterleave_tiles")
    return image
