#!/usr/bin/env python
"""bench.py -- Mrays/s of the ray-query hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg1|cfg3|cfg4|cfg5] [--impl reference]

A "step" is one frame() call over the workload's synthetic scene (one pass of ray generation +
closest_hit/any_hit traversal + shading + store).  Default workload = BASELINE.json configs[1]:
4096x4096 primary rays + Whitted shading vs the 1,002,528-triangle displaced grid.

  value     rays traced / device time of the timed region, scene + parameter arrays + framebuffer
            resident in HBM (rays are counted by the kernel: primary + shadow + bounce rays)
  e2e       the same metric through the host-buffer call a user of the reference makes (frame() into a
            host render target): parameter H2D + framebuffer D2H inside the region.  N > 1: ONE assembled
            frame in a pinned host buffer that all ranks map (rank 0 reads it)
  roofline  node-fetch memory roofline of SURVEY.md section 8d: B_ray = 64 N_inner + 68 N_prim + 64
            bytes per ray on the REFERENCE data layout (N_* counted by the oracle's instrumented
            traversal of the scene's BVH2 on a down-sampled frame of the same view), against
            MEASURED_PEAKS.json hbm_gbs.  The kernel is bound by issue slots, not HBM: `bound` says so,
            `issue` carries the binding figures from the committed ncu capture (profiles/issue.json)
  cpu_baseline  the reference's own tiled_sched<simd::ray8> (oracle/_ref, unmodified headers) on the
            host cores, on the SAME frame as the GPU arm
  extra     (N = 1, default run) the incoherent workloads measured the same way: cfg5 closest / any, cfg3;
            ploc_tree: cfg2 and cfg5 on the tree of the second GPU builder (Morton + PLOC), with build time and time to first frame;
            naive_gpu: the same frame by the reference's own cuda_sched kernel on this GPU (unmodified headers)
  strong    (N > 1) the fixed 4096^2 frame partitioned over the N GPUs: cfg2 and cfg4 (progressive path tracer), tiles
            stored into rank 0's frame over NVLink from inside the trace kernel, with v_N / (N v_1)

N > 1 (torchrun): the frame is partitioned into interleaved 64x64 tiles (tile k -> rank k % N), the
scene is replicated.  Headline = weak scaling: the same view at sqrt(N) x the resolution per axis
(4096^2 pixels per GPU).  The tree is the reference's binned-SAH tree re-encoded on the GPU as a BVH8, as BASELINE's
configs say; by default it is BUILT on the GPU too (vsb_scene_build, the host builder's tree node for node; --builder reference
uses the host builder, extra.ploc_tree measures the Morton + PLOC builder beside it).
"""
from __future__ import annotations

import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SMI_QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
             "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

KERNEL_NAMES = {0: "simple", 1: "whitted", 2: "ao", 3: "pathtracing"}
# name -> (scene, kernel, base width, base height, eye, center, frame kwargs, material kind)
WORKLOADS = {
    "cfg1": dict(scene="cfg1_sphere", kernel=0, w=1024, h=1024, eye=(0.0, 0.0, 3.0), mat=0,
                 desc="1024x1024 primary closest-hit, simple kernel, 99,904-tri UV sphere"),
    "cfg2": dict(scene="cfg2_grid", kernel=1, w=4096, h=4096, eye=(0.3, 1.2, 2.2), mat=0,
                 desc="4096x4096 primary + Whitted shading, 1,002,528-tri displaced grid"),
    "cfg3": dict(scene="cfg3_cubes", kernel=2, w=2048, h=2048, eye=(0.6, 1.4, 2.6), mat=0,
                 desc="2048x2048 AO any-hit 16 rays/pixel, 5,000,004-tri cube field"),
    "cfg4": dict(scene="cfg4_terrain", kernel=3, w=4096, h=4096, eye=(0.3, 1.2, 2.2), mat=1,
                 desc="4096x4096 path tracer 8 diffuse bounces, ~10M-tri terrain"),
}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={SMI_QUERY}", "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        rows = [l for (t, l) in self.lines if t0 <= t <= t1 + 0.1] or [l for (_, l) in self.lines[-3:]]
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in rows:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); smax.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_gbs():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            with open(path) as f:
                return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def _profile_json(name: str, key: str):
    try:
        with open(os.path.join(ROOT, "profiles", name)) as f:
            return json.load(f).get(key)
    except Exception:
        return None


def measured_traffic(key: str):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/traffic.json), or None."""
    e = _profile_json("traffic.json", key)
    return int(e["bytes"]) if e else None


def measured_issue(key: str):
    """What binds the kernel, from the committed ncu capture of the same command (profiles/issue.json):
    issue-slot utilisation, active threads per instruction, their product, the alu pipe."""
    return _profile_json("issue.json", key)


def frame_shape(wl: dict, n_gpus: int):
    """Weak scaling: the same view at sqrt(N) x the resolution per axis (rounded to whole 64-pixel
    tiles), i.e. base-frame pixels per GPU and the same mix of sky / geometry / secondary rays."""
    if n_gpus <= 1:
        return wl["w"], wl["h"]
    f = math.sqrt(n_gpus)
    return int(round(wl["w"] * f / 64.0)) * 64, int(round(wl["h"] * f / 64.0)) * 64


def make_frame_params(S, wl: dict, w: int, h: int, frame_num: int = 1):
    kernel = wl["kernel"]
    kw = dict(pixel_sampler=0, num_bounces=4, epsilon=1e-3)
    if kernel == 3:
        kw = dict(pixel_sampler=2, num_bounces=8, epsilon=1e-3, ambient=(1.0, 1.0, 1.0, 1.0))
    if kernel == 2:
        kw = dict(pixel_sampler=0, ao_samples=16, ao_radius=0.1)
    return S.frame_params(kernel, w, h, wl["eye"], (0.0, 0.0, 0.0), frame_num=frame_num, **kw)


def workload_config(wl: dict, w: int, h: int, scale: float, n_prims: int) -> dict:
    """The part of `config` that names the workload: identical in the GPU arm and in --impl reference."""
    return {"workload": wl["desc"], "frame": [w, h], "scale": scale, "kernel": KERNEL_NAMES[wl["kernel"]],
            "triangles": int(n_prims),
            "l2": f"inputs larger than L2: framebuffer {w * h * 16 >> 20} MiB rewritten every step + {n_prims * 64 >> 20} MiB of "
                  f"triangles (126 MB L2)"}


def raylist_config(n_prims: int, any_hit: bool, n_batch: int, scale: float) -> dict:
    return {"workload": f"incoherent random rays vs {n_prims}-tri soup, {'any' if any_hit else 'closest'}-hit, batches of {n_batch} rays per GPU",
            "scale": scale, "triangles": int(n_prims),
            "l2": f"inputs larger than L2: 4 rotating ray/hit batches ({4 * n_batch * 64 >> 20} MiB) + {n_prims * 64 >> 20} MiB of triangles (126 MB L2)"}


# ------------------------------------------------------------------------------------------------
# CPU side: algorithmic bytes per ray (oracle counters), cpu_baseline, and the --impl reference arm
# ------------------------------------------------------------------------------------------------

def algorithmic_bytes_per_ray(O, S, sc, wl, mats, lights, sample_wh=(512, 512)):
    """B_ray = 64 N_inner + 68 N_prim + 64 from the oracle's instrumented scalar
    traversal of the scene's BVH2 (reference layout) on a down-sampled frame of the same view."""
    fp = make_frame_params(S, wl, *sample_wh)
    c = O.Counters()
    O.oracle_frame(sc, fp, mats, lights, counters=c)
    n = max(int(c.n_rays), 1)
    n_inner, n_prim = c.n_inner / n, c.n_prim / n
    prim_bytes = 68.0 if sc.prim_kind == 0 else 52.0
    return 64.0 * n_inner + prim_bytes * n_prim + 64.0, dict(n_inner=n_inner, n_prim=n_prim, rays_per_pixel=n / (sample_wh[0] * sample_wh[1]))


def cpu_frame_seconds(O, S, sc, wl, mats, lights, w, h, reps, warmup, cores):
    """Seconds per frame() of the reference's tiled_sched<simd::ray8> (oracle/_ref, unmodified headers, -O3 -mavx2
    -mfma) on `cores` host threads; the oracle port (scalar C, pthreads) when _ref is absent.  -> (secs, kind)"""
    fp = make_frame_params(S, wl, w, h)
    if O.ref_available(fast=True):
        _, secs, _ = O.ref_frame(sc, fp, mats, lights, mode=1, reps=reps, warmup=warmup, fast=True, nthreads=cores)
        return secs, "reference"
    best = None
    for _ in range(max(reps, 1)):
        t0 = time.perf_counter()
        O.oracle_frame(sc, fp, mats, lights)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return best, "port"


def frame_ray_count(O, S, sc, wl, mats, lights, w, h):
    """Rays of one frame (primary + shadow + bounce), counted by the oracle on a frame of 1/4 the resolution per axis
    of the same view and scaled by 16 (the full-frame count is what the GPU arm's kernel counter reports; the CPU
    side uses this estimate so that counting does not cost more than the timed frames)."""
    dw, dh = max(w // 4, 64), max(h // 4, 64)
    c = O.Counters()
    O.oracle_frame(sc, make_frame_params(S, wl, dw, dh), mats, lights, counters=c)
    return int(round(int(c.n_rays) * (w * h) / float(dw * dh)))


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores, on the SAME frame
    as the GPU arm at this N (the weak-scaled frame for N > 1; rank 0 alone runs)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import __graft_entry__ as g
    g.build_host_only()
    from oracle import bindings as O
    from visionaray_b200 import scenes as S

    cores = os.cpu_count() or 1
    if args.workload == "cfg5":
        return run_reference_raylist(args, O, S, cores)
    wl = WORKLOADS[args.workload]
    w, h = frame_shape(wl, args.gpus)
    sc = S.make_scene(wl["scene"], args.scale)            # the reference's binned-SAH tree (host builder)
    mats, lights = S.default_material(wl["mat"]), S.point_light(wl["eye"])
    rays = frame_ray_count(O, S, sc, wl, mats, lights, w, h)
    kind = "reference" if O.ref_available(fast=True) else "port"
    times = []
    for i in range(args.warmup + args.steps):
        secs, kind = cpu_frame_seconds(O, S, sc, wl, mats, lights, w, h, reps=1, warmup=0, cores=cores)
        if i >= args.warmup:
            times.append(secs)
    ms = 1e3 * float(np.mean(times))
    value = rays / (ms * 1e-3) / 1e6
    how = "tiled_sched<simd::ray8> AVX2 (unmodified headers)" if kind == "reference" else "scalar C port, pthreads"
    sample = f"the whole {w}x{h} frame per step ({rays} rays, counted on a {max(w // 4, 64)}x{max(h // 4, 64)} frame of the view x 16), {how}"
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(wl, w, h, args.scale, sc.n_prims),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_reference_raylist(args, O, S, cores):
    """cfg5 on the host cores: scalar closest_hit()/any_hit() of the unmodified headers over std::threads."""
    sc = S.make_scene("cfg5_soup", args.scale)
    n = min(args.batch, 2_000_000)
    rays = S.random_rays(n, 600)
    times = []
    kind = "reference" if O.ref_available(fast=True) else "port"
    for i in range(args.warmup + args.steps):
        if kind == "reference":
            secs, _ = O.ref_query_timed(sc, rays, bool(args.any), nthreads=cores, reps=1, warmup=0)
        else:
            t0 = time.perf_counter()
            O.oracle_query(sc, rays, bool(args.any))
            secs = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(secs)
    ms = 1e3 * float(np.mean(times))
    value = n / (ms * 1e-3) / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": raylist_config(sc.n_prims, bool(args.any), args.batch, args.scale),
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind,
                         "sample": f"{n} rays of the batch per step, scalar closest_hit()/any_hit() on {cores} threads"},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm: scene preparation and the context every leg shares
# ------------------------------------------------------------------------------------------------

GPU_BUILDERS = {"gpu": (1, "reference binned SAH, built breadth-first on the GPU (vsb_scene_build, build_sah.cuh): the host builder's tree node for node"),
                "gpu_ploc": (0, "gpu (Morton sort + PLOC + SAH leaf collapse)")}


def prepare_scene(api, S, scene_name: str, scale: float, builder: str):
    """-> (host scene, device scene, builder info incl. time to a traceable scene).
    --builder gpu (default): primitives only on the host; vsb_scene_build builds the REFERENCE's binned-SAH tree on the
    GPU (the same nodes and indices as the host builder: tests/test_gpu_sah_builder.py) and re-encodes it as a BVH8.
    --builder reference: the same tree from the host builder, uploaded and re-encoded.  --builder gpu_ploc: the
    Morton + PLOC tree (about the same build time, traces 5-14 % slower)."""
    t0 = time.perf_counter()
    sc = S.make_scene(scene_name, scale, build_tree=(builder == "reference"))
    gen_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    if builder in GPU_BUILDERS:
        api.tune("build_algorithm", GPU_BUILDERS[builder][0])
        dev = api.Scene.build(sc.prim_kind, sc.prims, sc.normals)
        api.sync()
        ms = api.last_build_ms()
        info = {"kind": GPU_BUILDERS[builder][1], "build_ms": round(ms["total_ms"], 2), "iterations": ms["iterations"]}
    else:
        dev = api.Scene(sc)
        api.sync()
        info = {"kind": "reference binned SAH (host builder, inside scene generation)"}
    info["scene_generation_s"] = round(gen_s, 3)
    info["upload_and_build_s"] = round(time.perf_counter() - t0, 3)      # H2D of the primitives + build + re-encode, wall clock
    return sc, dev, info


def oracle_scene(S, sc, dev):
    """The scene with its BVH2 in the reference layout, for the oracle's instrumented traversal (the B_ray
    counters): the host tree if there is one, else the GPU-built tree downloaded (bvh_node[] + indices[])."""
    if sc.nodes is not None:
        return sc
    nodes, idx = dev.download_bvh()
    return S.HostScene(sc.prim_kind, sc.prims, nodes, idx, sc.normals, sc.name)


class Ctx:
    """What every GPU leg shares: modules, ranks, the main stream."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        self.torch, self.dist, self.args = torch, dist, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device -- the ray-query path has no CPU fallback")
        torch.cuda.set_device(self.local_rank)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local_rank))
        from visionaray_b200 import api, scenes as S

        self.api, self.S = api, S
        api.init(self.local_rank)
        # a real (non-legacy) stream: the C ABI treats a NULL stream as "the library's own stream"
        self.main = torch.cuda.Stream()
        torch.cuda.set_stream(self.main)
        self.trav = api.TRAVERSE_EXACT if args.traverse == "exact" else api.TRAVERSE_WIDE

    def fence(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, x: int) -> int:
        t = self.torch.tensor([x], dtype=self.torch.int64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return int(t.item())

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


# ------------------------------------------------------------------------------------------------
# GPU arm: one workload / frame size / assembly mode
# ------------------------------------------------------------------------------------------------

class FrameRig:
    """Device-resident buffers and step() of one frame workload.  Modes:
       single  one GPU renders the whole frame into its own buffer
       peer    N > 1: every rank stores its tiles straight into rank 0's frame over NVLink peer memory (CUDA IPC)
               from inside the trace kernel; a 4-byte all-reduce per step is the frame fence.  The progressive path
               tracer keeps its blend state in a local compact tile buffer and MIRRORS finished pixels into that frame
       nccl    N > 1: compact tile buffers + NCCL all-gather + de-interleave on a side stream, double-buffered
               (--gather nccl; kept as the comparison)"""

    def __init__(self, ctx: Ctx, wl: dict, dev, w: int, h: int, mode: str):
        torch, api = ctx.torch, ctx.api
        self.ctx, self.wl, self.dev, self.w, self.h, self.mode = ctx, wl, dev, w, h, mode
        self.mats, self.lights = ctx.S.default_material(wl["mat"]), ctx.S.point_light(wl["eye"])
        self.d_mats = torch.from_numpy(self.mats.view(np.float32).copy()).cuda()
        self.d_lights = torch.from_numpy(self.lights.view(np.float32).copy()).cuda()
        self.d_rays = torch.zeros(1, dtype=torch.int64, device="cuda")
        self.per = max(api.tiles_of_rank(w, h, r, ctx.world) for r in range(ctx.world))
        self.k = 0
        self.frame_buf = None
        self.local = None
        if mode == "single":
            self.image = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
        elif mode == "peer":
            handle = [None]
            if ctx.rank == 0:
                self.frame_buf = api.DeviceBuffer(w * h * 16)
                handle[0] = self.frame_buf.export()
            ctx.dist.broadcast_object_list(handle, src=0)
            if ctx.rank != 0:
                self.frame_buf = api.DeviceBuffer.open(handle[0], w * h * 16)
            self.done_flag = torch.zeros(1, dtype=torch.float32, device="cuda")
            # progressive blending re-reads the rank's own accumulated frame: it stays in a local compact tile buffer, and
            # every finished pixel is mirrored in place into rank 0's frame (vsb_frame_desc.mirror_color)
            self.local = torch.zeros((self.per, api.TILE * api.TILE, 4), dtype=torch.float32, device="cuda") if wl["kernel"] == 3 else None
        else:
            T = api.TILE * api.TILE
            self.tile_buf = [torch.zeros((self.per, T, 4), dtype=torch.float32, device="cuda") for _ in range(2)]
            self.gathered = torch.zeros((ctx.world, self.per, T, 4), dtype=torch.float32, device="cuda")
            self.image = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
            self.comm = torch.cuda.Stream()
            self.rendered = [torch.cuda.Event() for _ in range(2)]
            self.gathered_done = [torch.cuda.Event() for _ in range(2)]

    def params(self, k: int):
        return make_frame_params(self.ctx.S, self.wl, self.w, self.h, frame_num=k + 1 if self.wl["kernel"] == 3 else 1)

    def step(self):
        ctx, api = self.ctx, self.ctx.api
        k = self.k
        self.k += 1
        fp = self.params(k)
        st = ctx.main.cuda_stream
        if self.mode == "single":
            api.frame(self.dev, fp, None, self.d_mats, self.d_lights, self.image, traverse=ctx.trav, ray_count=self.d_rays, stream=st)
        elif self.mode == "peer":
            if self.local is not None:
                api.frame(self.dev, fp, None, self.d_mats, self.d_lights, self.local, traverse=ctx.trav, tile_rank=ctx.rank,
                          tile_nranks=ctx.world, ray_count=self.d_rays, stream=st, mirror=self.frame_buf)
            else:
                api.frame(self.dev, fp, None, self.d_mats, self.d_lights, self.frame_buf, traverse=ctx.trav, tile_rank=ctx.rank,
                          tile_nranks=ctx.world, ray_count=self.d_rays, stream=st, tiles_in_place=True)
            ctx.dist.all_reduce(self.done_flag)            # frame k is complete on rank 0 once this returns on its stream
        else:
            b = k & 1
            ctx.main.wait_event(self.gathered_done[b])     # buffer b was last read by the gather of frame k-2
            api.frame(self.dev, fp, None, self.d_mats, self.d_lights, self.tile_buf[b], traverse=ctx.trav, tile_rank=ctx.rank,
                      tile_nranks=ctx.world, ray_count=self.d_rays, stream=st)
            self.rendered[b].record(ctx.main)
            self.gather(b)

    def gather(self, b: int):
        ctx = self.ctx
        with ctx.torch.cuda.stream(self.comm):
            self.comm.wait_event(self.rendered[b])
            ctx.dist.all_gather_into_tensor(self.gathered.view(-1), self.tile_buf[b].view(-1))
            ctx.api.deinterleave_tiles(self.gathered, self.w, self.h, ctx.world, self.per, self.image, stream=self.comm.cuda_stream)
            self.gathered_done[b].record(self.comm)

    def end_of_region(self):
        if self.mode == "nccl":
            self.ctx.main.wait_stream(self.comm)           # the last gather belongs to the timed region

    def fence_only(self):
        """The exchange step of one frame WITHOUT the trace: what the assembly costs when nothing hides it."""
        if self.mode == "peer":
            self.ctx.dist.all_reduce(self.done_flag)
        elif self.mode == "nccl":
            self.rendered[0].record(self.ctx.main)
            self.gather(0)
            self.ctx.main.wait_stream(self.comm)

    def assembled(self):
        """Rank 0: the assembled frame as a device tensor."""
        torch, api = self.ctx.torch, self.ctx.api
        if self.mode != "peer":
            return self.image
        got = torch.empty((self.h, self.w, 4), dtype=torch.float32, device="cuda")
        api._check(api.N.cuda_lib().vsb_device_copy(got.data_ptr(), self.frame_buf.data_ptr(), self.w * self.h * 16, 0, 0), "vsb_device_copy")
        torch.cuda.synchronize()
        return got

    def close(self):
        if self.frame_buf is not None:
            self.ctx.fence()
            self.frame_buf.close()


def time_rig(ctx: Ctx, rig: FrameRig, steps: int, warmup: int, clocks: bool):
    """W untimed steps, then exactly K steps between barrier + synchronize on both sides, CUDA events on the launching
    stream, max over ranks.  Then K more steps for the dominant kernel's own launch time (events recorded by the
    library around each launch), and the exchange step alone."""
    torch, api = ctx.torch, ctx.api
    for _ in range(warmup):
        rig.step()
    rig.end_of_region()
    ctx.fence()
    rig.d_rays.zero_()
    launches0 = api.launch_count()
    sampler = ClockSampler(ctx.local_rank) if clocks else None
    if sampler:
        sampler.start()
        time.sleep(0.25)
        ctx.fence()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.time()
    e0.record(ctx.main)
    for _ in range(steps):
        rig.step()
    rig.end_of_region()
    e1.record(ctx.main)
    ctx.fence()
    tw1 = time.time()
    ms_total = ctx.max_over_ranks(e0.elapsed_time(e1))
    clk = sampler.stop(tw0, tw1) if sampler else None
    launches = ctx.sum_over_ranks(api.launch_count() - launches0)
    rays_total = ctx.sum_over_ranks(int(rig.d_rays.item()))
    kms = []
    rig.d_rays.zero_()
    for _ in range(steps):
        rig.step()
        kms.append(api.last_kernel_ms())
    rig.end_of_region()
    ctx.fence()
    rays_rank = int(rig.d_rays.item()) / steps
    fence_ms = None
    if ctx.world > 1:
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        rig.fence_only()
        ctx.fence()
        f0.record(ctx.main)
        for _ in range(steps):
            rig.fence_only()
        f1.record(ctx.main)
        ctx.fence()
        fence_ms = ctx.max_over_ranks(f0.elapsed_time(f1)) / steps
    return {"value": rays_total / (ms_total * 1e-3) / 1e6, "ms_per_step": ms_total / steps, "kernel_ms": float(np.mean(kms)),
            "rays_per_step": rays_total // steps, "rays_per_step_rank": rays_rank, "launches": launches, "clocks": clk,
            "exchange_ms": fence_ms}


class SharedHostFrame:
    """ONE host frame for all ranks of the node: a file in /dev/shm mapped by every process and page-locked + mapped
    into each process's CUDA address space, so that every rank's trace kernel stores its finished tiles straight into
    the frame the caller (rank 0) reads: no per-rank buffers, no gather, no CPU pass."""

    def __init__(self, ctx: Ctx, w: int, h: int):
        self.ctx = ctx
        name = [None]
        if ctx.rank == 0:
            name[0] = f"/dev/shm/vsb_frame_{os.getpid()}_{w}x{h}"
            with open(name[0], "wb") as f:
                f.truncate(w * h * 16)
        ctx.dist.broadcast_object_list(name, src=0)
        self.path = name[0]
        self.array = np.memmap(self.path, dtype=np.float32, mode="r+", shape=(h, w, 4))
        self.ptr, self.nbytes = self.array.ctypes.data, w * h * 16
        rt = ctx.torch.cuda.cudart()
        err = rt.cudaHostRegister(self.ptr, self.nbytes, 1 | 2)          # portable | mapped
        if int(err) != 0:
            raise RuntimeError(f"cudaHostRegister failed: {err}")
        ctx.fence()

    def close(self):
        self.ctx.fence()
        self.ctx.torch.cuda.cudart().cudaHostUnregister(self.ptr)
        del self.array
        self.ctx.fence()
        if self.ctx.rank == 0:
            try:
                os.unlink(self.path)
            except OSError:
                pass


def e2e_frames(ctx: Ctx, wl: dict, dev, w: int, h: int, steps: int, check_against=None):
    """The same metric through the host-buffer call: frame() into a pinned host render target, parameter H2D and
    framebuffer D2H inside the timed region (wall clock, max over ranks).  N > 1: ONE assembled frame in a host buffer
    shared by all ranks (SharedHostFrame); the progressive path tracer keeps per-rank tile buffers, because its blend
    re-reads the rank's own previous frame."""
    torch, api, S = ctx.torch, ctx.api, ctx.S
    mats, lights = S.default_material(wl["mat"]), S.point_light(wl["eye"])
    world, rank = ctx.world, ctx.rank
    shared = world > 1 and wl["kernel"] != 3
    per = max(api.tiles_of_rank(w, h, r, world) for r in range(world))
    frame = None
    if shared:
        frame = SharedHostFrame(ctx, w, h)
        host_np, n_px = frame.array, w * h // world
    else:
        n_px = per * api.TILE * api.TILE if world > 1 else w * h
        host_np = torch.zeros((n_px, 4), dtype=torch.float32).pin_memory().numpy()
    h_rays = np.zeros(1, dtype=np.uint64)

    def one(k):
        fp = make_frame_params(S, wl, w, h, frame_num=k + 1 if wl["kernel"] == 3 else 1)
        api.frame(dev, fp, None, mats, lights, host_np, traverse=ctx.trav, tile_rank=rank if world > 1 else 0,
                  tile_nranks=world if world > 1 else 0, ray_count=h_rays, tiles_in_place=shared)
        return int(h_rays[0])

    one(0)
    ctx.fence()
    t0 = time.perf_counter()
    rays = 0
    for k in range(steps):
        rays += one(k)
    api.sync()
    ctx.fence()                                       # N > 1: the frame is complete when every rank's call has returned
    secs = ctx.max_over_ranks(time.perf_counter() - t0)
    trace_ms = api.last_kernel_ms()
    rays = ctx.sum_over_ranks(rays)
    verified = None
    if shared and check_against is not None and rank == 0:
        verified = bool(np.array_equal(np.asarray(host_np), check_against.cpu().numpy()))
    h2d = (mats.nbytes + lights.nbytes + 256) * world
    if wl["kernel"] == 3:
        h2d += n_px * 16 * world                       # progressive state lives with the caller
    d2h = (n_px * 16 + 8) * world
    out = {"value": rays / secs / 1e6, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
           "steps": steps, "ms_per_step": secs * 1e3 / steps, "trace_ms": trace_ms,
           "d2h_gb_s_per_rank": n_px * 16 / (secs / steps) / 1e9,
           "target": ("one assembled frame in a /dev/shm buffer page-locked and mapped by all ranks; kernels store tiles in place over PCIe"
                      if shared else ("per-rank compact tile buffers (progressive blend state), pinned" if world > 1 else "pinned host frame, banded D2H overlapping the trace")),
           "assembled_frame_verified": verified}
    if frame is not None:
        frame.close()
    return out


# ------------------------------------------------------------------------------------------------
# GPU arm: ray lists (cfg5)
# ------------------------------------------------------------------------------------------------

def time_raylist(ctx: Ctx, dev, any_hit: bool, n_batch: int, steps: int, warmup: int, clocks: bool, e2e: bool):
    """Incoherent random-ray stress: rays uniform in the unit cube with uniform directions vs the triangle soup; a step
    = one vsb_closest_hit / vsb_any_hit pass over one batch of n_batch rays (4 rotating batches).  N > 1: every rank
    gets its own batches (contiguous chunks of the global list, different seeds), no collective."""
    torch, api, S = ctx.torch, ctx.api, ctx.S
    n_bufs = 4
    rays_h = [S.random_rays(n_batch, 600 + 16 * ctx.rank + i) for i in range(n_bufs)]
    d_rays = [torch.from_numpy(r).cuda() for r in rays_h]
    d_hits = [torch.empty((n_batch, 8), dtype=torch.int32, device="cuda") for _ in range(n_bufs)]
    fn = api.any_hit if any_hit else api.closest_hit
    k = [0]

    def step():
        i = k[0] % n_bufs
        k[0] += 1
        fn(dev, d_rays[i], hits=d_hits[i], traverse=ctx.trav, stream=ctx.main.cuda_stream)

    for _ in range(warmup):
        step()
    ctx.fence()
    launches0 = api.launch_count()
    sampler = ClockSampler(ctx.local_rank) if clocks else None
    if sampler:
        sampler.start()
        time.sleep(0.25)
        ctx.fence()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.time()
    e0.record(ctx.main)
    for _ in range(steps):
        step()
    e1.record(ctx.main)
    ctx.fence()
    tw1 = time.time()
    ms_total = ctx.max_over_ranks(e0.elapsed_time(e1))
    clk = sampler.stop(tw0, tw1) if sampler else None
    launches = ctx.sum_over_ranks(api.launch_count() - launches0)
    kms = []
    for _ in range(steps):
        step()
        kms.append(api.last_kernel_ms())
    ctx.fence()
    out = {"value": n_batch * steps * ctx.world / (ms_total * 1e-3) / 1e6, "ms_per_step": ms_total / steps,
           "kernel_ms": float(np.mean(kms)), "launches": launches, "clocks": clk,
           "hit_fraction": float((d_hits[0][:, 0] != 0).float().mean().item()), "sample_rays": rays_h[0][:200_000].copy(),
           "e2e": None}
    if e2e:
        # host ray list in, host hit records out (pinned), H2D + D2H inside the call
        h_rays = torch.from_numpy(rays_h[0].copy()).pin_memory().numpy()
        h_hits = torch.zeros((n_batch, 8), dtype=torch.int32).pin_memory().numpy().view(api.N.HIT_DTYPE).reshape(-1)
        fn(dev, h_rays, hits=h_hits, traverse=ctx.trav)
        ctx.fence()
        e2e_steps = max(1, min(steps, 5))
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            fn(dev, h_rays, hits=h_hits, traverse=ctx.trav)
        ctx.fence()
        secs = ctx.max_over_ranks(time.perf_counter() - t0)
        out["e2e"] = {"value": n_batch * e2e_steps * ctx.world / secs / 1e6, "unit": "Mrays/s",
                      "h2d_bytes_per_step": n_batch * 32 * ctx.world, "d2h_bytes_per_step": n_batch * 32 * ctx.world,
                      "steps": e2e_steps, "rays_per_step": n_batch, "ms_per_step": secs * 1e3 / e2e_steps,
                      "trace_ms": api.last_kernel_ms()}
    del d_rays, d_hits
    return out


def raylist_roofline(O, osc, sample, any_hit: bool, n_batch: int, kernel_ms: float, traffic_key):
    """B_ray from the oracle's instrumented traversal of the scene's BVH2 on 200 k rays of the list."""
    peak, peak_src = measured_peak_gbs()
    c = O.Counters()
    O.oracle_query(osc, sample, any_hit, counters=c)
    n_inner, n_prim = c.n_inner / c.n_rays, c.n_prim / c.n_rays
    b_ray = 64.0 * n_inner + 68.0 * n_prim + 64.0
    achieved = b_ray * n_batch / (kernel_ms * 1e-3) / 1e9
    return {"bound": "issue", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": measured_traffic(traffic_key) if traffic_key else None,
            "issue": measured_issue(traffic_key) if traffic_key else None,
            "kernel": "trace_kernel_q<RayListProgram>", "kernel_ms": kernel_ms, "bytes_per_ray": b_ray,
            "n_inner_per_ray": n_inner, "n_prim_per_ray": n_prim, "peak_source": peak_src,
            "roofline_mrays_s": peak * 1e9 / b_ray / 1e6}


def frame_roofline(O, S, osc, wl, rig_out: dict, traffic_key):
    peak, peak_src = measured_peak_gbs()
    mats, lights = S.default_material(wl["mat"]), S.point_light(wl["eye"])
    b_ray, visit = algorithmic_bytes_per_ray(O, S, osc, wl, mats, lights)
    achieved = b_ray * rig_out["rays_per_step_rank"] / (rig_out["kernel_ms"] * 1e-3) / 1e9
    return {"bound": "issue", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
            "traffic": measured_traffic(traffic_key) if traffic_key else None,
            "issue": measured_issue(traffic_key) if traffic_key else None,
            "kernel": "trace_kernel_q<FrameProgram>", "kernel_ms": rig_out["kernel_ms"], "bytes_per_ray": b_ray,
            "n_inner_per_ray": visit["n_inner"], "n_prim_per_ray": visit["n_prim"], "peak_source": peak_src,
            "roofline_mrays_s": peak * 1e9 / b_ray / 1e6}


# ------------------------------------------------------------------------------------------------
# GPU arm: the runs
# ------------------------------------------------------------------------------------------------

def extra_legs(ctx: Ctx, O, args, cfg2_prims):
    """N = 1, default run: the incoherent workloads in the same record, measured like the headline (short legs), on the
    reference's binned-SAH tree (BASELINE configs / north_star: "re-encoded on the GPU from the reference's binned-SAH
    tree"; built by --builder); and the second GPU builder (Morton + PLOC) beside it: build time and traced rate."""
    S, api = ctx.S, ctx.api
    out = {}
    steps, warmup = 5, 3
    n_batch = 1 << 24

    def raylist(dev, osc, any_hit, info, key):
        r = time_raylist(ctx, dev, any_hit, n_batch, steps, warmup, clocks=False, e2e=False)
        return {"value": r["value"], "unit": "Mrays/s", "kernel_ms": r["kernel_ms"], "ms_per_step": r["ms_per_step"],
                "steps": steps, "warmup": warmup, "hit_fraction": r["hit_fraction"], "builder": info,
                "config": raylist_config(osc.n_prims, any_hit, n_batch, 1.0),
                "roofline": raylist_roofline(O, osc, r["sample_rays"], any_hit, n_batch, r["kernel_ms"], key)}

    sc, dev, info = prepare_scene(api, S, "cfg5_soup", 1.0, args.builder)
    osc = oracle_scene(S, sc, dev)
    out["cfg5_closest"] = raylist(dev, osc, False, info, "cfg5_closest")
    out["cfg5_any"] = raylist(dev, osc, True, info, "cfg5_any")
    dev.close()
    del osc
    ploc = {}
    t0 = time.perf_counter()
    api.tune("build_algorithm", 0)
    dev = api.Scene.build(sc.prim_kind, sc.prims, None)
    api.sync()
    ms = api.last_build_ms()
    binfo = {"kind": GPU_BUILDERS["gpu_ploc"][1], "build_ms": round(ms["total_ms"], 2), "iterations": ms["iterations"],
             "upload_and_build_s": round(time.perf_counter() - t0, 3)}
    r = time_raylist(ctx, dev, False, n_batch, steps, warmup, clocks=False, e2e=False)
    ploc["cfg5_closest"] = {"value": r["value"], "unit": "Mrays/s", "kernel_ms": r["kernel_ms"], "builder": binfo,
                            "vs_headline_tree": r["value"] / out["cfg5_closest"]["value"]}
    dev.close()
    del sc

    wl = WORKLOADS["cfg3"]
    sc, dev, info = prepare_scene(api, S, wl["scene"], 1.0, args.builder)
    rig = FrameRig(ctx, wl, dev, wl["w"], wl["h"], "single")
    r = time_rig(ctx, rig, steps, warmup, clocks=False)
    out["cfg3"] = {"value": r["value"], "unit": "Mrays/s", "kernel_ms": r["kernel_ms"], "ms_per_step": r["ms_per_step"],
                   "steps": steps, "warmup": warmup, "rays_per_step": r["rays_per_step"], "builder": info,
                   "config": workload_config(wl, wl["w"], wl["h"], 1.0, sc.n_prims),
                   "roofline": frame_roofline(O, S, oracle_scene(S, sc, dev), wl, r, "cfg3")}
    rig.close()
    dev.close()
    del sc

    # the headline frame on the PLOC tree: primitives -> build -> first frame, then the same timing
    wl = WORKLOADS["cfg2"]
    t0 = time.perf_counter()
    api.tune("build_algorithm", 0)
    dev = api.Scene.build(0, cfg2_prims, S.face_normals(cfg2_prims))
    api.sync()
    ms = api.last_build_ms()
    rig = FrameRig(ctx, wl, dev, wl["w"], wl["h"], "single")
    rig.step()
    ctx.torch.cuda.synchronize()
    first = time.perf_counter() - t0
    r = time_rig(ctx, rig, steps, warmup, clocks=False)
    ploc["cfg2"] = {"value": r["value"], "unit": "Mrays/s", "kernel_ms": r["kernel_ms"],
                    "builder": {"kind": GPU_BUILDERS["gpu_ploc"][1], "build_ms": round(ms["total_ms"], 2),
                                "iterations": ms["iterations"], "primitives_to_first_frame_s": round(first, 3)}}
    rig.close()
    dev.close()
    api.tune("build_algorithm", 1)
    out["ploc_tree"] = ploc

    # the same frame by the REFERENCE'S OWN CUDA path on this GPU: cuda_sched<ray>::frame() + whitted::kernel over a
    # cuda_index_bvh, unmodified headers compiled for sm_100a (oracle/refcuda; none of this repository's kernels)
    if O.ref_cuda_available():
        try:
            wl = WORKLOADS["cfg2"]
            fp = make_frame_params(S, wl, wl["w"], wl["h"])
            host = S.HostScene(0, cfg2_prims, None, None, None, "cfg2")
            _, ms, bms = O.ref_cuda_whitted(host, fp, wl["eye"], reps=steps, warmup=warmup, want_image=False)
            out["naive_gpu"] = {"impl": "the reference's cuda_sched<basic_ray<float>> + whitted::kernel (one thread per pixel, BVH2, 16x16 blocks)",
                                "ms_per_step": ms, "frame": [wl["w"], wl["h"]], "steps": steps, "warmup": warmup,
                                "host_build_and_upload_ms": bms}
        except Exception as e:                         # a baseline, never a reason to lose the run's line
            out["naive_gpu"] = {"unavailable": str(e)[:200]}
    else:
        out["naive_gpu"] = {"unavailable": "oracle/_ref/libvsnray_ref_cuda.so was not built"}
    return out


def strong_legs(ctx: Ctx, args, cfg2_scene=None):
    """N > 1: the FIXED 4096^2 frame partitioned over the N GPUs -- cfg2, and cfg4 (configs[3]: the progressive path
    tracer), both assembled by peer stores into rank 0's frame -- against the same frame on one GPU alone, measured
    in the same run (every rank renders it alone at the same time; rank-local work, max over ranks)."""
    S, api = ctx.S, ctx.api
    out = {}
    steps, warmup = max(5, min(args.steps, 10)), 3
    mode = "nccl" if args.gather == "nccl" else "peer"
    for name in ("cfg2", "cfg4"):
        wl = WORKLOADS[name]
        if name == "cfg2" and cfg2_scene is not None:
            sc, dev, info, own = cfg2_scene[0], cfg2_scene[1], cfg2_scene[2], False
        else:
            sc, dev, info = prepare_scene(api, S, wl["scene"], 1.0, args.builder)
            own = True
        w, h = wl["w"], wl["h"]
        alone = FrameRig(ctx, wl, dev, w, h, "single")
        r1 = time_rig(ctx, alone, steps, warmup, clocks=False)
        v1 = r1["value"] / ctx.world                              # all ranks ran the same frame alone: per-GPU rate
        alone.close(); del alone
        rig = FrameRig(ctx, wl, dev, w, h, mode)
        rn = time_rig(ctx, rig, steps, warmup, clocks=False)
        rig.close(); del rig
        # the assembled frame equals the frame one GPU renders alone: fresh buffers on both sides, 3 frames each (the
        # progressive path tracer accumulates frame_num = 1..3)
        a2, r2 = FrameRig(ctx, wl, dev, w, h, "single"), FrameRig(ctx, wl, dev, w, h, mode)
        for _ in range(3):
            a2.step(); r2.step()
        r2.end_of_region()
        ctx.fence()
        verified = None
        if ctx.rank == 0:
            verified = bool(ctx.torch.equal(r2.assembled(), a2.image))
        ctx.fence()
        a2.close(); r2.close(); del a2, r2
        out[name] = {"frame": [w, h], "assembly": mode, "value": rn["value"], "unit": "Mrays/s", "ms_per_step": rn["ms_per_step"],
                     "trace_kernel_ms": rn["kernel_ms"], "exchange_alone_ms": rn["exchange_ms"],
                     "one_gpu_value": v1, "one_gpu_ms_per_step": r1["ms_per_step"], "efficiency": rn["value"] / (ctx.world * v1),
                     "steps": steps, "warmup": warmup, "assembled_frame_verified": verified,
                     "triangles": sc.n_prims, "builder": info}
        if own:
            dev.close()
    return out


def run_gpu(args):
    ctx = Ctx(args)
    S, api = ctx.S, ctx.api
    wl = WORKLOADS[args.workload]
    w, h = frame_shape(wl, ctx.world)
    t_start = time.perf_counter()
    sc, dev, builder = prepare_scene(api, S, wl["scene"], args.scale, args.builder)
    stats = dev.stats()
    mode = "single" if ctx.world == 1 else ("nccl" if args.gather == "nccl" else "peer")
    rig = FrameRig(ctx, wl, dev, w, h, mode)
    rig.step()                                         # generate -> build -> first frame (one more untimed warm-up step)
    rig.end_of_region()
    ctx.torch.cuda.synchronize()
    builder["time_to_first_frame_s"] = round(time.perf_counter() - t_start, 3)
    r = time_rig(ctx, rig, args.steps, args.warmup, clocks=True)

    # N > 1: the assembled frame on rank 0 equals the frame one GPU renders alone
    verified, alone_img = None, None
    if ctx.world > 1 and wl["kernel"] != 3:
        ctx.fence()
        if ctx.rank == 0:
            alone_img = ctx.torch.zeros((h, w, 4), dtype=ctx.torch.float32, device="cuda")
            api.frame(dev, rig.params(0), None, rig.d_mats, rig.d_lights, alone_img, traverse=ctx.trav, stream=ctx.main.cuda_stream)
            ctx.torch.cuda.synchronize()
            verified = bool(ctx.torch.equal(rig.assembled(), alone_img))
        ctx.fence()
    rig.close()
    e2e = e2e_frames(ctx, wl, dev, w, h, max(1, min(args.steps, 5)), check_against=alone_img)
    del alone_img

    strong = None
    if ctx.world > 1 and not args.no_extra and args.workload == "cfg2" and args.scale == 1.0:
        strong = strong_legs(ctx, args, cfg2_scene=(sc, dev, builder))

    line = None
    full = ctx.world == 1 and args.scale == 1.0
    if ctx.rank == 0:
        from oracle import bindings as O              # checker / CPU baseline only (never on the product path)
        osc = oracle_scene(S, sc, dev)
        roofline = frame_roofline(O, S, osc, wl, r, args.workload if full else None)
        cpu = None
        if ctx.world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            secs, kind = cpu_frame_seconds(O, S, osc if sc.nodes is None else sc, wl, rig.mats, rig.lights, w, h, reps=3, warmup=1, cores=cores)
            rays = r["rays_per_step"]
            cpu = {"value": rays / secs / 1e6, "unit": "Mrays/s", "cores": cores, "kind": kind,
                   "sample": f"the whole {w}x{h} frame ({rays} rays), best of 3 after 1 warm-up, "
                             + ("tiled_sched<simd::ray8> -O3 -mavx2 -mfma" if kind == "reference" else "scalar C port")
                             + (", on the GPU-built tree" if sc.nodes is None else "")}
        cfg = workload_config(wl, w, h, args.scale, sc.n_prims)
        cfg.update({"rays_per_step": r["rays_per_step"], "traverse": args.traverse,
                    "partition": f"interleaved 64x64 tiles over {ctx.world} GPU(s), scene replicated"
                                 + {"single": "", "peer": ", tiles stored in place into rank 0's frame over NVLink peer memory (CUDA IPC) + 1 tiny all-reduce per frame",
                                    "nccl": ", NCCL all-gather + de-interleave per frame on a side stream"}[mode],
                    "assembled_frame_verified": verified, "exchange_alone_ms": r["exchange_ms"],
                    "wide_nodes": stats["wide_nodes"], "bvh2_nodes": stats["bvh2_nodes"], "builder": builder})
        line = {"metric": "Mrays/s", "value": r["value"], "unit": "Mrays/s", "n_gpus": ctx.world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "roofline": roofline,
                "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(r["launches"]), "clocks": r["clocks"]}
        if strong is not None:
            line["strong"] = strong
    dev.close()
    if ctx.world == 1 and not args.no_extra and args.workload == "cfg2" and args.scale == 1.0:
        from oracle import bindings as O
        line["extra"] = extra_legs(ctx, O, args, sc.prims)
        line["extra"]["ploc_tree"]["cfg2"]["vs_headline_tree"] = line["extra"]["ploc_tree"]["cfg2"]["value"] / line["value"]
        ng = line["extra"].get("naive_gpu") or {}
        if "ms_per_step" in ng:
            # the same frame = the same rays (up to tie pixels): this run's kernel counter gives the rate
            ng["value"] = line["config"]["rays_per_step"] / (ng["ms_per_step"] * 1e-3) / 1e6
            ng["unit"] = "Mrays/s"
            ng["this_repo_over_naive_gpu"] = line["value"] / ng["value"]
    if ctx.rank == 0:
        print(json.dumps(line), flush=True)
    ctx.close()
    return 0


def run_gpu_raylist(args):
    """--workload cfg5 as the headline of a run (closest-hit, or any-hit with --any)."""
    ctx = Ctx(args)
    S, api = ctx.S, ctx.api
    sc, dev, builder = prepare_scene(api, S, "cfg5_soup", args.scale, args.builder)
    stats = dev.stats()
    r = time_raylist(ctx, dev, bool(args.any), args.batch, args.steps, args.warmup, clocks=True, e2e=True)
    if ctx.rank == 0:
        from oracle import bindings as O              # checker / CPU baseline only
        osc = oracle_scene(S, sc, dev)
        full = ctx.world == 1 and args.scale == 1.0 and args.batch == 1 << 24
        key = ("cfg5_any" if args.any else "cfg5_closest") if full else None
        roofline = raylist_roofline(O, osc, r["sample_rays"], bool(args.any), args.batch, r["kernel_ms"], key)
        cpu = None
        if ctx.world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            nref = min(args.batch, 2_000_000)
            rays = S.random_rays(nref, 600)
            if O.ref_available(fast=True):
                secs, _ = O.ref_query_timed(osc, rays, bool(args.any), nthreads=cores, reps=2, warmup=1)
                cpu = {"value": nref / secs / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                       "sample": f"{nref} rays of the same list, scalar closest_hit()/any_hit() of the unmodified headers on {cores} std::threads, best of 2"}
            else:
                t0 = time.perf_counter()
                O.oracle_query(osc, rays[:200_000], bool(args.any))
                cpu = {"value": 200_000 / (time.perf_counter() - t0) / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
                       "sample": "200000 rays of the same list, scalar C port, pthreads"}
        cfg = raylist_config(sc.n_prims, bool(args.any), args.batch, args.scale)
        cfg.update({"traverse": args.traverse, "builder": builder, "wide_nodes": stats["wide_nodes"], "bvh2_nodes": stats["bvh2_nodes"],
                    "bvh2_depth": stats["bvh2_depth"], "hit_fraction": r["hit_fraction"]})
        line = {"metric": "Mrays/s", "value": r["value"], "unit": "Mrays/s", "n_gpus": ctx.world, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg, "roofline": roofline,
                "cpu_baseline": cpu, "e2e": r["e2e"], "gpu_launches": int(r["launches"]), "clocks": r["clocks"]}
        print(json.dumps(line), flush=True)
    dev.close()
    ctx.close()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=sorted(WORKLOADS) + ["cfg5"])
    ap.add_argument("--any", action="store_true", help="cfg5: any-hit instead of closest-hit")
    ap.add_argument("--batch", type=int, default=1 << 24, help="cfg5: rays per batch")
    ap.add_argument("--traverse", default="wide", choices=["wide", "exact"])
    ap.add_argument("--builder", default="gpu", choices=["gpu", "reference", "gpu_ploc"],
                    help="who builds the BVH: gpu = the reference's binned-SAH tree built on the GPU (same nodes and indices as the host "
                         "builder, ~5 ms instead of 0.5 s at 1 M triangles); reference = the host builder; gpu_ploc = Morton + PLOC "
                         "(another tree: about the same build time, traces 5-14 %% slower)")
    ap.add_argument("--scale", type=float, default=1.0, help="scene size factor (1.0 = the BASELINE config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra (N = 1) / strong (N > 1) records of the default run")
    ap.add_argument("--gather", default="peer", choices=["peer", "nccl"], help="N > 1: how the frame is assembled")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "cfg5":
        return run_gpu_raylist(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
