#!/usr/bin/env python
"""bench.py -- Mrays/s of the ray-query hot path on B200 (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg2|cfg1|cfg3|cfg4|cfg5] [--impl reference]

A "step" is one frame() call over the workload's synthetic scene (one pass of ray generation +
closest_hit/any_hit traversal + shading + store).  Default workload = BASELINE.json configs[1]:
4096x4096 primary rays + Whitted shading vs the 1,002,528-triangle displaced grid.

  value     rays traced / device time of the timed region, scene + parameter arrays + framebuffer
            resident in HBM (rays are counted by the kernel: primary + shadow + bounce rays)
  e2e       the same metric through the host-buffer call a user of the reference makes
            (frame() into a host render target): parameter H2D + framebuffer D2H inside the region
  roofline  node-fetch memory roofline of SURVEY.md section 8d: B_ray = 64 N_inner + 68 N_prim + 64
            bytes per ray on the REFERENCE data structure (N_* counted by the oracle's instrumented
            traversal on a down-sampled frame of the same view), against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline  the reference's own tiled_sched<simd::ray8> (oracle/_ref, unmodified headers) on the
            host cores, bounded sample

N > 1 (torchrun): the frame is partitioned into interleaved 64x64 tiles (tile k -> rank k % N), the
scene is replicated, every step ends with one NCCL all-gather of the tile buffers + de-interleave.
Weak scaling: the same view at sqrt(N) x the resolution per axis (4096^2 pixels per GPU).
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


def measured_traffic(workload: str):
    """DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/), or None."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return int(json.load(f)[workload]["bytes"])
    except Exception:
        return None


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


# ------------------------------------------------------------------------------------------------
# CPU side: algorithmic bytes per ray (oracle counters), cpu_baseline, and the --impl reference arm
# ------------------------------------------------------------------------------------------------

def algorithmic_bytes_per_ray(O, S, sc, wl, mats, lights, sample_wh=(512, 512)):
    """B_ray = 64 N_inner + 68 N_prim + 64 from the oracle's instrumented scalar
    traversal of the reference BVH2 on a down-sampled frame of the same view."""
    fp = make_frame_params(S, wl, *sample_wh)
    c = O.Counters()
    O.oracle_frame(sc, fp, mats, lights, counters=c)
    n = max(int(c.n_rays), 1)
    n_inner, n_prim = c.n_inner / n, c.n_prim / n
    prim_bytes = 68.0 if sc.prim_kind == 0 else 52.0
    return 64.0 * n_inner + prim_bytes * n_prim + 64.0, dict(n_inner=n_inner, n_prim=n_prim, rays_per_pixel=n / (sample_wh[0] * sample_wh[1]))


def cpu_frame_rate(O, S, sc, wl, mats, lights, wh, reps, warmup):
    """Mrays/s of the reference's tiled_sched<simd::ray8>::frame() (oracle/_ref, unmodified headers,
    -O3 -mavx2 -mfma) on all host cores; the oracle port (scalar C, pthreads) when _ref is absent.
    Returns (mrays_per_s, seconds_per_frame, rays_per_frame, kind, cores)."""
    cores = os.cpu_count() or 1
    fp = make_frame_params(S, wl, *wh)
    c = O.Counters()
    t0 = time.perf_counter()
    O.oracle_frame(sc, fp, mats, lights, counters=c)       # ray count of this sample (and the port's time)
    port_s = time.perf_counter() - t0
    rays = int(c.n_rays)
    if O.ref_available(fast=True):
        _, secs, _ = O.ref_frame(sc, fp, mats, lights, mode=1, reps=reps, warmup=warmup, fast=True, nthreads=cores)
        return rays / secs / 1e6, secs, rays, "reference", cores
    best = port_s
    for _ in range(max(reps - 1, 0)):
        t0 = time.perf_counter()
        O.oracle_frame(sc, fp, mats, lights)
        best = min(best, time.perf_counter() - t0)
    return rays / best / 1e6, best, rays, "port", cores


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import __graft_entry__ as g
    g.build_host_only()
    from oracle import bindings as O
    from visionaray_b200 import scenes as S

    wl = WORKLOADS[args.workload]
    sc = S.make_scene(wl["scene"], args.scale)
    mats, lights = S.default_material(wl["mat"]), S.point_light(wl["eye"])
    wh = (1024, 1024) if wl["w"] >= 1024 else (wl["w"], wl["h"])
    # each "step" is one frame of the bounded sample; best-of is not used here: mean over K steps
    cores = os.cpu_count() or 1
    fp = make_frame_params(S, wl, *wh)
    c = O.Counters()
    O.oracle_frame(sc, fp, mats, lights, counters=c)
    rays = int(c.n_rays)
    kind = "reference" if O.ref_available(fast=True) else "port"
    times = []
    for i in range(args.warmup + args.steps):
        if kind == "reference":
            _, secs, _ = O.ref_frame(sc, fp, mats, lights, mode=1, reps=1, warmup=0, fast=True, nthreads=cores)
        else:
            t0 = time.perf_counter()
            O.oracle_frame(sc, fp, mats, lights)
            secs = time.perf_counter() - t0
        if i >= args.warmup:
            times.append(secs)
    ms = 1e3 * float(np.mean(times))
    value = rays / (ms * 1e-3) / 1e6
    sample = f"{wh[0]}x{wh[1]} frame of the same view ({rays} rays/step), tiled_sched<simd::ray8> AVX2" if kind == "reference" \
        else f"{wh[0]}x{wh[1]} frame of the same view ({rays} rays/step), scalar C port, pthreads"
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": wl["desc"], "scale": args.scale, "kernel": wl["kernel"]},
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------

def make_device_scene(api, sc, args):
    """The device scene of the run.  --builder reference (default): the tree of the reference's
    binned-SAH builder (host, include/visionaray_b200/bvh.h reproduces it byte for byte), uploaded and
    re-encoded.  --builder gpu: the tree is built on the GPU from the primitives (vsb_scene_build)."""
    if args.builder == "gpu":
        dev = api.Scene.build(sc.prim_kind, sc.prims, sc.normals)
        ms = api.last_build_ms()
        return dev, {"kind": "gpu (Morton sort + PLOC + SAH leaf collapse)", "build_ms": round(ms["total_ms"], 2),
                     "iterations": ms["iterations"]}
    return api.Scene(sc), {"kind": "reference binned SAH (host)"}


def run_gpu(args):
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the ray-query path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    from visionaray_b200 import api, scenes as S

    api.init(local_rank)
    wl = WORKLOADS[args.workload]
    w, h = frame_shape(wl, world)
    sc = S.make_scene(wl["scene"], args.scale)
    dev, builder = make_device_scene(api, sc, args)
    stats = dev.stats()
    mats, lights = S.default_material(wl["mat"]), S.point_light(wl["eye"])
    d_mats = torch.from_numpy(mats.view(np.float32).copy()).cuda()
    d_lights = torch.from_numpy(lights.view(np.float32).copy()).cuda()
    d_rays = torch.zeros(1, dtype=torch.int64, device="cuda")
    trav = api.TRAVERSE_EXACT if args.traverse == "exact" else api.TRAVERSE_WIDE

    per = max(api.tiles_of_rank(w, h, r, world) for r in range(world))
    # a real (non-legacy) stream: the C ABI treats a NULL stream as "the library's own stream"
    main = torch.cuda.Stream()
    torch.cuda.set_stream(main)
    # N > 1, how the frame is assembled:
    #   peer  (default) every rank stores its tiles straight into rank 0's frame over NVLink peer
    #         memory (CUDA IPC) from inside the traversal kernel; a tiny all-reduce per step marks
    #         completion.  No gather pass, no de-interleave.
    #   nccl  compact tile buffers + all-gather + de-interleave on a side stream (also used for the
    #         progressive path tracer, whose blend reads the previous frame locally)
    mode = "single" if world == 1 else ("nccl" if (args.gather == "nccl" or wl["kernel"] == 3) else "peer")
    image = None
    if mode == "single":
        image = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    elif mode == "peer":
        handle = [None]
        if rank == 0:
            frame_buf = api.DeviceBuffer(w * h * 16)
            handle[0] = frame_buf.export()
        dist.broadcast_object_list(handle, src=0)
        if rank != 0:
            frame_buf = api.DeviceBuffer.open(handle[0], w * h * 16)
        done_flag = torch.zeros(1, dtype=torch.float32, device="cuda")
    else:
        tile_buf = [torch.zeros((per, api.TILE * api.TILE, 4), dtype=torch.float32, device="cuda") for _ in range(2)]
        gathered = torch.zeros((world, per, api.TILE * api.TILE, 4), dtype=torch.float32, device="cuda")
        image = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
        comm = torch.cuda.Stream()
        rendered = [torch.cuda.Event() for _ in range(2)]
        gathered_done = [torch.cuda.Event() for _ in range(2)]

    step_no = [0]

    def step():
        """One frame: every rank renders its interleaved tiles (see `mode`)."""
        k = step_no[0]
        step_no[0] += 1
        fp = make_frame_params(S, wl, w, h, frame_num=k + 1 if wl["kernel"] == 3 else 1)
        if mode == "single":
            api.frame(dev, fp, None, d_mats, d_lights, image, traverse=trav, ray_count=d_rays, stream=main.cuda_stream)
            return
        if mode == "peer":
            api.frame(dev, fp, None, d_mats, d_lights, frame_buf, traverse=trav, tile_rank=rank, tile_nranks=world,
                      ray_count=d_rays, stream=main.cuda_stream, tiles_in_place=True)
            dist.all_reduce(done_flag)                 # frame k is complete on rank 0 once this returns on its stream
            return
        b = k & 1
        main.wait_event(gathered_done[b])          # buffer b was last read by the gather of frame k-2
        api.frame(dev, fp, None, d_mats, d_lights, tile_buf[b], traverse=trav, tile_rank=rank, tile_nranks=world,
                  ray_count=d_rays, stream=main.cuda_stream)
        rendered[b].record(main)
        with torch.cuda.stream(comm):
            comm.wait_event(rendered[b])
            dist.all_gather_into_tensor(gathered.view(-1), tile_buf[b].view(-1))
            api.deinterleave_tiles(gathered, w, h, world, per, image, stream=comm.cuda_stream)
            gathered_done[b].record(comm)

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    fence()
    d_rays.zero_()
    launches0 = api.launch_count()
    clocks = ClockSampler(local_rank)
    clocks.start()
    time.sleep(0.25)
    fence()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_wall0 = time.time()
    e0.record(main)
    for _ in range(args.steps):
        step()
    if mode == "nccl":
        main.wait_stream(comm)                      # the last gather belongs to the timed region
    e1.record(main)
    fence()
    t_wall1 = time.time()
    ms_total = e0.elapsed_time(e1)
    clk = clocks.stop(t_wall0, t_wall1)
    launches = api.launch_count() - launches0
    rays_total = d_rays.clone()
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays_total, op=dist.ReduceOp.SUM)
        lt = torch.tensor([launches], dtype=torch.int64, device="cuda")
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        launches = int(lt.item())
    ms_total = float(t.item())
    rays_total = int(rays_total.item())
    ms_per_step = ms_total / args.steps
    value = rays_total / (ms_total * 1e-3) / 1e6

    # ---- dominant kernel alone: CUDA events recorded by the library around each frame_kernel launch ----
    kms = []
    d_rays.zero_()
    for _ in range(args.steps):
        step()
        kms.append(api.last_kernel_ms())
    fence()
    kernel_ms = float(np.mean(kms))
    rays_per_step_rank = int(d_rays.item()) / args.steps

    # ---- e2e: host render target (pinned), parameter H2D + framebuffer D2H inside the region ----
    n_px = (per if world > 1 else 0) * api.TILE * api.TILE if world > 1 else w * h
    host_color = torch.zeros((n_px, 4), dtype=torch.float32).pin_memory()
    host_np = host_color.numpy()
    h_rays = np.zeros(1, dtype=np.uint64)
    e2e_steps = max(1, min(args.steps, 5))

    def e2e_step(k):
        fp = make_frame_params(S, wl, w, h, frame_num=k + 1 if wl["kernel"] == 3 else 1)
        api.frame(dev, fp, None, mats, lights, host_np, traverse=trav, tile_rank=rank if world > 1 else 0,
                  tile_nranks=world if world > 1 else 0, ray_count=h_rays)
        return int(h_rays[0])

    e2e_step(0)
    fence()
    t0 = time.perf_counter()
    e2e_rays = 0
    for k in range(e2e_steps):
        e2e_rays += e2e_step(k)
    api.sync()
    fence()
    e2e_s = time.perf_counter() - t0
    e2e_trace_ms = api.last_kernel_ms()           # device time of the last e2e frame's trace launches (all bands)
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    re = torch.tensor([e2e_rays], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        dist.all_reduce(re, op=dist.ReduceOp.SUM)
    e2e_value = int(re.item()) / float(te.item()) / 1e6
    h2d = (mats.nbytes + lights.nbytes + 256) * world            # parameter block + scene table per frame
    if wl["kernel"] == 3:
        h2d += n_px * 16 * world                                   # progressive state lives with the caller
    d2h = (n_px * 16 + 8) * world

    # N > 1: the assembled frame on rank 0 equals the frame one GPU renders alone
    verified = None
    if world > 1 and wl["kernel"] != 3:
        fence()
        if rank == 0:
            fp = make_frame_params(S, wl, w, h)
            alone = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
            api.frame(dev, fp, None, d_mats, d_lights, alone, traverse=trav, stream=main.cuda_stream)
            torch.cuda.synchronize()
            if mode == "peer":
                got = torch.empty_like(alone)
                N_ = api.N.cuda_lib()
                api._check(N_.vsb_device_copy(got.data_ptr(), frame_buf.data_ptr(), w * h * 16, 0, 0), "vsb_device_copy")
                torch.cuda.synchronize()
            else:
                got = image
            verified = bool(torch.equal(got, alone))
        fence()

    line = None
    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        from oracle import bindings as O              # checker / CPU baseline only (never on the product path)
        b_ray, visit = algorithmic_bytes_per_ray(O, S, sc, wl, mats, lights)
        achieved = b_ray * rays_per_step_rank / (kernel_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": measured_traffic(args.workload) if world == 1 and args.scale == 1.0 else None,
                    "kernel": "trace_kernel<FrameProgram>", "kernel_ms": kernel_ms, "bytes_per_ray": b_ray,
                    "n_inner_per_ray": visit["n_inner"], "n_prim_per_ray": visit["n_prim"], "peak_source": peak_src,
                    "roofline_mrays_s": peak * 1e9 / b_ray / 1e6}
        cpu = None
        if world == 1 and not args.no_cpu:
            wh = (1024, 1024) if wl["w"] >= 1024 else (wl["w"], wl["h"])
            v, secs, rays, kind, cores = cpu_frame_rate(O, S, sc, wl, mats, lights, wh, reps=3, warmup=1)
            cpu = {"value": v, "unit": "Mrays/s", "cores": cores, "kind": kind,
                   "sample": f"{wh[0]}x{wh[1]} frame of the same view, {rays} rays, best of 3 after 1 warm-up, "
                             + ("tiled_sched<simd::ray8> -O3 -mavx2 -mfma" if kind == "reference" else "scalar C port")}
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["desc"], "frame": [w, h], "rays_per_step": rays_total // args.steps,
                       "traverse": args.traverse, "scale": args.scale,
                       "partition": f"interleaved 64x64 tiles over {world} GPU(s), scene replicated"
                                    + {"single": "", "peer": ", tiles stored in place into rank 0's frame over NVLink peer memory (CUDA IPC) + 1 tiny all-reduce per frame",
                                       "nccl": ", NCCL all-gather + de-interleave per frame on a side stream"}[mode],
                       "assembled_frame_verified": verified,
                       "l2": f"framebuffer {w * h * 16 >> 20} MiB rewritten every step + scene {(stats['wide_node_bytes'] + stats['wide_prim_bytes']) >> 20} MiB (> 126 MB L2 together)",
                       "wide_nodes": stats["wide_nodes"], "bvh2_nodes": stats["bvh2_nodes"], "triangles": sc.n_prims,
                       "builder": builder},
            "roofline": roofline, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "steps": e2e_steps, "ms_per_step": float(te.item()) * 1e3 / e2e_steps, "trace_ms": e2e_trace_ms},
            "gpu_launches": int(launches), "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    dev.close()
    return 0


def run_gpu_raylist(args):
    """cfg5: incoherent random-ray stress.  Rays uniform in the unit cube with uniform directions
    vs the triangle soup, processed in batches of 2^24; a step = one vsb_closest_hit (or
    vsb_any_hit with --any) pass over one batch.  N > 1: every rank gets its own batches
    (contiguous chunks of the global ray list, different seeds), no collective."""
    import torch
    import torch.distributed as dist

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the ray-query path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    from visionaray_b200 import api, scenes as S

    api.init(local_rank)
    t0 = time.time()
    sc = S.make_scene("cfg5_soup", args.scale)
    build_s = time.time() - t0
    dev, builder = make_device_scene(api, sc, args)
    stats = dev.stats()
    n_batch = args.batch
    n_bufs = 4                                            # rotate batches: 4 x (512 MiB rays + 512 MiB hits) >> L2
    main = torch.cuda.Stream()
    torch.cuda.set_stream(main)
    rays_h = [S.random_rays(n_batch, 600 + 16 * rank + i) for i in range(n_bufs)]
    d_rays = [torch.from_numpy(r).cuda() for r in rays_h]
    d_hits = [torch.empty((n_batch, 8), dtype=torch.int32, device="cuda") for _ in range(n_bufs)]
    fn = api.any_hit if args.any else api.closest_hit
    trav = api.TRAVERSE_EXACT if args.traverse == "exact" else api.TRAVERSE_WIDE
    k = [0]

    def step():
        i = k[0] % n_bufs
        k[0] += 1
        fn(dev, d_rays[i], hits=d_hits[i], traverse=trav, stream=main.cuda_stream)

    def fence():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step()
    fence()
    launches0 = api.launch_count()
    clocks = ClockSampler(local_rank)
    clocks.start()
    time.sleep(0.25)
    fence()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    tw0 = time.time()
    e0.record(main)
    for _ in range(args.steps):
        step()
    e1.record(main)
    fence()
    tw1 = time.time()
    t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total = float(t.item())
    clk = clocks.stop(tw0, tw1)
    launches = (api.launch_count() - launches0) * world
    value = n_batch * args.steps * world / (ms_total * 1e-3) / 1e6
    kms = []
    for _ in range(args.steps):
        step()
        kms.append(api.last_kernel_ms())
    fence()
    kernel_ms = float(np.mean(kms))
    hit_frac = float((d_hits[0][:, 0] != 0).float().mean().item())

    # e2e: host ray list in, host hit records out (pinned), H2D + D2H inside the call
    n_e2e = n_batch
    h_rays = torch.from_numpy(rays_h[0][:n_e2e].copy()).pin_memory().numpy()
    h_hits = torch.zeros((n_e2e, 8), dtype=torch.int32).pin_memory().numpy().view(api.N.HIT_DTYPE).reshape(-1)
    fn(dev, h_rays, hits=h_hits, traverse=trav)
    fence()
    e2e_steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        fn(dev, h_rays, hits=h_hits, traverse=trav)
    fence()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_value = n_e2e * e2e_steps * world / float(te.item()) / 1e6
    e2e_trace_ms = api.last_kernel_ms()           # device time from the first to the last piece's trace of the last call

    if rank == 0:
        from oracle import bindings as O              # checker / CPU baseline only
        peak, peak_src = measured_peak_gbs()
        sample = rays_h[0][:200_000]
        c = O.Counters()
        t0 = time.perf_counter()
        O.oracle_query(sc, sample, bool(args.any), counters=c)
        port_s = time.perf_counter() - t0
        n_inner, n_prim = c.n_inner / c.n_rays, c.n_prim / c.n_rays
        b_ray = 64.0 * n_inner + 68.0 * n_prim + 64.0
        achieved = b_ray * n_batch / (kernel_ms * 1e-3) / 1e9
        cpu = None
        if world == 1 and not args.no_cpu:
            cores = os.cpu_count() or 1
            if O.ref_available(fast=True):
                nref = 2_000_000
                secs, _ = O.ref_query_timed(sc, rays_h[0][:nref], bool(args.any), nthreads=cores, reps=2, warmup=1)
                cpu = {"value": nref / secs / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "reference",
                       "sample": f"{nref} rays of the same list, scalar closest_hit()/any_hit() of the unmodified headers on {cores} std::threads, best of 2"}
            else:
                cpu = {"value": 200_000 / port_s / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
                       "sample": "200000 rays of the same list, scalar C port, pthreads"}
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"incoherent random rays vs {sc.n_prims}-tri soup, {'any' if args.any else 'closest'}-hit, "
                                   f"batches of {n_batch} rays per GPU", "traverse": args.traverse, "scale": args.scale,
                       "l2": f"{n_bufs} rotating ray/hit batches ({n_bufs * n_batch * 64 >> 20} MiB) + scene "
                             f"{(stats['wide_node_bytes'] + stats['wide_prim_bytes']) >> 20} MiB, all > 126 MB L2",
                       "builder": builder,
                       "wide_nodes": stats["wide_nodes"], "bvh2_nodes": stats["bvh2_nodes"], "bvh2_depth": stats["bvh2_depth"],
                       "host_build_s": build_s, "hit_fraction": hit_frac},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": measured_traffic("cfg5") if world == 1 and args.scale == 1.0 and not args.any else None,
                         "kernel": "trace_kernel<RayListProgram>", "kernel_ms": kernel_ms, "bytes_per_ray": b_ray,
                         "n_inner_per_ray": n_inner, "n_prim_per_ray": n_prim, "peak_source": peak_src,
                         "roofline_mrays_s": peak * 1e9 / b_ray / 1e6},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": "Mrays/s", "h2d_bytes_per_step": n_e2e * 32 * world,
                    "d2h_bytes_per_step": n_e2e * 32 * world, "steps": e2e_steps, "rays_per_step": n_e2e,
                    "ms_per_step": float(te.item()) * 1e3 / e2e_steps, "trace_ms": e2e_trace_ms},
            "gpu_launches": int(launches), "clocks": clk,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    dev.close()
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
    ap.add_argument("--builder", default="reference", choices=["reference", "gpu"],
                    help="whose BVH the scene uses: the reference's host builder, or the GPU builder (vsb_scene_build)")
    ap.add_argument("--scale", type=float, default=1.0, help="scene size factor (1.0 = the BASELINE config)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
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
