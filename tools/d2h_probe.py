"""Host-side ceiling of the end-to-end frame path at N GPUs of one box (run under torchrun, or alone for N = 1):
every rank moves one 4096^2 RGBA32F frame (268 MB) per step from its GPU to host memory, all ranks at the same time,
  (a) copy engine -> the rank's own pinned buffer (what per-rank host targets cost),
  (b) copy engine -> the rank's slice of ONE /dev/shm frame registered by all ranks (what an assembled frame costs).
Prints GB/s per rank and aggregate; rank 0 also prints the NUMA node of each GPU (sysfs), when readable."""
import os
import time

import numpy as np
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def fence():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    fence()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    fence()
    dt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return float(dt.item())


n_bytes = 4096 * 4096 * 16
x = torch.zeros(n_bytes // 4, dtype=torch.float32, device="cuda")
own = torch.zeros(n_bytes // 4, dtype=torch.float32).pin_memory()
name = [f"/dev/shm/vsb_probe_{os.getpid()}"] if rank == 0 else [None]
if world > 1:
    dist.broadcast_object_list(name, src=0)
if rank == 0:
    with open(name[0], "wb") as f:
        f.truncate(n_bytes * world)
fence()
shared = np.memmap(name[0], dtype=np.float32, mode="r+", shape=(world, n_bytes // 4))
rt = torch.cuda.cudart()
assert int(rt.cudaHostRegister(shared.ctypes.data, n_bytes * world, 1 | 2)) == 0
mine = torch.from_numpy(shared[rank])
res = {}
res["copy engine -> own pinned buffer"] = timed(lambda: own.copy_(x, non_blocking=True))
res["copy engine -> slice of the shared registered frame"] = timed(lambda: mine.copy_(x, non_blocking=True))
if rank == 0:
    for k, dt in res.items():
        print(f"N={world}  {k:55s} {dt * 1e3:7.2f} ms  {n_bytes / dt / 1e9:6.1f} GB/s per rank  {world * n_bytes / dt / 1e9:7.1f} GB/s aggregate")
    for i in range(world):
        try:
            bus = torch.cuda.get_device_properties(i).pci_bus_id
        except Exception:
            bus = None
        print(f"  gpu {i}: pci {bus}")
    try:
        print("  host:", os.cpu_count(), "cpus;", open("/sys/devices/system/node/online").read().strip(), "NUMA nodes online")
    except OSError:
        pass
fence()
rt.cudaHostUnregister(shared.ctypes.data)
del mine, shared
fence()
if rank == 0:
    os.unlink(name[0])
if world > 1:
    dist.destroy_process_group()
