"""Host side of the multi-GPU path on CPU: world_size-2 gloo all-gather of interleaved tile buffers
+ de-interleave rebuilds the frame; tile counts agree with the C ABI's vsb_tiles_of_rank."""
import ctypes as C
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from visionaray_b200 import _native as N
from visionaray_b200 import partition as P


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, w, h, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(7)                      # same image on every rank
    image = rng.random((h, w, 4), dtype=np.float32)
    mine = torch.from_numpy(P.pack_tiles(image, rank, world))
    per = P.padded_tiles_per_rank(w, h, world)
    gathered = torch.zeros((world, per, P.TILE * P.TILE, 4), dtype=torch.float32)
    dist.all_gather_into_tensor(gathered.view(-1), mine.view(-1))
    out = P.deinterleave(gathered.numpy(), w, h, world)
    ok = bool(np.array_equal(out, image))
    # every pixel is owned by exactly one rank
    counts = torch.tensor([P.tiles_of_rank(w, h, rank, world)], dtype=torch.int64)
    dist.all_reduce(counts)
    tx, ty = P.tiles_xy(w, h)
    ok = ok and int(counts.item()) == tx * ty
    if rank == 0:
        q.put(ok)
    dist.destroy_process_group()


def test_two_rank_gather_rebuilds_the_frame():
    ctx = mp.get_context("spawn")
    for (w, h) in ((200, 136), (256, 128)):
        q = ctx.Queue()
        port = _free_port()
        procs = [ctx.Process(target=_worker, args=(r, 2, port, w, h, q)) for r in range(2)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        assert q.get(timeout=10) is True


def test_tile_counts_match_c_abi():
    lib = N.cuda_lib()      # loads without a GPU; vsb_tiles_of_rank is pure host arithmetic
    for (w, h, n) in ((200, 136, 3), (4096, 4096, 8), (16384, 8192, 8), (65, 1, 2)):
        for r in range(n):
            out = C.c_int(0)
            assert lib.vsb_tiles_of_rank(w, h, r, n, C.byref(out)) == 0
            assert out.value == P.tiles_of_rank(w, h, r, n)
    out = C.c_int(0)
    assert lib.vsb_tiles_of_rank(64, 64, 2, 2, C.byref(out)) != 0      # rank out of range


def test_abi_exports_every_declared_symbol():
    import re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "vsnray_b200.h")).read()
    declared = set(re.findall(r"\b(vsb_[a-z0-9_]+)\s*\(", hdr))
    lib = N.cuda_lib()
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, missing
    assert declared == set(N.ABI_SYMBOLS), declared ^ set(N.ABI_SYMBOLS)


def test_compute_entry_points_fail_without_device():
    if torch.cuda.is_available():
        return
    lib = N.cuda_lib()
    assert lib.vsb_init(0) == -1                         # VSB_ERR_NO_DEVICE
    assert lib.vsb_closest_hit(None, 0, None, 1, C.c_float(1.0), None, None, 1, 0, None) == -4   # NOT_INIT
    assert b"no CUDA device" in lib.vsb_strerror(-1)
