"""vsb_closest_hit_static through the C ABI: the no-allocation variant of vsb_closest_hit (scratch in a file-scope
__device__ array; include/vsnray_b200.h) gives the hit records of the existing entry point, on two calls in a row
(the array keeps the last call's contents), and refuses lists beyond its fixed maximum.  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from visionaray_b200 import _native as N
from visionaray_b200 import scenes as S

pytestmark = pytest.mark.gpu

STATIC_MAX_RAYS = 1 << 20
ERR_TOO_LARGE = -6


def _static_closest_hit(gpu, dev, d_rays, n, max_t=3.4028234663852886e38):
    import torch

    hits = torch.empty((n, 8), dtype=torch.int32, device="cuda")       # the CALLER's output buffer
    stream = torch.cuda.current_stream().cuda_stream or 1
    rc = N.cuda_lib().vsb_closest_hit_static(dev.handle, C.c_void_p(d_rays.data_ptr()), n, max_t, C.c_void_p(hits.data_ptr()), stream)
    return rc, hits


def test_static_scratch_variant_matches_vsb_closest_hit_twice_in_a_row(gpu):
    import torch

    sc = S.make_scene("cfg3_cubes", 0.01)
    dev = gpu.Scene(sc)
    # two lists in a row through the same scratch array: a large sorted one (>= 64 Ki rays: coherence sort inside the
    # array), then a smaller unsorted one with a finite max_t that would go wrong on stale counters / order / scene table
    lists = [(S.random_rays(300_000, 11, -1.2, 1.2), 3.4028234663852886e38), (S.random_rays(40_001, 12, -1.2, 1.2), 0.75)]
    for k in range(2):
        for rays, max_t in lists:
            d_rays = torch.from_numpy(rays).cuda()
            want = gpu.closest_hit(dev, d_rays, max_t=max_t)
            rc, got = _static_closest_hit(gpu, dev, d_rays, rays.shape[0], max_t)
            assert rc == 0, N.cuda_lib().vsb_strerror(rc)
            torch.cuda.synchronize()
            gpu.sync()
            # same kernels, same (t, record) tie rule: the records are identical, not merely within the t / u / v tolerance
            assert torch.equal(got, want), (k, rays.shape[0])
            assert int((got[:, 0] != 0).sum().item()) > 0
    # back to back without any synchronisation in between: the library orders the calls of a context
    d_a = torch.from_numpy(lists[0][0]).cuda()
    d_b = torch.from_numpy(lists[1][0]).cuda()
    rc1, h1 = _static_closest_hit(gpu, dev, d_a, d_a.shape[0])
    rc2, h2 = _static_closest_hit(gpu, dev, d_b, d_b.shape[0])
    assert rc1 == 0 and rc2 == 0
    torch.cuda.synchronize()
    assert torch.equal(h1, gpu.closest_hit(dev, d_a)) and torch.equal(h2, gpu.closest_hit(dev, d_b))
    torch.cuda.synchronize()
    gpu.sync()
    dev.close()


def test_static_variant_rejects_lists_beyond_its_maximum(gpu):
    import torch

    sc = S.make_scene("cfg1_sphere", 0.02)
    dev = gpu.Scene(sc)
    d_rays = torch.zeros((16, 8), dtype=torch.float32, device="cuda")
    rc, _ = _static_closest_hit(gpu, dev, d_rays, 16)
    assert rc == 0
    hits = torch.empty((16, 8), dtype=torch.int32, device="cuda")
    rc = N.cuda_lib().vsb_closest_hit_static(dev.handle, C.c_void_p(d_rays.data_ptr()), STATIC_MAX_RAYS + 1, 1.0, C.c_void_p(hits.data_ptr()), None)
    assert rc == ERR_TOO_LARGE
    assert b"fixed maximum" in N.cuda_lib().vsb_strerror(rc)
    assert N.cuda_lib().vsb_closest_hit_static(dev.handle, None, 16, 1.0, C.c_void_p(hits.data_ptr()), None) == -2     # bad argument
    torch.cuda.synchronize()
    gpu.sync()
    dev.close()
