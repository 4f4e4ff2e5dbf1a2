"""GPU tier: frames and multi_hit over a RANGE of bvhs (detail/traverse_linear.inl:71-130, get_normal.h:28-37:
normals[prim_id] is one array over the whole range), calls on mixed streams, and what the Python layer refuses."""
import dataclasses

import numpy as np
import pytest

from oracle import bindings as O
from test_hostsim import scene_rays
from visionaray_b200 import scenes as S

pytestmark = pytest.mark.gpu


def _two_scenes():
    """A sphere mesh and a displaced grid as two bvhs with globally unique prim_ids, plus the same primitives
    merged into ONE bvh (what the oracle renders) and the combined per-face normals."""
    a = S.make_scene("cfg1_sphere", 0.05)
    b = S.make_scene("cfg2_grid", 0.01)
    na = a.n_prims
    pb = b.prims.copy()
    pb.view(np.uint32)[:, 1] += na                                     # prim_id continues after scene a's
    b2 = dataclasses.replace(b, prims=pb, normals=None)
    a2 = dataclasses.replace(a, normals=None)
    normals = np.concatenate([a.normals, b.normals])
    prims = np.concatenate([a.prims, pb])
    nodes, idx = S.build_bvh(0, prims)
    merged = S.HostScene(0, prims, nodes, idx, normals, "merged")
    return a2, b2, merged, normals


def test_frame_over_two_bvhs_indexes_one_normals_array(gpu):
    a, b, merged, normals = _two_scenes()
    da, db = gpu.Scene(a), gpu.Scene(b)
    eye = (0.3, 1.2, 2.2)
    mats, lights = S.default_material(0), S.point_light(eye)
    for kernel in (0, 1):
        fp = S.frame_params(kernel, 192, 128, eye, (0, 0, 0), num_bounces=3)
        want = O.oracle_frame(merged, fp, mats, lights)
        for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
            img = np.zeros((128, 192, 4), dtype=np.float32)
            gpu.frame([da, db], fp, normals, mats, lights, img, traverse=trav)        # normals passed with the call
            bad = float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())
            assert bad <= 2e-4, (kernel, trav, bad)                                      # t-ties between the two bvhs only
    # the same with the array attached to scenes[0]; an array that does not reach scene b's prim_ids is refused
    fp = S.frame_params(1, 192, 128, eye, (0, 0, 0), num_bounces=3)
    want = O.oracle_frame(merged, fp, mats, lights)
    da.set_normals(normals)
    img = np.zeros((128, 192, 4), dtype=np.float32)
    gpu.frame([da, db], fp, None, mats, lights, img)
    assert float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean()) <= 2e-4
    da.set_normals(normals[:a.n_prims].copy())
    with pytest.raises(RuntimeError):
        gpu.frame([da, db], fp, None, mats, lights, img)
    da.close(); db.close()


def test_multi_hit_reports_the_same_geometry_in_two_bvhs_twice(gpu):
    sc = S.make_scene("cfg3_cubes", 0.001)
    d1, d2 = gpu.Scene(sc), gpu.Scene(sc)
    rays = scene_rays(sc, 20_000, 7)
    want = O.oracle_multi_hit([sc, sc], rays, 8)
    exact = gpu.multi_hit([d1, d2], rays, 8, traverse=gpu.TRAVERSE_EXACT)
    assert exact.tobytes() == want.tobytes()
    wide = gpu.multi_hit([d1, d2], rays, 8)
    assert (wide["hit"] == want["hit"]).all()
    assert np.array_equal(wide["t"], want["t"])                        # every hit twice, as in the reference
    assert (wide["pad"] == 0).all()
    d1.close(); d2.close()


def test_calls_on_mixed_streams_are_ordered(gpu):
    """Device-buffer calls alternate between torch's current stream, a side stream and the library's own stream
    without any synchronisation in between: the library orders them, so every result equals the synchronous one."""
    import torch

    sc = S.make_scene("cfg3_cubes", 0.004)
    dev = gpu.Scene(sc)
    rays = scene_rays(sc, 400_000, 3)
    want = gpu.closest_hit(dev, rays)
    d_rays = torch.from_numpy(rays).cuda()
    side = torch.cuda.Stream()
    outs = []
    for k in range(6):
        hits = torch.empty((rays.shape[0], 8), dtype=torch.int32, device="cuda")
        if k % 3 == 0:
            gpu.closest_hit(dev, d_rays, hits=hits)                                    # torch's current stream
        elif k % 3 == 1:
            gpu.closest_hit(dev, d_rays, hits=hits, stream=side.cuda_stream)
        else:
            gpu.any_hit(dev, d_rays, hits=hits, max_t=0.3)                             # different work in between
        outs.append(hits)
    torch.cuda.synchronize()
    gpu.sync()
    for k in (0, 1, 3, 4):
        assert outs[k].cpu().numpy().view(O.HIT_DTYPE).reshape(-1).tobytes() == want.tobytes(), k
    # a tensor produced on the current stream right before the call is consumed in order (no explicit sync)
    d2 = d_rays.clone()
    h2 = gpu.closest_hit(dev, d2)
    torch.cuda.synchronize()
    assert h2.cpu().numpy().view(O.HIT_DTYPE).reshape(-1).tobytes() == want.tobytes()
    dev.close()


def test_python_layer_refuses_buffers_it_cannot_pass_as_pointers(gpu):
    sc = S.make_scene("cfg1_sphere", 0.02)
    dev = gpu.Scene(sc)
    rays = S.random_rays(1000, 3, -1.5, 1.5)
    with pytest.raises(ValueError):
        gpu.closest_hit(dev, rays[::2])                                # not contiguous
    with pytest.raises(ValueError):
        gpu.closest_hit(dev, rays[:, :6].copy())                       # not 32-byte records
    with pytest.raises(ValueError):
        gpu.closest_hit(dev, rays, max_t_per_ray=np.ones(1000, dtype=np.float64))
    dev.close()


def test_overlapping_leaf_ranges_are_rejected(gpu):
    """A caller-supplied tree whose leaves cover more list positions than the list has (overlaps) would overflow the
    re-encode's leaf list: VSB_ERR_BAD_TREE at vsb_scene_create."""
    sc = S.make_scene("cfg1_sphere", 0.02)
    nodes = sc.nodes.copy()
    leaves = np.flatnonzero(nodes["num_prims"] > 0)
    small = [i for i in leaves if nodes["num_prims"][i] < 4 and nodes["first"][i] > 0]
    i = small[0]
    nodes["first"][i] -= 1                                             # now overlaps its left neighbour's range ...
    nodes["num_prims"][i] += 1                                         # ... and the total exceeds the list length
    with pytest.raises(RuntimeError):
        gpu.Scene(dataclasses.replace(sc, nodes=nodes))


def test_mirror_frame_of_compact_progressive_tiles(gpu):
    """Progressive blending over 3 'ranks' (one after the other on one GPU): each keeps its accumulation in its compact
    tile buffer and mirrors every finished pixel into ONE row-major frame (what peers do into rank 0's buffer over
    NVLink): after 3 accumulated frames the mirror equals the frame one rank accumulates alone, byte for byte."""
    import torch

    sc = S.make_scene("cfg2_grid", 0.02)
    dev = gpu.Scene(sc)
    eye = (0.3, 1.2, 2.2)
    w, h, nr = 328, 200, 3                             # ragged right / top tiles
    mats, lights = S.default_material(1), S.point_light(eye)
    d_m = torch.from_numpy(mats.view(np.float32).copy()).cuda()
    d_l = torch.from_numpy(lights.view(np.float32).copy()).cuda()
    alone = torch.zeros((h, w, 4), dtype=torch.float32, device="cuda")
    per = max(gpu.tiles_of_rank(w, h, r, nr) for r in range(nr))
    local = [torch.zeros((per, 64 * 64, 4), dtype=torch.float32, device="cuda") for _ in range(nr)]
    mirror = torch.full((h, w, 4), -1.0, dtype=torch.float32, device="cuda")
    for frame_num in (1, 2, 3):
        fp = S.frame_params(3, w, h, eye, (0, 0, 0), pixel_sampler=2, frame_num=frame_num, num_bounces=3, ambient=(1, 1, 1, 1))
        gpu.frame(dev, fp, None, d_m, d_l, alone)
        for r in range(nr):
            gpu.frame(dev, fp, None, d_m, d_l, local[r], tile_rank=r, tile_nranks=nr, mirror=mirror)
    torch.cuda.synchronize()
    gpu.sync()
    assert torch.equal(mirror, alone)
    fp = S.frame_params(3, w, h, eye, (0, 0, 0), pixel_sampler=2, frame_num=1, num_bounces=3, ambient=(1, 1, 1, 1))
    with pytest.raises(RuntimeError):                  # a mirror needs compact tile buffers
        gpu.frame(dev, fp, None, d_m, d_l, alone, mirror=mirror)
    dev.close()

