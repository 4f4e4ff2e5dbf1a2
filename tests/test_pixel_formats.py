"""Render targets of other colour formats: PF_RGBA8, PF_RGB32F, PF_RGB8 next to
PF_RGBA32F -- pixel_access::store / get / blend with the reference's colour conversion
(detail/pixel_access.h:38-58,668-689,1153-1175, detail/color_conversion.h:168-265, math/detail/unorm.inl:18-31):
unorm8 by truncation, RGBA -> RGB times alpha, RGB -> RGBA with alpha 1, blends and SSAA accumulate through
the stored format.  CPU: oracle vs fixtures generated from the reference headers (and vs the headers where
oracle/_ref exists).  GPU: vsb_frame with color_format vs the oracle."""
import os
import sys

import numpy as np
import pytest

from oracle import bindings as O
from visionaray_b200 import scenes as S

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))
import make_golden as MG  # noqa: E402

FORMATS = ["PF_RGBA8", "PF_RGB32F", "PF_RGB8"]


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_formats.npz"))


def stored(buf, fmt):
    """The channels the format stores (the 4th word of a PF_RGB32F pixel is never written)."""
    return buf[..., :3] if fmt == "PF_RGB32F" else buf


def accumulate(fn, sc, fp, mats, light, fmt, ps, **kw):
    buf = O.format_buffer(fmt, fp.width, fp.height)
    for frame_num in ((1, 2, 3) if ps == 2 else (1,)):
        fp.frame_num = frame_num
        fn(sc, fp, mats, light, fmt, buf, **kw)
    return buf


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("kernel,ps,mat", MG.FORMAT_CASES)
def test_oracle_matches_reference_fixture(golden, fmt, kernel, ps, mat):
    sc, fp, mats, light = MG.format_case(kernel, ps, mat)
    got = accumulate(O.oracle_frame_format, sc, fp, mats, light, fmt, ps)
    want = golden[f"{fmt}_k{kernel}_ps{ps}"]
    a, b = stored(got, fmt), stored(want, fmt)
    if a.dtype == np.uint8:
        # same libm in the build container: identical; another libm may move a value across a truncation boundary
        assert int(np.abs(a.astype(np.int32) - b.astype(np.int32)).max()) <= 1 and float((a != b).mean()) <= 1e-3
    else:
        assert np.abs(a - b).max() <= 1e-6
    assert b.std() > 0


def test_conversion_rules():
    """Known answers of the conversion: truncation (not rounding), clamp, alpha."""
    import ctypes as C
    lib = O.oracle_lib()
    lib.vso_store_pixel.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    lib.vso_get_pixel.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
    src = np.array([0.999, 0.5, -0.25, 1.5], dtype=np.float32)
    q = np.zeros(4, dtype=np.uint8)
    lib.vso_store_pixel(1, src.ctypes.data, q.ctypes.data)
    assert q.tolist() == [254, 127, 0, 255]                      # 0.999 * 255 = 254.7 -> 254; 0.5 * 255 = 127.5 -> 127
    back = np.zeros(4, dtype=np.float32)
    lib.vso_get_pixel(1, q.ctypes.data, back.ctypes.data)
    assert back.tolist() == [np.float32(254 / 255), np.float32(127 / 255), 0.0, 1.0]
    rgb = np.zeros(4, dtype=np.float32)
    lib.vso_store_pixel(2, np.array([0.2, 0.4, 0.8, 0.5], dtype=np.float32).ctypes.data, rgb.ctypes.data)
    assert rgb.tolist() == [np.float32(0.2) * np.float32(0.5), np.float32(0.4) * np.float32(0.5), np.float32(0.8) * np.float32(0.5), 0.0]
    lib.vso_get_pixel(2, rgb.ctypes.data, back.ctypes.data)
    assert back[3] == 1.0
    q3 = np.zeros(3, dtype=np.uint8)
    lib.vso_store_pixel(3, np.array([1.0, 0.5, 0.25, 0.5], dtype=np.float32).ctypes.data, q3.ctypes.data)
    assert q3.tolist() == [127, 63, 31]


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref is only built where /root/reference exists")
@pytest.mark.parametrize("fmt", ["PF_RGBA32F"] + FORMATS)
@pytest.mark.parametrize("kernel,ps,mat", [(0, 1, 0), (1, 3, 0), (3, 2, 0), (1, 0, 1)])
def test_oracle_matches_reference_headers(fmt, kernel, ps, mat):
    sc, fp, mats, light = MG.format_case(kernel, ps, mat, 64, 48)
    got = accumulate(O.oracle_frame_format, sc, fp, mats, light, fmt, ps)
    want = accumulate(O.ref_frame_format, sc, fp, mats, light, fmt, ps)
    a, b = stored(got, fmt), stored(want, fmt)
    if kernel != 3:
        assert a.tobytes() == b.tobytes()
        return
    # path tracer: the oracle's sin / cos / pow of sampled directions are the correctly rounded values
    # (oracle/vso_raymath.h), the reference's are libm's -- one ulp apart in ~1 % of the draws
    if a.dtype == np.uint8:
        assert np.abs(a.astype(np.int32) - b.astype(np.int32)).max() <= 1
    else:
        assert np.abs(a - b).max() <= 1e-6


# ---- GPU ------------------------------------------------------------------------------------------

FMT_ID = {k: v[0] for k, v in O.PIXEL_FORMATS.items()}


def check(got, want, fmt, limit):
    a, b = stored(got, fmt), stored(want, fmt)
    if a.dtype == np.uint8:
        d = np.abs(a.astype(np.int32) - b.astype(np.int32)).max(axis=2)
        bad = float((d > 1).mean())            # one step of the format = 1/255, the north star's pixel tolerance
    else:
        bad = float((np.abs(a - b).max(axis=2) > 1.0 / 255.0).mean())
    assert bad <= limit, (fmt, bad)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("kernel,ps,mat", MG.FORMAT_CASES)
def test_gpu_host_targets(gpu, fmt, kernel, ps, mat):
    sc, fp, mats, light = MG.format_case(kernel, ps, mat, 192, 128)
    want = accumulate(O.oracle_frame_format, sc, fp, mats, light, fmt, ps)
    dev = gpu.Scene(sc)
    try:
        for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
            buf = O.format_buffer(fmt, fp.width, fp.height)
            if fmt == "PF_RGB32F":
                buf[..., 3] = 7.0                                  # the unused word survives
            for frame_num in ((1, 2, 3) if ps == 2 else (1,)):
                fp.frame_num = frame_num
                gpu.frame(dev, fp, None, mats, light, buf, traverse=trav, color_format=FMT_ID[fmt])
            check(buf, want, fmt, (0.0 if trav == gpu.TRAVERSE_EXACT else 1e-3) if kernel in (0, 1) else 4e-3)
            if fmt == "PF_RGB32F":
                assert (buf[..., 3] == 7.0).all()
    finally:
        dev.close()


@pytest.mark.gpu
@pytest.mark.parametrize("fmt", FORMATS)
def test_gpu_device_targets_full_size(gpu, fmt):
    """cfg2 at 2048 x 2048 into device targets: a formatted frame == the RGBA32F frame converted on the host
    with the same rules (size-independent property), byte for byte."""
    import torch

    sc = S.make_scene("cfg2_grid", 1.0)
    fp = S.frame_params(1, 2048, 2048, (0.3, 1.2, 2.2), (0, 0, 0), num_bounces=4, bg=(0.1, 0.4, 1.0, 0.8))
    mats, light = S.default_material(0), S.point_light((0.3, 1.2, 2.2))
    dev = gpu.Scene(sc)
    try:
        d_m = torch.from_numpy(mats.view(np.uint8)).cuda(); d_l = torch.from_numpy(light.view(np.uint8)).cuda()
        ref = torch.zeros((2048, 2048, 4), dtype=torch.float32, device="cuda")
        gpu.frame(dev, fp, None, d_m, d_l, ref)
        _, bpp, dt, ch = O.PIXEL_FORMATS[fmt]
        out = torch.zeros((2048, 2048, ch), dtype=torch.float32 if dt == np.float32 else torch.uint8, device="cuda")
        gpu.frame(dev, fp, None, d_m, d_l, out, color_format=FMT_ID[fmt])
        gpu.sync(); torch.cuda.synchronize()
        r = ref.cpu().numpy(); o = out.cpu().numpy()

        def unorm(x):
            return (np.clip(x, 0.0, 1.0).astype(np.float32).astype(np.float64) * 255.0).astype(np.uint32).astype(np.uint8)

        if fmt == "PF_RGBA8":
            want = unorm(r)
        elif fmt == "PF_RGB32F":
            want = r[..., :3] * r[..., 3:4]; o = o[..., :3]
        else:
            want = unorm(r[..., :3] * r[..., 3:4])
        assert o.tobytes() == np.ascontiguousarray(want).tobytes()
    finally:
        dev.close()
