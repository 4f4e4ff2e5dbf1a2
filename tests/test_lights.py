"""spot_light<T> lists (spot_light.h, detail/spot_light.inl:14-38) next to point lights: intensity =
cl kl att spot, spot = dot(spot_direction, normalize(pos - position)) > cos(cutoff) ? pow(spot, exponent) : 0.
CPU: oracle vs the reference headers (where oracle/_ref exists) and vs a fixture generated from them.
GPU: vsb_frame vs the oracle."""
import math
import os

import numpy as np
import pytest

from oracle import bindings as O
from visionaray_b200 import scenes as S

ANGLES = (25.0, 40.0)


def case(kernel, w=96, h=64):
    sc = S.make_scene("cfg2_grid", 0.004)
    fp = S.frame_params(kernel, w, h, (0.9, 0.75, 1.3), (0, 0, 0), num_bounces=4, ambient=(0.1, 0.1, 0.1, 1.0), n_lights=2)
    lights = np.concatenate([S.spot_light((0.4, 1.2, 0.6), (0.0, 0.0, 0.0), ANGLES[0], 8.0),
                             S.spot_light((-0.6, 0.9, -0.3), (0.2, 0.0, 0.1), ANGLES[1], 2.0, cl=(1.0, 0.6, 0.3))])
    return sc, fp, S.default_material(0), lights


@pytest.mark.skipif(not O.ref_available(), reason="oracle/_ref is only built where /root/reference exists")
@pytest.mark.parametrize("kernel", [0, 1])
def test_oracle_matches_reference_headers(kernel):
    sc, fp, mats, lights = case(kernel)
    got = O.oracle_frame(sc, fp, mats, lights)
    want = O.ref_frame_spot(sc, fp, mats, lights, [math.radians(a) for a in ANGLES])
    assert np.abs(got - want).max() <= 1e-6
    as_points = lights.copy(); as_points["kind"] = 0
    assert np.abs(got - O.oracle_frame(sc, fp, mats, as_points)).max() > 0.2       # the cone matters


@pytest.mark.parametrize("kernel", [0, 1])
def test_oracle_matches_reference_fixture(kernel):
    g = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_surface.npz"))
    sc, fp, mats, lights = case(kernel, 40, 28)
    assert np.abs(O.oracle_frame(sc, fp, mats, lights) - g[f"spot_k{kernel}"]).max() <= 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [0, 1])
def test_gpu_spot_lights(gpu, kernel):
    sc, fp, mats, lights = case(kernel, 192, 128)
    want = O.oracle_frame(sc, fp, mats, lights)
    dev = gpu.Scene(sc)
    try:
        for trav in (gpu.TRAVERSE_EXACT, gpu.TRAVERSE_WIDE):
            img = np.zeros((fp.height, fp.width, 4), dtype=np.float32)
            gpu.frame(dev, fp, None, mats, lights, img, traverse=trav)
            bad = float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())
            # a pixel on the rim of a cone can fall on the other side of the cutoff by an ulp of the dot product
            assert bad <= 1e-3, (kernel, trav, bad)
    finally:
        dev.close()
