"""The reference's OWN CUDA path (cuda_sched<basic_ray<float>>::frame() + whitted::kernel over a cuda_index_bvh, unmodified
headers compiled for sm_100a: oracle/refcuda) as a second checker: its frame agrees with the oracle -- the host scalar
instantiation -- except where the device min / max trick of math/detail/math.h:176-201 or a t-tie
picks another primitive, and with this repository's kernels likewise.  It is the `naive_gpu` baseline of bench.py."""
import numpy as np
import pytest

from oracle import bindings as O
from visionaray_b200 import scenes as S

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(not O.ref_cuda_available(), reason="oracle/_ref/libvsnray_ref_cuda.so is built where /root/reference and nvcc exist")
def test_reference_cuda_frame_agrees_with_oracle_and_with_this_repo(gpu):
    sc = S.make_scene("cfg2_grid", 0.05)
    eye = (0.3, 1.2, 2.2)
    fp = S.frame_params(1, 512, 384, eye, (0, 0, 0), num_bounces=4, epsilon=1e-3)
    mats, lights = S.default_material(0), S.point_light(eye)
    want = O.oracle_frame(sc, fp, mats, lights)
    img, ms, build_ms = O.ref_cuda_whitted(sc, fp, eye, reps=2, warmup=1)
    assert ms > 0 and np.isfinite(img).all()
    bad = float((np.abs(img - want).max(axis=2) > 1.0 / 255.0).mean())
    assert bad <= 5e-3, bad
    dev = gpu.Scene(sc)
    mine = np.zeros((384, 512, 4), dtype=np.float32)
    gpu.frame(dev, fp, sc.normals, mats, lights, mine)
    assert float((np.abs(img - mine).max(axis=2) > 1.0 / 255.0).mean()) <= 5e-3
    dev.close()
