"""sin / cos / pow of sampled directions: the device code's sequence (csrc/shade.cuh ray_sincosf / ray_powf,
host instantiation through tests/hostsim) and the oracle's restatement (oracle/vso_raymath.h) give the same
bits, both are the correctly rounded float, and libm -- what the reference calls (sampling.h:61-71,
brdf.h:121-150) -- is within one ulp of them."""
import ctypes as C

import numpy as np

from oracle import bindings as O
import hostsim_bindings as H


def _p(a):
    return C.c_void_p(a.ctypes.data)


def _both_sincos(x):
    o = O.oracle_lib()
    h = H.lib()
    so, co, sh, ch = (np.empty_like(x) for _ in range(4))
    o.vso_ray_sincosf(_p(x), C.c_size_t(x.size), _p(so), _p(co))
    h.hs_ray_sincosf(_p(x), C.c_size_t(x.size), _p(sh), _p(ch))
    return so, co, sh, ch


def _within_one_ulp(a, b):
    return bool((np.abs(a.astype(np.float64) - b.astype(np.float64)) <= np.spacing(np.abs(b)).astype(np.float64)).all())


def test_sincos_device_sequence_equals_oracle_and_is_correctly_rounded():
    rng = np.random.default_rng(11)
    # theta = 2 pi u, u a 24-bit sampler output (vso_sampler_next), plus a spread of other angles
    u = (rng.integers(0, 1 << 24, 2_000_000).astype(np.float32)) * np.float32(1.0 / 16777216.0)
    x = np.concatenate([np.float32(6.28318530717958647692) * u, rng.uniform(-50.0, 50.0, 200_000).astype(np.float32),
                        np.array([0.0, -0.0, 1e-30, 3.14159274, 1.57079637, 6.28318548], dtype=np.float32)])
    so, co, sh, ch = _both_sincos(x)
    assert so.tobytes() == sh.tobytes() and co.tobytes() == ch.tobytes()
    xd = x.astype(np.float64)
    assert np.array_equal(so, np.sin(xd).astype(np.float32))
    assert np.array_equal(co, np.cos(xd).astype(np.float32))
    # float-precision sin / cos (what the reference calls through libm): never further than one ulp
    assert _within_one_ulp(so, np.sin(x)) and _within_one_ulp(co, np.cos(x))


def test_pow_device_sequence_equals_oracle_and_is_correctly_rounded():
    rng = np.random.default_rng(12)
    n = 1_000_000
    x = (rng.integers(0, 1 << 24, n).astype(np.float32)) * np.float32(1.0 / 16777216.0)
    e = rng.uniform(0.0, 200.0, n).astype(np.float32)
    y = np.where(np.arange(n) & 1, np.float32(1.0) / (e + np.float32(1.0)), e).astype(np.float32)
    x = np.concatenate([x, np.array([0.0, 0.0, 1.0, 0.5, 2.0, 1e-38, 3.0e38], dtype=np.float32)])
    y = np.concatenate([y, np.array([2.0, 0.0, 7.0, 2.0, 10.0, 0.5, 0.25], dtype=np.float32)])
    a, b = np.empty_like(x), np.empty_like(x)
    O.oracle_lib().vso_ray_powf(_p(x), _p(y), C.c_size_t(x.size), _p(a))
    H.lib().hs_ray_powf(_p(x), _p(y), C.c_size_t(x.size), _p(b))
    assert a.tobytes() == b.tobytes()
    want = np.power(x.astype(np.float64), y.astype(np.float64)).astype(np.float32)
    assert np.array_equal(a, want)
    with np.errstate(all="ignore"):
        lm = np.power(x, y)
    ok = np.isfinite(lm) & (lm > 1e-37)
    assert _within_one_ulp(a[ok], lm[ok])
