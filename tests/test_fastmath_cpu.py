"""The product's fp64 polynomial sincos/atan (csrc/mpcb_fastmath.cuh), compiled for the host, against libm."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def fm():
    src = os.path.join(HERE, "native", "fastmath_host.cpp")
    out_dir = os.path.join(HERE, "native", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libfastmath_host.so")
    hdrs = [os.path.join(HERE, "..", "model-predictive-control_b200", "csrc", f)
            for f in ("mpcb_fastmath.cuh", "mpcb_fastmath_coef.h")]
    if not os.path.exists(so) or any(os.path.getmtime(p) > os.path.getmtime(so) for p in [src] + hdrs):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-mfma", "-ffp-contract=off",
                               "-o", so, src])
    lib = C.CDLL(so)
    dp = C.POINTER(C.c_double)
    lib.fmh_sincos.argtypes = [dp, dp, dp, C.c_long]
    lib.fmh_atan.argtypes = [dp, dp, C.c_long]
    lib.fmh_sin.argtypes = [dp, dp, C.c_long]
    lib.fmh_sincos_x2.argtypes = [dp, dp, dp, C.c_long]
    lib.fmh_sin_x2.argtypes = [dp, dp, C.c_long]
    lib.fmh_atan_x2.argtypes = [dp, dp, C.c_long, C.c_int]
    lib.fmh_atan_any_x2.argtypes = [dp, dp, C.c_long, C.c_int]
    return lib


def _ptr(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ulps(got, ref):
    return np.abs(got - ref) / np.spacing(np.maximum(np.abs(ref), 1e-300))


def test_sincos_matches_libm(fm):
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-8, 8, 400000), rng.uniform(-1e4, 1e4, 100000), np.linspace(-np.pi, np.pi, 10001),
                        np.array([0.0, -0.0, 1e-300, np.pi / 4, np.pi / 2, 99999.9, 2e5, -3e7])])
    s, c = np.empty_like(x), np.empty_like(x)
    fm.fmh_sincos(_ptr(x), _ptr(s), _ptr(c), x.size)
    # absolute error <= 2 ulp of 1 everywhere (results near a zero of sin/cos are limited by the reduction)
    assert np.max(np.abs(s - np.sin(x))) < 3e-16
    assert np.max(np.abs(c - np.cos(x))) < 3e-16
    big = np.abs(np.sin(x)) > 0.1
    assert _ulps(s[big], np.sin(x[big])).max() <= 2.0
    assert np.all(s * s + c * c - 1.0 < 1e-15)
    s1 = np.empty_like(x)
    fm.fmh_sin(_ptr(x), _ptr(s1), x.size)
    assert np.array_equal(s1, s)  # fm_sin is the sine half of fm_sincos, bit for bit
    n = np.array([np.nan, np.inf])
    fm.fmh_sincos(_ptr(n), _ptr(s), _ptr(c), 2)
    assert np.isnan(s[:2]).all() and np.isnan(c[:2]).all()


def test_atan_matches_libm(fm):
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.uniform(-0.5, 0.5, 300000), rng.uniform(-1, 1, 200000), rng.uniform(-50, 50, 200000),
                        10.0 ** rng.uniform(-12, 12, 100000), np.array([0.0, -0.0, 0.5, 1.0, -1.0, 1e300, 1e308, np.inf, -np.inf])])
    y = np.empty_like(x)
    fm.fmh_atan(_ptr(x), _ptr(y), x.size)
    ref = np.arctan(x)
    ok = ref != 0
    assert _ulps(y[ok], ref[ok]).max() <= 2.0
    assert np.all(np.signbit(y) == np.signbit(x))
    n = np.array([np.nan])
    fm.fmh_atan(_ptr(n), _ptr(y), 1)
    assert np.isnan(y[0])


def test_paired_branch_free_cores(fm):
    """The *_x2 cores the model call uses on its fast path (ranges guaranteed by the caller)."""
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.uniform(-8, 8, 200000), rng.uniform(-1e5, 1e5, 100000), np.array([0.0, -0.0, np.pi / 2, -np.pi])])
    s, c, s1 = np.empty_like(x), np.empty_like(x), np.empty_like(x)
    fm.fmh_sincos_x2(_ptr(x), _ptr(s), _ptr(c), x.size)
    assert np.max(np.abs(s - np.sin(x))) < 3e-16 and np.max(np.abs(c - np.cos(x))) < 3e-16
    fm.fmh_sin_x2(_ptr(x), _ptr(s1), x.size)
    assert np.max(np.abs(s1 - np.sin(x))) < 3e-16
    big = np.abs(np.sin(x)) > 0.1
    assert _ulps(s1[big], np.sin(x[big])).max() <= 2.0
    for small, lim in ((1, 0.5), (0, 1.0)):
        a = np.concatenate([rng.uniform(-lim, lim, 300000), np.array([0.0, -0.0, lim, -lim, 1e-200, -1e-9])])
        y = np.empty_like(a)
        fm.fmh_atan_x2(_ptr(a), _ptr(y), a.size, small)
        ref = np.arctan(a)
        ok = ref != 0
        assert _ulps(y[ok], ref[ok]).max() <= 2.0
        nz = a != 0  # the branch-free form returns +0 for -0 (irrelevant to the model)
        assert np.all(np.signbit(y[nz]) == np.signbit(a[nz]))


def test_wide_path_atan_is_the_general_atan_and_the_fast_atan(fm):
    """fm_atan_any_x2 (the model call's wide path): bit-identical to the branchy fm_atan for every argument, and —
    with force_small — bit-identical to the fast path's fm_atan_nr_x2<true>, which is what lets a warp route its
    fast lanes through the wide path without changing their results."""
    rng = np.random.default_rng(4)
    x = np.concatenate([rng.uniform(-0.5, 0.5, 100000), rng.uniform(-1, 1, 100000), rng.uniform(-50, 50, 100000),
                        10.0 ** rng.uniform(-300, 300, 100000) * rng.choice([-1, 1], 100000),
                        np.array([0.0, -0.0, 0.5, -0.5, 1.0, -1.0, 1e300, -1e308, np.inf, -np.inf, 5e-324, 1.0000000000000002])])
    if x.size % 2:
        x = x[:-1]
    y_any, y_gen = np.empty_like(x), np.empty_like(x)
    fm.fmh_atan_any_x2(_ptr(x), _ptr(y_any), x.size, 0)
    fm.fmh_atan(_ptr(x), _ptr(y_gen), x.size)
    assert np.array_equal(y_any.view(np.uint64), y_gen.view(np.uint64))
    ref = np.arctan(x)
    ok = (ref != 0) & (np.abs(x) > 1e-300)
    assert _ulps(y_any[ok], ref[ok]).max() <= 2.0
    n = np.array([np.nan, 0.3])
    out = np.empty(2)
    fm.fmh_atan_any_x2(_ptr(n), _ptr(out), 2, 0)
    assert np.isnan(out[0]) and abs(out[1] - np.arctan(0.3)) < 1e-16
    a = np.concatenate([rng.uniform(-0.5, 0.5, 200000), np.array([0.0, -0.0, 0.5, -0.5, 1e-200, -1e-9])])
    y_forced, y_fast = np.empty_like(a), np.empty_like(a)
    fm.fmh_atan_any_x2(_ptr(a), _ptr(y_forced), a.size, 1)
    fm.fmh_atan_x2(_ptr(a), _ptr(y_fast), a.size, 1)
    assert np.array_equal(y_forced.view(np.uint64), y_fast.view(np.uint64))
