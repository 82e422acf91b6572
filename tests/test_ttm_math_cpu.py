"""Host-side accuracy check of imd_b200/csrc/ttm_math.cuh (the branch-free log/exp/pow the FD-lattice
kernel uses for the wide-range models, imd_ttm.c:2191-2399).  The header is __host__ __device__:
compiled here with g++ (-ffp-contract=off, fma spelled out exactly as in the device build) and
compared with libm over the argument ranges do_FILLMESH produces."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r"""
#include "%s/imd_b200/csrc/ttm_math.cuh"
extern "C" void v_plog(int n, const double *x, double *o) { for (int i = 0; i < n; i++) o[i] = plog(x[i]); }
extern "C" void v_pexp(int n, const double *x, double *o) { for (int i = 0; i < n; i++) o[i] = pexp(x[i]); }
extern "C" void v_ppow(int n, const double *x, const double *y, double *o) { for (int i = 0; i < n; i++) o[i] = ppow(x[i], y[i]); }
""" % ROOT


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("ttm_math")
    src = d / "t.cpp"
    src.write_text(SRC)
    so = d / "libt.so"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", str(so), str(src)], check=True)
    return ctypes.CDLL(str(so))


def _dp(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _ulps(got, ref):
    return np.abs(got - ref) / np.spacing(np.abs(ref))


def test_plog_within_2_ulp(lib):
    rng = np.random.default_rng(1)
    n = 400_000
    for x in (np.exp(rng.uniform(-300, 300, n)), rng.uniform(0.5, 2.0, n), 1 + rng.uniform(-1e-6, 1e-6, n)):
        o = np.empty(n)
        lib.v_plog(n, _dp(x), _dp(o))
        assert _ulps(o, np.log(x)).max() <= 2.0


def test_pexp_within_1_ulp(lib):
    rng = np.random.default_rng(2)
    n = 400_000
    x = rng.uniform(-700, 700, n)
    o = np.empty(n)
    lib.v_pexp(n, _dp(x), _dp(o))
    assert _ulps(o, np.exp(x)).max() <= 1.0


def test_ppow_relative_error(lib):
    """rel. error <= 2^-51 (1 + |y log x|): 1e-14 class for the wide-range model arguments."""
    rng = np.random.default_rng(3)
    n = 400_000
    x = np.exp(rng.uniform(-80, 80, n))
    y = rng.uniform(-4, 4, n)
    o = np.empty(n)
    lib.v_ppow(n, _dp(x), _dp(y), _dp(o))
    ref = np.power(x, y)
    bound = 2.0 ** -51 * (1 + np.abs(y * np.log(x)))
    assert (np.abs(o - ref) / ref <= bound).all()
    # the constant exponents of the models
    for ex in (2.0 / 3.0, 1.0 / 3.0, 0.971832, 3.10165, 0.6624):
        x = np.exp(rng.uniform(-20, 70, n))
        yy = np.full(n, ex)
        lib.v_ppow(n, _dp(x), _dp(yy), _dp(o))
        ref = np.power(x, ex)
        assert (np.abs(o - ref) / ref <= 2.0 ** -51 * (1 + np.abs(ex * np.log(x)))).all()
        assert (np.abs(o - ref) / ref).max() < 1e-13          # |y log x| up to ~220 here


def test_out_of_domain_is_nan_not_garbage(lib):
    x = np.array([0.0, -1.0, np.inf, np.nan, 5e-324])
    o = np.empty(5)
    lib.v_plog(5, _dp(x), _dp(o))
    assert np.isnan(o).all()
    x = np.array([np.nan])
    o = np.empty(1)
    lib.v_pexp(1, _dp(x), _dp(o))
    assert np.isnan(o).all()
