"""Host-side check of imd_b200/csrc/tmm.cuh, the device functions of the transfer-matrix laser source
(do_tmm): the header is __host__ __device__, compiled here with g++ and compared with the oracle's
restatement of imd_tmm.c (C99 complex arithmetic, libm) -- which is itself pinned to the reference's own
do_tmm (tests/test_ttm_oracle_vs_reference.py).  Covers the permittivity, the interface matrix, csqrt and
the adaptive quadrature; the whole source on the device is compared with the reference in test_ttm_gpu.py."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from imd_b200 import synth
from oracle import binding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SRC = r"""
#include <cmath>
#include "%s/imd_b200/csrc/tmm.cuh"
extern "C" void v_eps(int n, int nbb, const double *bb, int nk, const double *k12, double lambda, const double *te,
                      const double *ti, const double *z, const double *ne, double *out)
{
  TmmTables t; t.nbb = nbb; t.nk = nk; t.eps_bb = bb; t.K12 = k12;
  const cplx ebb = tmm_eps_bb(t, lambda);
  for (int i = 0; i < n; i++) {
    const cplx e = tmm_get_epsilon(t, lambda, te[i], ti[i], z[i], ne[i], ebb);
    out[2 * i] = e.re; out[2 * i + 1] = e.im;
  }
}
extern "C" void v_interface(double dx, const double *kl, const double *kr, double *out)
{
  cplx M[4];
  const cplx k = cmk(kl[0], kl[1]);
  tmm_interface(tmm_eiphi(dx, k), k, cmk(kr[0], kr[1]), M);
  for (int j = 0; j < 4; j++) { out[2 * j] = M[j].re; out[2 * j + 1] = M[j].im; }
}
extern "C" void v_csqrt(const double *z, double *out) { const cplx r = csqrtc(cmk(z[0], z[1])); out[0] = r.re; out[1] = r.im; }
extern "C" double v_runge5(double delta, const double *kl, const double *bp, const double *bm)
{
  return tmm_runge5(delta, cmk(kl[0], kl[1]), cmk(bp[0], bp[1]), cmk(bm[0], bm[1]));
}
""" % ROOT

dp = C.POINTER(C.c_double)


def _dp(a):
    return a.ctypes.data_as(dp)


@pytest.fixture(scope="module")
def lib(tmp_path_factory):
    d = tmp_path_factory.mktemp("tmm")
    src = d / "t.cpp"
    src.write_text(SRC)
    so = d / "libt.so"
    subprocess.run(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-o", str(so), str(src)], check=True)
    L = C.CDLL(str(so))
    L.v_runge5.restype = C.c_double
    L.v_runge5.argtypes = [C.c_double, dp, dp, dp]
    L.v_interface.argtypes = [C.c_double, dp, dp, dp]
    L.v_csqrt.argtypes = [dp, dp]
    L.v_eps.argtypes = [C.c_int, C.c_int, dp, C.c_int, dp, C.c_double, dp, dp, dp, dp, dp]
    return L


@pytest.fixture(scope="module")
def orc():
    tp = synth.TtmParams(ttmdimx=8, use_tmm=1, tmm_threshold=5.0)
    eos = synth.make_eos_tables()
    t = binding.TtmOracle(tp, eos, np.diag([200.0, 20.0, 20.0]), 39.69, 0.1, 1)
    O = t.L
    O.orc_tmm_probe_epsilon.argtypes = [C.c_void_p, C.c_double, C.c_double, C.c_double, C.c_double, dp]
    O.orc_tmm_probe_runge5.restype = C.c_double
    O.orc_tmm_probe_runge5.argtypes = [C.c_double, dp, dp, dp]
    O.orc_tmm_probe_interface.argtypes = [C.c_double, dp, dp, dp]
    O.orc_tmm_probe_csqrt.argtypes = [dp, dp]
    yield t
    t.close()


def test_permittivity(lib, orc):
    rng = np.random.default_rng(3)
    n = 4000
    te = np.exp(rng.uniform(np.log(0.02), np.log(30.0), n))            # eV
    ti = np.exp(rng.uniform(np.log(0.01), np.log(1.5), n))
    z = rng.uniform(2.0, 11.0, n)
    ne = z * np.exp(rng.uniform(np.log(5e27), np.log(8e28), n))         # 1/m^3
    tabs = synth.make_tmm_tables()
    bb, k12 = np.ascontiguousarray(tabs["eps_bb"]), np.ascontiguousarray(tabs["K12"])
    got = np.zeros((n, 2))
    lib.v_eps(n, bb.shape[0], _dp(bb), k12.shape[0], _dp(k12), 800e-9, _dp(te), _dp(ti), _dp(z), _dp(ne), _dp(got))
    ref = np.zeros((n, 2))
    o2 = np.zeros(2)
    for i in range(n):
        orc.L.orc_tmm_probe_epsilon(orc.t, te[i], ti[i], z[i], ne[i], _dp(o2))
        ref[i] = o2
    scale = np.abs(ref[:, 0] + 1j * ref[:, 1])
    assert np.max(np.abs(got - ref).max(axis=1) / scale) < 1e-12


def test_interface_matrix_csqrt_and_quadrature(lib, orc):
    rng = np.random.default_rng(5)
    k0 = 2 * np.pi / 800e-9 / 1e10
    for _ in range(300):
        eps = np.array([rng.uniform(-80, 2), rng.uniform(0.0, 60)])
        r1, r2 = np.zeros(2), np.zeros(2)
        lib.v_csqrt(_dp(eps), _dp(r1)); orc.L.orc_tmm_probe_csqrt(_dp(eps), _dp(r2))
        assert np.abs(r1 - r2).max() <= 4e-16 * np.abs(r2).max()
        kl, kr = k0 * r2, k0 * np.array([rng.uniform(0.2, 3), rng.uniform(0.0, 9)])
        dx = rng.uniform(8.0, 40.0)
        m1, m2 = np.zeros(8), np.zeros(8)
        lib.v_interface(dx, _dp(kl), _dp(kr), _dp(m1)); orc.L.orc_tmm_probe_interface(dx, _dp(kl), _dp(kr), _dp(m2))
        assert np.abs(m1 - m2).max() <= 1e-13 * np.abs(m2).max()
        bp = rng.normal(0, 1, 2) * 0.2
        bm = rng.normal(0, 1, 2) * 0.2
        a = lib.v_runge5(dx, _dp(kl), _dp(bp), _dp(bm)); b = orc.L.orc_tmm_probe_runge5(dx, _dp(kl), _dp(bp), _dp(bm))
        assert abs(a - b) <= 1e-12 * abs(b) + 1e-300
    # the vacuum cell in front of the sample: dx = 1e10 A, a phase of 7.85e6 rad
    kl = np.array([k0, 0.0]); kr = k0 * np.array([0.7, 6.4])
    m1, m2 = np.zeros(8), np.zeros(8)
    lib.v_interface(1e10, _dp(kl), _dp(kr), _dp(m1)); orc.L.orc_tmm_probe_interface(1e10, _dp(kl), _dp(kr), _dp(m2))
    assert np.abs(m1 - m2).max() <= 1e-13 * np.abs(m2).max()
