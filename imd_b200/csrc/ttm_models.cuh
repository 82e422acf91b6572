// imd_b200/csrc/ttm_models.cuh -- the wide-range models of do_FILLMESH for ONE FD cell:
// MeanCharge (imd_ttm.c:2241-2280), getKappa (:2356-2399), getGamma (:2310-2335) and their helpers
// fermi_E/fermi_T/omega_pl/r0/bMin/bMax/coulomb_log/numet/numax/nupl (:2191-2308), fused.
//
// k_ttm_grid is a latency chain (very few threads, >= 250 sub-steps per MD step, every sub-step a
// sequence of dependent transcendental calls per cell), so the formulation minimises the NUMBER of
// dependent log/exp evaluations per sub-step.  The reference makes 22 of them per cell; here 13:
//   * what depends only on the cell's density is evaluated once per MD step (wide_range_inv): density
//     does not change during the sub-step loop (imd_ttm.c:111-121 never touches dens).  That is
//     log(rho/(z0 am)) of MeanCharge, Ni = Ne/Z = rho/(am AMU), r0(Ni) and the two A3/r0 quotients;
//   * fermi_T(Ni*Z) is fermi_T(Ne): Ni*Z = (Ne/Z)*Z;
//   * pow(pow(s, 1/yc), YBE) = exp((YBE/yc) log s) and pow(ya exp(yb lr0), yc) = exp(yc (log ya + yb lr0))
//     in MeanCharge.
// These are identities in exact arithmetic; in FP64 they move the result by a few ulp (the reference's
// own pow() already carries |y log x| ulp), far inside the 1e-10 parity tolerance of the FD-cell
// fields: tests/test_ttm_models_cpu.py compares this header, compiled for the host, with the test suite's
// line-by-line CPU restatement of the reference functions over the whole argument range (<= 5e-14, kappa <= 1e-12),
// tests/test_ttm_gpu.py compares the lattice with dumps of the reference itself.
#pragma once

#include "ttm_math.cuh"

#define TMC_ECONST   8.841941282883075e-12
#define TMC_HBAR     1.054571800139113e-34
#define TMC_EMASS    9.109383560000000e-31
#define TMC_BOLTZMAN 1.380648520000000e-23
#define TMC_ECHARGE  1.602176620000000e-19
#define TMC_AMU      1.660539040200000e-27
#define TMC_PI       3.14159265358979323846

#define TM_MAX(a, b) ((a) > (b) ? (a) : (b))
#define TM_MIN(a, b) ((a) < (b) ? (a) : (b))

struct WideRange { double Z, ne, kappa, gamma; };

// per cell and MD step: everything that depends on the density alone
struct WideRangeInv { double lr0, ni, r0n, g_r0n, t_r0n; };

TM_HD WideRangeInv wide_range_inv(double dens, double z0, double am)
{
  WideRangeInv v;
  v.lr0 = plog(pdiv(dens * 1e-3, z0 * am));                            // MeanCharge: r0 = rho 1e-3/(z0 am)
  v.ni = pdiv(dens, am * TMC_AMU);                                      // Ne/Z
  v.r0n = ppow(pdiv(3.0, 4 * TMC_PI * v.ni), 1.0 / 3.0);                // r0(Ni)
  v.g_r0n = pdiv(0.25, v.r0n);                                          // getGamma's A3/r0
  v.t_r0n = pdiv(0.16, v.r0n);                                          // getKappa's A3/r0
  return v;
}

// z0p43 = pow(z0, 4/3) (MeanCharge), hoisted by the caller
TM_HD WideRange wide_range(double temp_ev, double ti_ev, const WideRangeInv &v, double z0, double z0p43,
                           double omega_las, int la)
{
  WideRange w;
  // ---- MeanCharge(temp [K], rho, z0, am)
  {
    double temp = temp_ev * 11604.5;
    temp *= 1e-3;
    const double YA1 = 0.003323, YA2 = 0.971832, YA3 = 0.926148e-5, YA4 = 3.10165, YB0 = -1.7630, YB1 = 1.43175,
                 YB2 = 0.315463, YC1 = 0.036666667, YC2 = 0.983333333, YAL = 0.143139e2, YBE = 0.6624e0;
    const double t0 = pdiv(temp, z0p43);
    const double lt0 = plog(t0);
    const double ytf = pdiv(t0, 1.0 + t0);
    const double ya = YA1 * pexp(YA2 * lt0) + YA3 * pexp(YA4 * lt0);
    const double yb = -pexp(YB0 + YB1 * ytf + YB2 * ytf * ytf);
    const double yc = YC1 * ytf + YC2;
    const double lq1 = plog(ya) + yb * v.lr0;                           // log(ya exp(yb lr0))
    const double s = pexp(yc * v.lr0) + pexp(yc * lq1);                 // r0^yc + yq1^yc
    const double yx = YAL * pexp(pdiv(YBE, yc) * plog(s));              // YAL (s^(1/yc))^YBE
    w.Z = pdiv(z0 * yx, 1.0 + yx + psqrt(1.0 + 2.0 * yx));
  }
  const double Z = w.Z;
  const double Ne = w.ne = Z * v.ni;
  // ---- shared by getKappa and getGamma (both start with Te *= 11604.5; Ti *= 11604.5)
  const double Te = temp_ev * 11604.5, Ti = ti_ev * 11604.5;
  const double EF = pdiv(TMC_HBAR * TMC_HBAR * ppow(3.0 * TMC_PI * TMC_PI * Ne, 2.0 / 3.0) / 2.0, TMC_EMASS);   // fermi_E(Ne)
  const double TF = 4.828648689433765e+22 * EF;
  const double VF = 1.481734876966785e+15 * psqrt(EF);
  const double kTe = TMC_BOLTZMAN * Te;
  const double kTe_m = pdiv(kTe, TMC_EMASS);
  const double vth = psqrt(VF * VF + kTe_m);                                                                    // numax's root
  // coulomb_log(omega_las, Z, Ni, Te, la) = MAX(2, log(bMax/bMin))
  const double ft = pdiv(2.0 * EF, 3.0 * TMC_BOLTZMAN);                                                         // fermi_T(Ni Z)
  const double teff = psqrt(Te * Te + ft * ft);
  const double wpl = psqrt(pdiv(pdiv(Ne * TMC_ECHARGE * TMC_ECHARGE, TMC_EMASS), TMC_ECONST));                  // omega_pl(Ni Z)
  const double vte = psqrt(pdiv(TMC_BOLTZMAN * teff, TMC_EMASS));
  const double bmax = la ? TM_MAX(v.r0n, pdiv(vte, TM_MAX(omega_las, wpl))) : TM_MAX(v.r0n, pdiv(vte, wpl));
  const double bmin = TM_MAX(pdiv(Z * TMC_ECHARGE * TMC_ECHARGE, kTe), pdiv(TMC_HBAR, psqrt(2 * TMC_EMASS * TMC_BOLTZMAN * Te)));
  const double coulomb = TM_MAX(2.0, plog(pdiv(bmax, bmin)));
  const double skTe = psqrt(kTe);
  const double kTe15 = kTe * skTe, kTe25 = kTe * kTe * skTe;                                                    // pow(kTe,1.5), pow(kTe,2.5)
  const double te_tf = pdiv(Te, TF), te2_tf = Te * te_tf;                                                        // Te/TF, Te^2/TF
  // ---- getGamma
  {
    const double A1g = 50, A2g = 20, m_atom = 26.9815;
    const double numet = 1.309202957843054e+11 * (A1g * Ti + A2g * te2_tf);
    const double numax = v.g_r0n * vth;
    const double nupl = pdiv(1.863839656495274e-40 * Z * Ne * coulomb, kTe15);
    const double nu_eff = TM_MIN(numet, TM_MIN(numax, nupl));
    w.gamma = pdiv(2.783313120645740e-74, m_atom * TMC_AMU) * Ne * nu_eff;
  }
  // ---- getKappa
  {
    const double A1t = 2.95, A2t = 0.5, A4t = 1.2;
    const double numet = 1.309202957843054e+11 * (A1t * Ti + A2t * te2_tf);
    const double numax = v.t_r0n * vth;
    const double nu_eff = TM_MIN(numet, numax);
    const double kappa_met = pdiv(6.884236239621913e-16 * Ne * Te, nu_eff);
    const double kappa_pl = pdiv(pdiv(4.428788911416808e+43 * kTe25, Z), coulomb);
    const double kappa_wr = kappa_pl + (kappa_met - kappa_pl) * pexp(-A4t * te_tf);
    w.kappa = kappa_wr * 0.073768115942029;
  }
  return w;
}
