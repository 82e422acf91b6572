// imd_b200/csrc/tmm.cuh -- the reference's laser source: transfer-matrix solution of the Helmholtz
// equation over the FD cells (do_tmm, imd_tmm.c:193-445) with the wide-range permittivity of
// tmm_get_epsilon (:503-570), the collision integrals tmm_K1/K2 (:667-696) and the adaptive quadrature
// Runge5 / EE (:37-98).  Device functions for k_ttm_grid (ttm.cu); one call per diffusion sub-step.
//
// Complex arithmetic is written out with the formulas C99 / gcc use for finite operands (the product
// (a+bi)(c+di) = (ac-bd) + (ad+bc)i, the quotient by c^2+d^2 -- __divdc3 only rescales by powers of two
// around it -- and glibc's csqrt), so that the results agree with the reference to round-off.
#pragma once

#include "ttm_models.cuh"

#ifndef __CUDACC__                   /* host build of the header (tests/test_tmm_cpu.py) */
#include <cmath>
#endif

struct cplx { double re, im; };
TM_HD cplx cmk(double re, double im) { cplx z; z.re = re; z.im = im; return z; }
TM_HD cplx cadd(cplx a, cplx b) { return cmk(a.re + b.re, a.im + b.im); }
TM_HD cplx csub(cplx a, cplx b) { return cmk(a.re - b.re, a.im - b.im); }
TM_HD cplx cneg(cplx a) { return cmk(-a.re, -a.im); }
TM_HD cplx cmul(cplx a, cplx b) { return cmk(a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re); }
TM_HD cplx cscale(double s, cplx a) { return cmk(s * a.re, s * a.im); }
TM_HD cplx cdivc(cplx a, cplx b)
{
  const double r = prcp(b.re * b.re + b.im * b.im);          // one reciprocal for both parts (<= 1 ulp each)
  return cmk((a.re * b.re + a.im * b.im) * r, (a.im * b.re - a.re * b.im) * r);
}
TM_HD cplx crdiv(double s, cplx b)          // real / complex
{
  const double r = prcp(b.re * b.re + b.im * b.im);
  return cmk((s * b.re) * r, (-s * b.im) * r);
}
TM_HD double cabs2(cplx a) { return a.re * a.re + a.im * a.im; }
TM_HD cplx csqrtc(cplx z)                   // glibc csqrt, finite non-zero operands
{
  if (z.im == 0.0) return z.re < 0.0 ? cmk(0.0, copysign(sqrt(-z.re), z.im)) : cmk(fabs(sqrt(z.re)), copysign(0.0, z.im));
  const double d = psqrt(z.re * z.re + z.im * z.im);         // permittivities: no overflow to guard against
  double r, s;
  if (z.re > 0) { r = psqrt(0.5 * (d + z.re)); s = 0.5 * pdiv(z.im, r); }
  else { s = psqrt(0.5 * (d - z.re)); r = fabs(0.5 * pdiv(z.im, s)); }
  return cmk(r, copysign(s, z.im));
}
// exp(phi1) cos(phi2) + I exp(phi1) sin(phi2) with phi1 = creal(dx kl I) = -dx Im kl, phi2 = dx Re kl
TM_HD cplx tmm_eiphi(double dx, cplx kl)
{
  const double phi1 = -(dx * kl.im), phi2 = dx * kl.re;
  double sn, cs;
  sincos(phi2, &sn, &cs);
  const double e = pexp(phi1);
  return cmk(e * cs, e * sn);
}

struct TmmTables {                   // device pointers
  int nbb, nk;
  const double *eps_bb;              // [nbb][3]: lambda (um), Re, Im   (alu_eps_bb.dat)
  const double *K12;                 // [nk][2]                          (K12.dat)
};

// tmm_K1 (which = 0) / tmm_K2 (which = 1); (int) nu * 30 - 1 as written in the reference
TM_HD double tmm_K(const TmmTables &t, double nu, int which)
{
  if (nu < 1.0 / 30.0) return 1;
  if (nu < 50) {
    int i = (int) nu * 30 - 1;
    if (i < 0) i = 0;
    if (i > 1498) i = 1498;
    const double a = t.K12[2 * i + which], b = t.K12[2 * (i + 1) + which];
    return a + (30 * nu - i - 1) * (b - a);
  }
  const double n2 = nu * nu;
  if (which == 0) return 315.0 / 8.0 / n2 - 675675.0 / 64.0 / (n2 * n2);
  return 6.0 / n2 - 720.0 / (n2 * n2);
}

// tmm_get_epsilon (imd_tmm.c:503-570); `phase` = (int) EOS_phase_from_r_ti(Ni 26.9815 AMU, Ti [K]) from the caller.
// Helpers fermi_E / fermi_T / omega_pl / r0 / bMin / bMax / coulomb_log / numet / numax / nupl are the file-level
// ones of imd_ttm.c:2191-2308, written out; laser_active is 1 here (do_tmm has just set it).
TM_HD double tmm_fermi_E(double ne)
{
  return pdiv(TMC_HBAR * TMC_HBAR * ppow(3.0 * TMC_PI * TMC_PI * ne, 2.0 / 3.0) / 2.0, TMC_EMASS);
}
// the interband permittivity at the laser wavelength: table walk and linear interpolation of
// tmm_get_epsilon (:524-527, :547), the same in every call of a run
TM_HD cplx tmm_eps_bb(const TmmTables &t, double lambda)
{
  int il = 0, ir = 0;
  for (int id = 0; lambda * 1e6 > t.eps_bb[3 * id]; id++) { il = id; ir = id + 1; }
  const cplx ebl = cmk(t.eps_bb[3 * il + 1], t.eps_bb[3 * il + 2]), ebr = cmk(t.eps_bb[3 * ir + 1], t.eps_bb[3 * ir + 2]);
  const double dl = t.eps_bb[3 * ir] - t.eps_bb[3 * il], xl = lambda * 1e6 - t.eps_bb[3 * il];
  cplx d = csub(ebr, ebl);
  d = cmk(d.re / dl, d.im / dl);
  return cadd(ebl, cmk(d.re * xl, d.im * xl));
}
// epsilon_bb: tmm_eps_bb, or 0 where the phase table says 3, 4 or 5 (:549-552).  Divisions, roots, pow/log/exp
// are the branch-free versions of ttm_math.cuh (<= 1-2 ulp), like in the wide-range models of do_FILLMESH.
TM_HD cplx tmm_get_epsilon(const TmmTables &t, double lambda, double Te, double Ti, double Z, double Ne,
                                       cplx epsilon_bb)
{
  Te *= 11604.52500617; Ti *= 11604.52500617;
  const double omega_las = (2.0 * TMC_PI * 2.998e8 / lambda);                    // LIGHTSPEED, globals.h:325
  const double Ncr = TMC_ECONST * TMC_EMASS * omega_las * omega_las / TMC_ECHARGE / TMC_ECHARGE;
  const double Ni = pdiv(Ne, Z);
  const double EF = tmm_fermi_E(Ne);
  const double TF = pdiv(2.0 * EF, 3.0 * TMC_BOLTZMAN);
  const double VF = psqrt(pdiv(2 * EF, TMC_EMASS));
  const double A1p = 4.41, A2p = 0.8, A3p = 0.7, A4p = 0.2;
  const double nu_met = 1.309202957843054e+11 * (A1p * Ti + pdiv(A2p * Te * Te, TF));
  const double r0n = ppow(pdiv(3.0, 4 * TMC_PI * Ni), 1.0 / 3.0);
  const double kTe = TMC_BOLTZMAN * Te;
  const double nu_max = pdiv(A3p, r0n) * psqrt(VF * VF + pdiv(kTe, TMC_EMASS));
  // coulomb_log(omega_las, Z, Ni, Te) with laser_active
  const double ft = pdiv(2.0 * tmm_fermi_E(Ni * Z), 3.0 * TMC_BOLTZMAN);
  const double teff = psqrt(Te * Te + ft * ft);
  const double wpl = psqrt(pdiv(pdiv(Ni * Z * TMC_ECHARGE * TMC_ECHARGE, TMC_EMASS), TMC_ECONST));
  const double bmax = TM_MAX(r0n, pdiv(psqrt(pdiv(TMC_BOLTZMAN * teff, TMC_EMASS)), TM_MAX(omega_las, wpl)));
  const double bmin = TM_MAX(pdiv(Z * TMC_ECHARGE * TMC_ECHARGE, kTe), pdiv(TMC_HBAR, psqrt(2 * TMC_EMASS * kTe)));
  const double coulomb = TM_MAX(2.0, plog(pdiv(bmax, bmin)));
  const double nu_pl = pdiv(1.863839656495274e-40 * Z * Ne * coulomb, kTe * psqrt(kTe));      // pow(kTe, 1.5)
  const double ksi = pdiv(0.75 * sqrt(TMC_PI) * nu_pl, omega_las);
  const double nn = pdiv(Ne, Ncr);
  // epsilon_met = epsilon_bb + 1 - Ne/Ncr / (1 + i MIN(nu_met, nu_max)/omega_las)
  const cplx met = csub(cadd(epsilon_bb, cmk(1.0, 0.0)), crdiv(nn, cmk(1.0, pdiv(TM_MIN(nu_met, nu_max), omega_las))));
  // epsilon_pl = 1 - Ne/Ncr (K1(ksi) - i nu_pl/omega_las K2(ksi))
  const cplx pl = csub(cmk(1.0, 0.0), cscale(nn, cmk(tmm_K(t, ksi, 0), -(pdiv(nu_pl, omega_las) * tmm_K(t, ksi, 1)))));
  const double w = pexp(pdiv(-A4p * Te, TF));
  return cadd(pl, cscale(w, csub(met, pl)));
}

// EE (imd_tmm.c:72-98): |Bplus e^{i k z delta} + Bminus e^{-i k z delta}|^2
TM_HD double tmm_EE(double z, double delta, cplx kl, cplx Bplus, cplx Bminus)
{
  const cplx e = tmm_eiphi(delta * z, kl);
  const cplx Ez = cadd(cmul(Bplus, e), cdivc(Bminus, e));
  return Ez.re * Ez.re + Ez.im * Ez.im;
}

// Runge5 (imd_tmm.c:37-69)
TM_HD double tmm_runge5(double delta, cplx kl, cplx Bplus, cplx Bminus)
{
  double dx = 1, cur_xpos = 0, result = 0;
  const double errval = 1e-5;
  while (cur_xpos < 1 && dx > 1e-5) {
    const double k1 = dx / 3.0 * tmm_EE(cur_xpos, delta, kl, Bplus, Bminus);
    const double k3 = dx / 3.0 * tmm_EE(cur_xpos + dx / 3.0, delta, kl, Bplus, Bminus);
    const double k4 = dx / 3.0 * tmm_EE(cur_xpos + 0.5 * dx, delta, kl, Bplus, Bminus);
    const double k5 = dx / 3.0 * tmm_EE(cur_xpos + dx, delta, kl, Bplus, Bminus);
    const double ERR = k1 - 4.5 * k3 + 4 * k4 - 0.5 * k5;
    if (ERR < 5.0 / 32 * errval) { result += 0.5 * (k1 + 4 * k4 + k5); cur_xpos += dx; dx *= 1.1; }
    else if (ERR > 5 * errval) { dx *= 0.5; continue; }
    else { result += 0.5 * (k1 + 4 * k4 + k5); cur_xpos += dx; }
    if (cur_xpos + dx > 1) dx = 1 - cur_xpos;
  }
  return result;
}

// tmpmat = Cinv * Pmat of one interface (imd_tmm.c:299-313): M = {m0, m1, m2, m3}, column-major like matmul
TM_HD void tmm_interface(cplx e, cplx kl, cplx kr, cplx M[4])      // e = tmm_eiphi(dx, kl)
{
  const cplx P0 = e, P2 = crdiv(1.0, e), P1 = e, P3 = cneg(P2);
  const cplx C0 = cmk(0.5, 0.0), C1 = cmk(0.5, 0.0);
  const cplx C2 = cdivc(cscale(0.5, kl), kr), C3 = cneg(C2);
  M[0] = cadd(cmul(C0, P0), cmul(C2, P1));
  M[1] = cadd(cmul(C1, P0), cmul(C3, P1));
  M[2] = cadd(cmul(C0, P2), cmul(C2, P3));
  M[3] = cadd(cmul(C1, P2), cmul(C3, P3));
}
