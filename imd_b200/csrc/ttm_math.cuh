// imd_b200/csrc/ttm_math.cuh -- branch-free FP64 log / exp / pow for the FD-lattice kernel.
//
// The wide-range models of do_FILLMESH (imd_ttm.c:2191-2399) are a chain of ~18 pow() calls per
// FD cell and sub-step, evaluated by very few threads: k_ttm_grid is a latency chain, not a
// throughput kernel.  libdevice's pow() costs ~250 dependent instructions and its special-case
// branches stop ptxas from interleaving independent calls.  These versions are straight-line
// code (every operation explicit, fma() spelled out because ttm.cu is compiled -fmad=false), so
// the independent calls of one cell are scheduled into each other's latency shadows.
//
// Domain: x > 0, finite and normal; |y log x| < 700.  Outside it the result is NaN (log) or
// saturates (exp) -- do_FILLMESH then raises the same "NaN" errors as the reference.
// Accuracy (checked on the host against libm, tests/test_ttm_math_cpu.py): pexp <= 1 ulp, plog <= 2 ulp
// on their own; ppow(x, y) relative error <= 2^-51 * (1 + |y log x|), i.e. ~1e-14 for the
// arguments met here -- far inside the 1e-10 parity tolerance of the FD-cell fields.
#pragma once

#if defined(__CUDACC__)
#define TM_HD __host__ __device__ __forceinline__
#else
#define TM_HD static inline
#endif

#if defined(__CUDA_ARCH__)
#define TM_D2LL(x) __double_as_longlong(x)
#define TM_LL2D(x) __longlong_as_double(x)
#define TM_FMA(a, b, c) fma((a), (b), (c))
#define TM_RINT(x) rint(x)
#else
#include <cmath>
#include <cstring>
static inline long long TM_D2LL(double x) { long long r; memcpy(&r, &x, 8); return r; }
static inline double TM_LL2D(long long x) { double r; memcpy(&r, &x, 8); return r; }
#define TM_FMA(a, b, c) std::fma((a), (b), (c))
#define TM_RINT(x) std::nearbyint(x)
#endif

// 1/d, a/d and sqrt(x) without the IEEE slow paths (the hardware approximation refined by Newton
// steps, a final residual correction): ~10 straight-line instructions instead of the ~35 of the
// library routines with their special-case calls.  Arguments must be normal and finite with a
// normal result (true for every use in the wide-range models); results agree with the IEEE
// operations to <= 1 ulp (tests/test_ttm_gpu.py::test_device_math).  Host build: the plain operators.
TM_HD double prcp(double d)
{
#if defined(__CUDA_ARCH__)
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double e = TM_FMA(-d, y, 1.0);               // the approximation carries ~20 bits: 40 after one step, 80 after two
  y = TM_FMA(y, e, y);
  e = TM_FMA(-d, y, 1.0);
  return TM_FMA(y, e, y);
#else
  return 1.0 / d;
#endif
}

TM_HD double pdiv(double a, double d)
{
#if defined(__CUDA_ARCH__)
  const double y = prcp(d);
  const double q = a * y;
  const double r = TM_FMA(-q, d, a);
  return TM_FMA(r, y, q);
#else
  return a / d;
#endif
}

TM_HD double psqrt(double x)
{
#if defined(__CUDA_ARCH__)
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double g = x * y, h = 0.5 * y;
  double r = TM_FMA(-h, g, 0.5);               // coupled Newton steps for g -> sqrt(x), h -> 1/(2 sqrt(x))
  g = TM_FMA(g, r, g); h = TM_FMA(h, r, h);
  r = TM_FMA(-h, g, 0.5);
  g = TM_FMA(g, r, g); h = TM_FMA(h, r, h);
  const double d = TM_FMA(-g, g, x);
  return TM_FMA(d, h, g);
#else
  return std::sqrt(x);
#endif
}

// natural logarithm, x normal positive.  m in [sqrt(1/2), sqrt(2)), f = (m-1)/(m+1),
// log m = 2 atanh f = 2f + 2f^3/3 + ... ; the quotient is refined once so that f carries
// (almost) no rounding error of its own.
TM_HD double plog(double x)
{
  const long long ix = TM_D2LL(x);
  int e = (int) (ix >> 52) - 1023;
  double m = TM_LL2D((ix & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);       // [1, 2)
  const bool up = m > 1.4142135623730951;
  m = up ? m * 0.5 : m;
  e = up ? e + 1 : e;
  const double num = m - 1.0, den = m + 1.0;                                       // num exact
  const double rden = prcp(den);                                                   // den in [1.7, 2.42)
  double f = num * rden;
  // residual of the quotient: num - f*den, with den = m + 1 exact to 1 rounding (err <= 2^-53)
  const double r = TM_FMA(-f, den, num);
  const double fl = r * rden;                                                      // tail of f
  const double f2 = f * f;
  // sum_{k=0..10} 2/(2k+3) f2^k by Estrin's scheme: dependent depth 4 instead of 11 (the kernel is a
  // latency chain; the independent partial sums fill each other's latency shadows)
  const double z2 = f2 * f2, z4 = z2 * z2, z8 = z4 * z4;
  const double a0 = TM_FMA(2.0 / 5.0, f2, 2.0 / 3.0), a1 = TM_FMA(2.0 / 9.0, f2, 2.0 / 7.0);
  const double a2 = TM_FMA(2.0 / 13.0, f2, 2.0 / 11.0), a3 = TM_FMA(2.0 / 17.0, f2, 2.0 / 15.0);
  const double a4 = TM_FMA(2.0 / 21.0, f2, 2.0 / 19.0);
  const double b0 = TM_FMA(a1, z2, a0), b1 = TM_FMA(a3, z2, a2), b2 = TM_FMA(2.0 / 23.0, z2, a4);
  const double p = TM_FMA(b2, z8, TM_FMA(b1, z4, b0));
  const double de = (double) e;
  // log x = e ln2_hi + 2f + (f f2 p + 2 fl + e ln2_lo)
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
  const double tail = TM_FMA(f * f2, p, TM_FMA(de, ln2_lo, 2.0 * fl));
  const double res = TM_FMA(de, ln2_hi, 2.0 * f) + tail;
  const bool ok = (ix > 0x000fffffffffffffLL) && (ix < 0x7ff0000000000000LL);   // positive, normal, finite
  return ok ? res : TM_LL2D(0x7ff8000000000000LL);
}

// e^x, |x| < 700 (clamped beyond: the models never get there with finite inputs)
TM_HD double pexp(double x)
{
  const bool isn = !(x == x);
  x = x < -700.0 ? -700.0 : x;
  x = x > 700.0 ? 700.0 : x;
  const double k = TM_RINT(x * 1.4426950408889634);
  const double ln2_hi = 6.93147180369123816490e-01, ln2_lo = 1.90821492927058770002e-10;
  double r = TM_FMA(-k, ln2_hi, x);
  r = TM_FMA(-k, ln2_lo, r);                                                       // |r| <= 0.3466
  // e^r = 1 + (r + r^2 Q(r)), Q = sum_{k=0..12} r^k/(k+2)! by Estrin's scheme (depth 5 instead of 14)
  const double r2 = r * r, r4 = r2 * r2, r8 = r4 * r4;
  const double a0 = TM_FMA(1.0 / 6.0, r, 0.5), a1 = TM_FMA(1.0 / 120.0, r, 1.0 / 24.0);
  const double a2 = TM_FMA(1.0 / 5040.0, r, 1.0 / 720.0), a3 = TM_FMA(1.0 / 362880.0, r, 1.0 / 40320.0);
  const double a4 = TM_FMA(1.0 / 39916800.0, r, 1.0 / 3628800.0), a5 = TM_FMA(1.0 / 6227020800.0, r, 1.0 / 479001600.0);
  const double b0 = TM_FMA(a1, r2, a0), b1 = TM_FMA(a3, r2, a2), b2 = TM_FMA(a5, r2, a4);
  const double d0 = TM_FMA(b1, r4, b0), d1 = TM_FMA(1.0 / 87178291200.0, r4, b2);
  const double q = TM_FMA(d1, r8, d0);
  const double p = TM_FMA(r2, q, r) + 1.0;
  // scale by 2^k in two exact steps (k in [-1010, 1010])
  const int ki = (int) k;
  const int k1 = ki / 2, k2 = ki - k1;
  const double s1 = TM_LL2D((long long) (k1 + 1023) << 52), s2 = TM_LL2D((long long) (k2 + 1023) << 52);
  const double res = p * s1 * s2;
  return isn ? x : res;
}

// x^y for x > 0
TM_HD double ppow(double x, double y) { return pexp(y * plog(x)); }
