// pow(x, y) and log(x) for the Exner function and the hydrostatic integral (calc_ph, calc_t, calc_gz_lev:
// operators_mod.F90:136-190,433-469), restricted to what those loops feed them: x a positive normal number.
//
// k_ph and k_gz_col are bound by instruction issue, not by bytes (ncu, profiles/r02_ncu_full_summary_*.txt: 230 and 237
// instructions per point, a third of them FP64, 11-16 % UMOV materialising the library routine's coefficients); the
// general pow() spends more than half of its instructions on cases these kernels never produce (negative bases,
// integer exponents, infinities, subnormals) and on loading 64-bit immediates.  The routines below keep the
// double-double core and nothing else; the coefficients sit in constant memory and enter the FMAs as operands.
//
// Accuracy (tests/test_fastmath_cpu.py, against libquadmath on the CPU): the SAME source compiles for the host, and
// every operation in it is an IEEE-754 operation with one rounding (+, -, *, correctly rounded 1/x, fma, rint), so the
// host results are bit for bit the device results and the measured bounds carry over: pow_pos < 1 ulp, log_pos < 1 ulp.
// The library routines they replace are documented at 2 ulp (pow) and 1 ulp (log); the CPU checker uses glibc's.
// Outside the domain (x not a positive normal number, |y ln x| > 700) they defer to pow() / log().
#pragma once
#include <cmath>
#include <cstring>

#if defined(__CUDACC__)
#define FM_HD __host__ __device__ __forceinline__
#else
#define FM_HD inline
#endif

// 2/(2k+1), k = 1..11 (atanh series of ln m), then 1/k!, k = 2..13 (Taylor series of exp)
#define FM_COEFFS { \
  2.0 / 3.0, 2.0 / 5.0, 2.0 / 7.0, 2.0 / 9.0, 2.0 / 11.0, 2.0 / 13.0, 2.0 / 15.0, 2.0 / 17.0, 2.0 / 19.0, 2.0 / 21.0, 2.0 / 23.0, \
  1.0 / 2.0, 1.0 / 6.0, 1.0 / 24.0, 1.0 / 120.0, 1.0 / 720.0, 1.0 / 5040.0, 1.0 / 40320.0, 1.0 / 362880.0, 1.0 / 3628800.0, \
  1.0 / 39916800.0, 1.0 / 479001600.0, 1.0 / 6227020800.0 }
#if defined(__CUDACC__)
static __constant__ double fm_cd[23] = FM_COEFFS;
#endif
static const double fm_ch[23] = FM_COEFFS;
#if defined(__CUDA_ARCH__)
#define FM_C(i) fm_cd[i]
#else
#define FM_C(i) fm_ch[i]
#endif

// ln 2 = FM_LN2_HI + FM_LN2_LO; the high part has 32 significant bits, so e * FM_LN2_HI is exact for |e| < 2^21
#define FM_LN2_HI 6.93147180369123816490e-01
#define FM_LN2_LO 1.90821492927058770002e-10
#define FM_INV_LN2 1.44269504088896338700e+00
#define FM_SQRT2 1.41421356237309514547e+00

FM_HD long long fm_bits(double x) {
#if defined(__CUDA_ARCH__)
  return __double_as_longlong(x);
#else
  long long i; std::memcpy(&i, &x, 8); return i;
#endif
}
FM_HD double fm_from_bits(long long i) {
#if defined(__CUDA_ARCH__)
  return __longlong_as_double(i);
#else
  double x; std::memcpy(&x, &i, 8); return x;
#endif
}
FM_HD bool fm_normal_pos(double x) { return x >= 2.2250738585072014e-308 && x <= 1.7976931348623157e308; }

// ln x = hi + lo for a positive normal x; |hi + lo - ln x| < 2^-60 max(|ln x|, 2^-3 |ln m|)
FM_HD void fm_log_dd(double x, double& hi, double& lo) {
  const long long ix = fm_bits(x);
  int e = (int)(ix >> 52) - 1023;
  double m = fm_from_bits((ix & 0x000fffffffffffffLL) | 0x3ff0000000000000LL);      // [1, 2)
  if (m > FM_SQRT2) { m *= 0.5; e += 1; }                                           // [sqrt(1/2), sqrt(2)]
  // s = (m - 1) / (m + 1) as s_hi + s_lo; ln m = 2 atanh(s), |s| <= 0.1716
  const double num = m - 1.0;                  // exact
  const double den = m + 1.0;
  const double den_lo = m - (den - 1.0);       // exact: den + den_lo = m + 1
  const double r = 1.0 / den;
  const double s_hi = num * r;
  const double s_lo = (fma(-s_hi, den, num) - s_hi * den_lo) * r;
  const double s2 = s_hi * s_hi;
  double p = FM_C(10);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int q = 9; q >= 0; --q) p = fma(p, s2, FM_C(q));
  const double t = s_hi * s2 * p;              // 2 s^3/3 + 2 s^5/5 + ... + 2 s^23/23
  const double a = 2.0 * s_hi;
  const double b = fma(2.0, s_lo, t);
  const double ef = (double)e;
  const double c_hi = ef * FM_LN2_HI;          // exact
  const double c_lo = ef * FM_LN2_LO;
  const double h = c_hi + a;                   // |c_hi| >= ln 2 > |a| or c_hi == 0: fast two-sum
  const double l = (a - (h - c_hi)) + (b + c_lo);
  hi = h + l;
  lo = l - (hi - h);
}

FM_HD double log_pos(double x) {
  if (!fm_normal_pos(x)) return log(x);
  double hi, lo; fm_log_dd(x, hi, lo);
  return hi;
}

FM_HD double pow_pos(double x, double y) {
  if (!fm_normal_pos(x)) return pow(x, y);
  double lh, ll; fm_log_dd(x, lh, ll);
  const double ph = y * lh;
  if (!(fabs(ph) < 700.0)) return pow(x, y);
  const double pl = fma(y, lh, -ph) + y * ll;  // y ln x = ph + pl
  const double n = rint(ph * FM_INV_LN2);
  const double r1 = fma(-n, FM_LN2_HI, ph);    // exact: n * FM_LN2_HI has <= 43 bits and cancels the leading bits of ph
  const double t = fma(n, FM_LN2_LO, -pl);     // |t| < 2^-20
  const double r = r1 - t;                     // y ln x - n ln 2 = r + r_lo, |r| <= 0.3466 + 2^-20
  const double r_lo = (r1 - r) - t;
  double q = FM_C(22);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int c = 21; c >= 11; --c) q = fma(q, r, FM_C(c));
  // exp(r + r_lo) = 1 + r + (r^2 q + r_lo) summed so that only the last addition rounds at the 2^-53 level
  const double inner = fma(r * r, q, r_lo);
  const double s1 = 1.0 + r;
  const double e1 = r - (s1 - 1.0);
  const double ex = s1 + (e1 + inner);         // in [0.70, 1.42]
  return fm_from_bits(fm_bits(ex) + (long long)(int)n * 4503599627370496LL);      // * 2^n through the exponent field
}
