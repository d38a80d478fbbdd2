// Host build of gmcore_b200/csrc/fastmath.cuh for the tests (test infrastructure, like oracle/): the same source the
// device compiles, evaluated on the CPU, plus its error against libquadmath in units in the last place.
// Built by `make oracle` into oracle/_build/libfastmath_host.so with -ffp-contract=off (the device build uses
// -fmad=false: no operation of the source is fused or split on either side).
#include "../gmcore_b200/csrc/fastmath.cuh"
#include <quadmath.h>

static double ulp_of(double v) { const double a = fabs(v); return nextafter(a, INFINITY) - a; }

extern "C" {
// out[i] = pow_pos(x[i], y) (what 0) | log_pos(x[i]) (what 1)
void fm_eval(int what, const double* x, double y, double* out, int n) {
  for (int i = 0; i < n; ++i) out[i] = what == 0 ? pow_pos(x[i], y) : log_pos(x[i]);
}
// largest error of `got` against the 113-bit pow / log of the same arguments, in ulps of the exact result
double fm_max_ulp(int what, const double* x, double y, const double* got, int n) {
  double worst = 0;
  for (int i = 0; i < n; ++i) {
    const __float128 ex = what == 0 ? powq((__float128)x[i], (__float128)y) : logq((__float128)x[i]);
    if (ex == 0) { if (got[i] != 0) worst = INFINITY; continue; }
    const double err = (double)fabsq((__float128)got[i] - ex) / ulp_of((double)ex);
    if (err > worst) worst = err;
  }
  return worst;
}
// the C library's routines on the same arguments (what the oracle calls)
void fm_libm(int what, const double* x, double y, double* out, int n) {
  for (int i = 0; i < n; ++i) out[i] = what == 0 ? pow(x[i], y) : log(x[i]);
}
}
