// Host build of the product's fast-math header (test infrastructure): lets the CPU suite check the
// polynomial kernels against libm without a GPU.  Built by tests/test_fastmath_cpu.py into tests/native/_build/.
#include "../../model-predictive-control_b200/csrc/mpcb_fastmath.cuh"

extern "C" {
void fmh_sincos(const double* x, double* s, double* c, long n) {
  for (long i = 0; i < n; ++i) mpcb::fm_sincos(x[i], &s[i], &c[i]);
}
void fmh_sin(const double* x, double* y, long n) {
  for (long i = 0; i < n; ++i) y[i] = mpcb::fm_sin(x[i]);
}
void fmh_atan(const double* x, double* y, long n) {
  for (long i = 0; i < n; ++i) y[i] = mpcb::fm_atan(x[i]);
}
// paired branch-free cores (n even)
void fmh_sincos_x2(const double* x, double* s, double* c, long n) {
  for (long i = 0; i + 1 < n; i += 2) mpcb::fm_sincos_nr_x2(x[i], x[i + 1], &s[i], &c[i], &s[i + 1], &c[i + 1]);
}
void fmh_sin_x2(const double* x, double* s, long n) {
  for (long i = 0; i + 1 < n; i += 2) mpcb::fm_sin_nr_x2(x[i], x[i + 1], &s[i], &s[i + 1]);
}
void fmh_atan_x2(const double* x, double* y, long n, int small) {
  for (long i = 0; i + 1 < n; i += 2) {
    if (small) mpcb::fm_atan_nr_x2<true>(x[i], x[i + 1], &y[i], &y[i + 1]);
    else mpcb::fm_atan_nr_x2<false>(x[i], x[i + 1], &y[i], &y[i + 1]);
  }
}
void fmh_atan_any_x2(const double* x, double* y, long n, int force_small) {
  for (long i = 0; i + 1 < n; i += 2) mpcb::fm_atan_any_x2(x[i], x[i + 1], force_small != 0, &y[i], &y[i + 1]);
}
}
