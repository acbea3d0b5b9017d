// Host build of the product's fast-math header (test infrastructure): lets the CPU suite check the
// polynomial kernels against libm without a GPU.  Built by tests/conftest.py into tests/native/_build/.
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
}
