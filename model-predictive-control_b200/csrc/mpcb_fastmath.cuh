// mpcb_fastmath.cuh — short, branch-light fp64 sincos / atan / reciprocal for the solver kernels.
//
// Why: the solve is bound by the number of FP64-pipe instructions one model call issues (a lone warp
// issues one DFMA per ~3.75 cycles on B200, profiles/r01_ubench.json).  libdevice's atan2/atan/sin
// cost ~60-120 fp64 instructions each with slow-path branches; the versions below cost 18-27 and are
// accurate to ~1 ulp (approximation error < 6e-18, tools/gen_fastmath.py), far inside the 1e-9
// rollout tolerance against the oracle's libm.
//
// The same header compiles as plain host C++ (tests/test_fastmath_cpu.py checks it against libm on
// the CPU): fm_rcp falls back to 1/x there, everything else is identical arithmetic.
#pragma once
#include <math.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define MPCB_FM_FN __host__ __device__ __forceinline__
#else
#define MPCB_FM_FN inline
#endif

// How device code gets polynomial coefficients (sm_100a DFMA takes register, uniform-register or
// 32-bit-immediate operands only): 0 = literals (two UMOV each), 1 = __constant__ tables,
// 2 = one shared-memory table filled by fm_load_smem() at kernel start.  Measured with
// tools/ubench_model.cu; see profiles/r01_ubench_model.json.
#ifndef MPCB_FM_MODE
#define MPCB_FM_MODE 0
#endif
#if defined(__CUDACC__)
#define MPCB_FM_TABLE(name, n, ...)                  \
  static const double name##_h[n] = {__VA_ARGS__};   \
  static __constant__ double name##_d[n] = {__VA_ARGS__};
#else
#define MPCB_FM_TABLE(name, n, ...) static const double name##_h[n] = {__VA_ARGS__};
#endif
#if !defined(__CUDA_ARCH__)
#define MPCB_FM_C(name) name##_h
#elif MPCB_FM_MODE == 2
#define MPCB_FM_C(name) (fm_smem + name##_off)
#else
#define MPCB_FM_C(name) name##_d
#endif

namespace mpcb {

#if defined(__CUDACC__) && MPCB_FM_MODE == 2
static __shared__ double fm_smem[64];
#endif

#include "mpcb_fastmath_coef.h"

#if defined(__CUDACC__)
// Mode 2: every kernel that uses the fm_* functions calls this once (all threads) before them.
__device__ __forceinline__ void fm_load_smem() {
#if MPCB_FM_MODE == 2
  for (int i = threadIdx.x; i < MPCB_FM_SMEM_DOUBLES; i += blockDim.x) {
    double v;
    if (i < fm_c_atan_full_off) v = fm_c_atan_small_d[i];
    else if (i < fm_c_sin_off) v = fm_c_atan_full_d[i - fm_c_atan_full_off];
    else if (i < fm_c_cos_off) v = fm_c_sin_d[i - fm_c_sin_off];
    else if (i < fm_c_misc_off) v = fm_c_cos_d[i - fm_c_cos_off];
    else v = fm_c_misc_d[i - fm_c_misc_off];
    fm_smem[i] = v;
  }
  __syncthreads();
#endif
}
#endif

// 1/x to <= 1 ulp: MUFU seed (20 bits, measured 9.9e-7) + one third-order correction (3 DFMA).
// Finite non-zero normal x only (callers guard).
MPCB_FM_FN double fm_rcp(double x) {
#if defined(__CUDA_ARCH__)
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  double e = fma(-x, r, 1.0);
  double t = fma(e, e, e);
  return fma(r, t, r);
#else
  return 1.0 / x;
#endif
}

MPCB_FM_FN int fm_lo32(double v) {
#if defined(__CUDA_ARCH__)
  return __double2loint(v);
#else
  uint64_t u;
  memcpy(&u, &v, 8);
  return (int)(uint32_t)u;
#endif
}

// sin and cos of x.  Cody-Waite reduction with three FMA (valid for |x| < 2^17), 7-term kernels.
MPCB_FM_FN void fm_sincos(double x, double* s, double* c) {
  if (!(fabs(x) < 131072.0)) {  // huge / NaN / Inf: library path (never taken by sane headings)
    *s = sin(x);
    *c = cos(x);
    return;
  }
  const double magic = 6755399441055744.0;  // 1.5 * 2^52: adding it rounds to nearest integer
  double kd = fma(x, MPCB_FM_TWO_OVER_PI, magic);
  int q = fm_lo32(kd);
  double k = kd - magic;
  double r = fma(-k, MPCB_FM_PIO2_1, x);
  r = fma(-k, MPCB_FM_PIO2_2, r);
  r = fma(-k, MPCB_FM_PIO2_3, r);
  double z = r * r;
  double sr = fma(r * z, fm_poly_sin(z), r);
  double cr = fma(z * z, fm_poly_cos(z), fma(-0.5, z, 1.0));
  double ss = (q & 1) ? cr : sr;
  double cc = (q & 1) ? sr : cr;
  *s = (q & 2) ? -ss : ss;
  *c = ((q + 1) & 2) ? -cc : cc;
}

// sin(x) alone (value-only evaluations skip the cosine kernel unless the quadrant asks for it).
MPCB_FM_FN double fm_sin(double x) {
  if (!(fabs(x) < 131072.0)) return sin(x);
  const double magic = 6755399441055744.0;
  double kd = fma(x, MPCB_FM_TWO_OVER_PI, magic);
  int q = fm_lo32(kd);
  double k = kd - magic;
  double r = fma(-k, MPCB_FM_PIO2_1, x);
  r = fma(-k, MPCB_FM_PIO2_2, r);
  r = fma(-k, MPCB_FM_PIO2_3, r);
  double z = r * r;
  double v;
  if (q & 1) v = fma(z * z, fm_poly_cos(z), fma(-0.5, z, 1.0));
  else v = fma(r * z, fm_poly_sin(z), r);
  return (q & 2) ? -v : v;
}

// atan(x), all finite x (and +-Inf -> +-pi/2, NaN -> NaN).
MPCB_FM_FN double fm_atan(double x) {
  double ax = fabs(x);
  double r;
  if (ax <= 0.5) {
    double z = ax * ax;
    r = fma(ax * z, fm_poly_atan_small(z), ax);
  } else {
    bool big = ax > 1.0;
    double t = ax;
    if (big) t = (ax < 0x1p1000) ? fm_rcp(ax) : 0.0;
    double z = t * t;
    r = fma(t * z, fm_poly_atan_full(z), t);
    if (big) r = (MPCB_FM_PIO2_1 - r) + MPCB_FM_PIO2_2;
  }
  return copysign(r, x);
}

}  // namespace mpcb
