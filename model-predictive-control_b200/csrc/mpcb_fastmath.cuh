// mpcb_fastmath.cuh — short fp64 sincos / atan / reciprocal kernels for the solver kernels.
//
// Why: one Pacejka model call is 4 atan + 3 sincos + a division, and the solve does 16 N of them in
// a dependent chain per evaluation.  With libdevice (or any branchy implementation) the call is
// latency-bound at ~2050 cycles for a lone warp: every transcendental is its own serial Horner chain
// and the branches keep ptxas from interleaving the independent front/rear tyre chains
// (profiles/r01_ubench_model_v2a.json).  The *_x2 kernels below are branch-free, evaluate two
// arguments statement by statement (ILP 2, one materialised coefficient feeds two DFMAs), and split
// every polynomial into even/odd halves (half the dependent depth).  Callers check the argument
// ranges once per model call and fall back to the general (branchy) versions otherwise.
// Accuracy ~1 ulp (approximation error < 2.3e-17, tools/gen_fastmath.py), far inside the 1e-9
// rollout tolerance against the oracle's libm.
//
// The same header compiles as plain host C++ (tests/test_fastmath_cpu.py checks every kernel against
// libm on the CPU): fm_rcp falls back to 1/x there, everything else is identical arithmetic.
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
// 32-bit-immediate operands only, no constant-bank operand): MPCB_FM_MODE 0 = literals (two UMOV per
// coefficient, shared by the paired evaluators), 2 = one shared-memory table filled by fm_load_smem()
// at kernel start (one LDS.64 per coefficient).  Measured with tools/ubench_model.cu.
#ifndef MPCB_FM_MODE
#define MPCB_FM_MODE 0
#endif
#if defined(__CUDACC__)
#define MPCB_FM_TABLE(name, n, ...) static __constant__ double name##_d[n] = {__VA_ARGS__};
#else
#define MPCB_FM_TABLE(name, n, ...)
#endif
#if defined(__CUDA_ARCH__) && MPCB_FM_MODE == 2
#define MPCB_FM_K(name, i, lit) (fm_smem[name##_off + (i)])
#else
#define MPCB_FM_K(name, i, lit) (lit)
#endif

namespace mpcb {

#if defined(__CUDACC__) && MPCB_FM_MODE == 2
static __shared__ double fm_smem[80];
#endif

#include "mpcb_fastmath_coef.h"

#if defined(__CUDACC__)
// Mode 2: every kernel that uses the fm_* functions calls this once (all threads) before them.
__device__ __forceinline__ void fm_load_smem() {
#if MPCB_FM_MODE == 2
  for (int i = threadIdx.x; i < MPCB_FM_SMEM_DOUBLES; i += blockDim.x) fm_smem[i] = fm_c_all_d[i];
  __syncthreads();
#endif
}
#endif

#define MPCB_FM_MAGIC 6755399441055744.0 /* 1.5 * 2^52: adding it rounds to the nearest integer */
#define MPCB_FM_TRIG_MAX 131072.0        /* Cody-Waite reduction below is exact enough for |x| < 2^17 */

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

MPCB_FM_FN double fm_flip(double v, int neg) {  // neg != 0 -> -v (integer sign-bit flip, no fp64 slot)
#if defined(__CUDA_ARCH__)
  return __hiloint2double(__double2hiint(v) ^ (neg ? (int)0x80000000 : 0), __double2loint(v));
#else
  return neg ? -v : v;
#endif
}

// ---- branch-free cores: the caller guarantees |x| < MPCB_FM_TRIG_MAX -------------------------------
MPCB_FM_FN void fm_sincos_nr(double x, double* s, double* c) {
  double kd = fma(x, MPCB_FM_TWO_OVER_PI, MPCB_FM_MAGIC);
  int q = fm_lo32(kd);
  double k = kd - MPCB_FM_MAGIC;
  double r = fma(-k, MPCB_FM_PIO2_1, x);
  r = fma(-k, MPCB_FM_PIO2_2, r);
  r = fma(-k, MPCB_FM_PIO2_3, r);
  double z = r * r;
  double sr = fma(r * z, fm_poly_sin(z), r);
  double cr = fma(z * z, fm_poly_cos(z), fma(-0.5, z, 1.0));
  double ss = (q & 1) ? cr : sr;
  double cc = (q & 1) ? sr : cr;
  *s = fm_flip(ss, q & 2);
  *c = fm_flip(cc, (q + 1) & 2);
}

MPCB_FM_FN void fm_sincos_nr_x2(double x1, double x2, double* s1, double* c1, double* s2, double* c2) {
  double kd1 = fma(x1, MPCB_FM_TWO_OVER_PI, MPCB_FM_MAGIC), kd2 = fma(x2, MPCB_FM_TWO_OVER_PI, MPCB_FM_MAGIC);
  int q1 = fm_lo32(kd1), q2 = fm_lo32(kd2);
  double k1 = kd1 - MPCB_FM_MAGIC, k2 = kd2 - MPCB_FM_MAGIC;
  double r1 = fma(-k1, MPCB_FM_PIO2_1, x1), r2 = fma(-k2, MPCB_FM_PIO2_1, x2);
  r1 = fma(-k1, MPCB_FM_PIO2_2, r1); r2 = fma(-k2, MPCB_FM_PIO2_2, r2);
  r1 = fma(-k1, MPCB_FM_PIO2_3, r1); r2 = fma(-k2, MPCB_FM_PIO2_3, r2);
  double z1 = r1 * r1, z2 = r2 * r2;
  double ps1, ps2, pc1, pc2;
  fm_poly_sin_x2(z1, z2, &ps1, &ps2);
  fm_poly_cos_x2(z1, z2, &pc1, &pc2);
  double sr1 = fma(r1 * z1, ps1, r1), sr2 = fma(r2 * z2, ps2, r2);
  double cr1 = fma(z1 * z1, pc1, fma(-0.5, z1, 1.0)), cr2 = fma(z2 * z2, pc2, fma(-0.5, z2, 1.0));
  double ss1 = (q1 & 1) ? cr1 : sr1, cc1 = (q1 & 1) ? sr1 : cr1;
  double ss2 = (q2 & 1) ? cr2 : sr2, cc2 = (q2 & 1) ? sr2 : cr2;
  *s1 = fm_flip(ss1, q1 & 2); *c1 = fm_flip(cc1, (q1 + 1) & 2);
  *s2 = fm_flip(ss2, q2 & 2); *c2 = fm_flip(cc2, (q2 + 1) & 2);
}

// sin only, two arguments: reduction modulo pi and one odd kernel on [-pi/2, pi/2] (no cosine kernel,
// no quadrant select) — what a value-only model call needs for the two Pacejka forces.
MPCB_FM_FN void fm_sin_nr_x2(double x1, double x2, double* s1, double* s2) {
  double kd1 = fma(x1, MPCB_FM_ONE_OVER_PI, MPCB_FM_MAGIC), kd2 = fma(x2, MPCB_FM_ONE_OVER_PI, MPCB_FM_MAGIC);
  int q1 = fm_lo32(kd1), q2 = fm_lo32(kd2);
  double k1 = kd1 - MPCB_FM_MAGIC, k2 = kd2 - MPCB_FM_MAGIC;
  const double pi_hi = 2.0 * MPCB_FM_PIO2_1;
  double r1 = fma(-k1, pi_hi, x1), r2 = fma(-k2, pi_hi, x2);
  r1 = fma(-k1, MPCB_FM_PI_LO, r1); r2 = fma(-k2, MPCB_FM_PI_LO, r2);
  double z1 = r1 * r1, z2 = r2 * r2;
  double p1, p2;
  fm_poly_sin_wide_x2(z1, z2, &p1, &p2);
  *s1 = fm_flip(fma(r1 * z1, p1, r1), q1 & 1);
  *s2 = fm_flip(fma(r2 * z2, p2, r2), q2 & 1);
}

// atan of two arguments with |x| <= 0.5 (SMALL) or |x| <= 1.
template <bool SMALL>
MPCB_FM_FN void fm_atan_nr_x2(double x1, double x2, double* y1, double* y2) {
  double z1 = x1 * x1, z2 = x2 * x2;
  double p1, p2;
  if (SMALL) fm_poly_atan_small_x2(z1, z2, &p1, &p2);
  else fm_poly_atan_full_x2(z1, z2, &p1, &p2);
  *y1 = fma(x1 * z1, p1, x1);
  *y2 = fma(x2 * z2, p2, x2);
}

// atan of two arguments of any magnitude, branch-free (selects only): |x| > 1 goes through 1/|x|, and each
// argument takes the short polynomial when |x| <= 0.5 (or when the caller forces it: `force_small` lanes get
// exactly fm_atan_nr_x2<true>'s bits) and the full one otherwise — the same arithmetic, argument by argument,
// as fm_atan below.  NaN -> NaN, +-Inf -> +-pi/2.
MPCB_FM_FN void fm_atan_any_x2(double x1, double x2, bool force_small, double* y1, double* y2) {
  const double ax1 = fabs(x1), ax2 = fabs(x2);
  const bool big1 = !force_small && ax1 > 1.0, big2 = !force_small && ax2 > 1.0;
  const bool sm1 = force_small || ax1 <= 0.5, sm2 = force_small || ax2 <= 0.5;
  const double rc1 = fm_rcp(fmin(fmax(ax1, 1.0), 0x1p1000)), rc2 = fm_rcp(fmin(fmax(ax2, 1.0), 0x1p1000));
  // forced lanes keep the signed argument: fma(x z, p, x) exactly as fm_atan_nr_x2 (the sign of a zero result included)
  const double t1 = force_small ? x1 : (big1 ? (ax1 < 0x1p1000 ? rc1 : 0.0) : ax1);
  const double t2 = force_small ? x2 : (big2 ? (ax2 < 0x1p1000 ? rc2 : 0.0) : ax2);
  const double z1 = t1 * t1, z2 = t2 * t2;
  double ps1, ps2, pf1, pf2;
  fm_poly_atan_small_x2(z1, z2, &ps1, &ps2);
  fm_poly_atan_full_x2(z1, z2, &pf1, &pf2);
  double r1 = fma(t1 * z1, sm1 ? ps1 : pf1, t1), r2 = fma(t2 * z2, sm2 ? ps2 : pf2, t2);
  const double c1 = (MPCB_FM_PIO2_1 - r1) + MPCB_FM_PIO2_2, c2 = (MPCB_FM_PIO2_1 - r2) + MPCB_FM_PIO2_2;
  r1 = big1 ? c1 : r1;
  r2 = big2 ? c2 : r2;
  *y1 = force_small ? r1 : copysign(r1, x1);
  *y2 = force_small ? r2 : copysign(r2, x2);
}

// ---- general versions (any argument; branchy: slow path of the model call, utility kernels) -------
MPCB_FM_FN void fm_sincos(double x, double* s, double* c) {
  if (!(fabs(x) < MPCB_FM_TRIG_MAX)) {  // huge / NaN / Inf: library path
    *s = sin(x);
    *c = cos(x);
    return;
  }
  fm_sincos_nr(x, s, c);
}

MPCB_FM_FN double fm_sin(double x) {
  if (!(fabs(x) < MPCB_FM_TRIG_MAX)) return sin(x);
  double s, c;
  fm_sincos_nr(x, &s, &c);
  return s;
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
