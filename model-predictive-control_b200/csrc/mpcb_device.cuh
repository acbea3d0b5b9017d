// mpcb_device.cuh — sm_100a device maths for the batched MPC hot path (fp64).
//
// Reference formulas: vehicle ODEs car_dynamics.py:93-129 / dynamics.py:144-173, RK4 elements
// car_dynamics.py:134-147, nearest vertex + errors car_dynamics.py:174-228, stage cost
// car_dynamics.py:230-258, potentials dpf_test.py:45-102, obstacle motion event_triggered.py:26.
// Library-maths model calls used by the utility kernels (mpcb_dynamics); the solver kernels use the
// forward-mode evaluation of mpcb_fwd.cuh.  Gradient conventions (what CasADi's AD generates for
// grad_f, main.py:54): selection indices constant, d sign/dx = 0, d wrap/da = 1.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "../../include/mpcb200.h"

namespace mpcb {

#define MPCB_DEV __device__ __forceinline__

// Everything the kernels need from MpcProblemDesc, passed by value (constant bank).
struct KDesc {
  int model_id, N, S, n_sub, nx, n, m, n_obs, flags, road_mode, n_roads, bounded, debug;
  int exact;       // MPCB_EXACT / MPCB_FLAG_EXACT: the solver's shortcuts are off (early line-search rejection, lemma
                   // steps decided without psi(x̂+), evaluations cut at a blown-up state) — for the tests that show
                   // the iterates do not depend on them
  int psi_nonneg;  // every cost term is a non-negative weight times a square / a Gaussian: psi >= 0 everywhere
  double h, Ts, v_ref, c[6];
  double kin_acc, kin_mu;
  double lane_a, lane_y_t, lane_b, lane_y_br, lane_y_bl;
  double obs_Af, obs_b, obs_sx, obs_sy, obs_adiv;
  double g_off[6], D_lo[6], D_hi[6];
  int rows;  // rows of g per stage: nx state rows (main.py:46-51) + n_obs safe-distance ellipses (MPCB_FLAG_SAFE_DIST); m = rows N
  double safe_x0, safe_y0, safe_ax, safe_ay;  // dpf_test.py:6
};

// Vehicle constants staged once per CTA in shared memory (runtime parameter p[22], main.py:119).
struct VehP {
  double lf, lr, m, iz, inv_m, inv_iz, cm1, cm2, cr0, cr2, bf, cf, df, br, cr, dr;
  double max_steer, max_drive, acc, mu, inv_lr, lflr;
  int trig_ok;  // |cf|, |cr| small enough for the unguarded sincos reduction of cf*atan(.)
};

MPCB_DEV void load_vehicle(VehP* v, const double* __restrict__ p, const KDesc& d) {
  v->lf = p[1]; v->lr = p[2]; v->m = p[7]; v->iz = p[8];
  v->inv_m = 1.0 / p[7]; v->inv_iz = 1.0 / p[8];
  v->max_steer = p[9]; v->max_drive = p[10];
  v->bf = p[11]; v->cf = p[12]; v->df = p[13]; v->br = p[14]; v->cr = p[15]; v->dr = p[16];
  v->cm1 = p[17]; v->cm2 = p[18]; v->cr0 = p[19]; v->cr2 = p[21];
  v->acc = d.kin_acc; v->mu = d.kin_mu; v->inv_lr = 1.0 / p[2]; v->lflr = p[1] + p[2];
  v->trig_ok = fabs(p[12]) < 8.0e4 && fabs(p[15]) < 8.0e4;
}

// Problem description and vehicle constants of the running kernel, at file scope in shared memory so the
// non-inlined evaluation function (mpcb_fwd.cuh) reads them with LDS instead of through generic pointers.
// MPCB_DEBUG_TIMERS diagnostics: model calls that took the fast / general path (lane granularity)
__device__ unsigned long long g_dbg_calls[8];  // fast, general (= wide + nan + slow) lanes; wide, nan, slow lanes; warp calls, with wide, with slow
static __shared__ VehP s_veh;
static __shared__ KDesc s_desc;
// Called by every thread of a kernel before it evaluates anything (ends with a block barrier).
MPCB_DEV void stage_problem(const KDesc& d, const double* __restrict__ p) {
  if (threadIdx.x == 0) {
    s_desc = d;
    load_vehicle(&s_veh, p, d);
  }
  __syncthreads();
}

// Stage input with its hoisted trigonometry (delta is constant over the 16 model calls of a stage).
struct StageU {
  double d, dl, sdl, cdl;
  // Pacejka-6 fast path (mpcb_fwd.cuh): 2 = the second-stage atan arguments bf*af, br*ar are within 0.5
  // whenever the slip tangents are, 1 = within 1 whenever the slip tangents are within 1, 0 = neither
  int z2_mode;
  // Kinematic-4 only: beta = atan2(lf tan(dl), lf+lr) and its derivative wrt dl
  double beta, sb, cb, dbeta;
};

template <int MODEL>
MPCB_DEV void make_stage_u(StageU& su, double d, double dl, const VehP& v, bool need_grad) {
  su.d = d; su.dl = dl;
  if constexpr (MODEL == MPCB_MODEL_PACEJKA6) {
    sincos(dl, &su.sdl, &su.cdl);
    const double adl = fabs(dl), abf = fabs(v.bf), abr = fabs(v.br);
    su.z2_mode = 0;
    if (v.trig_ok) {
      if (abf * (adl + 0.46365) <= 0.5 && abr * 0.46365 <= 0.5) su.z2_mode = 2;         // atan(0.5) = 0.46365
      else if (abf * (adl + 0.78540) <= 1.0 && abr * 0.78540 <= 1.0) su.z2_mode = 1;    // atan(1) = pi/4
    }
  } else {
    double tl = tan(dl);
    double num = v.lf * tl;
    su.beta = atan2(num, v.lflr);
    sincos(su.beta, &su.sb, &su.cb);
    su.dbeta = need_grad ? v.lflr / (num * num + v.lflr * v.lflr) * v.lf * (1.0 + tl * tl) : 0.0;
  }
}

// ---- Pacejka-6 ----------------------------------------------------------------------------------
MPCB_DEV void pacejka_f(const double* y, const StageU& su, const VehP& v, double* out) {
  double vx = y[3], vy = y[4], w = y[5];
  double sphi, cphi;
  sincos(y[2], &sphi, &cphi);
  double af = su.dl - atan2(fma(w, v.lf, vy), vx);
  double ar = atan2(fma(w, v.lr, -vy), vx);
  double sgn = vx > 0.0 ? 1.0 : (vx < 0.0 ? -1.0 : 0.0);
  double frx = (v.cm1 - v.cm2 * vx) * su.d - v.cr0 * sgn - v.cr2 * vx * vx;
  double ffy = v.df * sin(v.cf * atan(v.bf * af));
  double fry = v.dr * sin(v.cr * atan(v.br * ar));
  out[0] = vx * cphi - vy * sphi;
  out[1] = vx * sphi + vy * cphi;
  out[2] = w;
  out[3] = (frx - ffy * su.sdl + v.m * vy * w) * v.inv_m;
  out[4] = (fry + ffy * su.cdl - v.m * vx * w) * v.inv_m;
  out[5] = (ffy * v.lf * su.cdl - fry * v.lr) * v.inv_iz;
}

// ---- Kinematic-4 --------------------------------------------------------------------------------
MPCB_DEV void kinematic_f(const double* y, const StageU& su, const VehP& v, double* out) {
  double s, c;
  sincos(y[2] + su.beta, &s, &c);
  out[0] = y[3] * c;
  out[1] = y[3] * s;
  out[2] = y[3] * su.sb * v.inv_lr;
  out[3] = v.acc * su.d - v.mu * y[3];
}

template <int MODEL> struct ModelTraits;
template <> struct ModelTraits<MPCB_MODEL_PACEJKA6> {
  static constexpr int NX = 6;
  static MPCB_DEV void f(const double* y, const StageU& su, const VehP& v, double* o) { pacejka_f(y, su, v, o); }
};
template <> struct ModelTraits<MPCB_MODEL_KINEMATIC4> {
  static constexpr int NX = 4;
  static MPCB_DEV void f(const double* y, const StageU& su, const VehP& v, double* o) { kinematic_f(y, su, v, o); }
};

}  // namespace mpcb
