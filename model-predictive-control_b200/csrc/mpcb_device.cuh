// mpcb_device.cuh — sm_100a device maths for the batched MPC hot path (fp64).
//
// Reference formulas: vehicle ODEs car_dynamics.py:93-129 / dynamics.py:144-173, RK4 elements
// car_dynamics.py:134-147, nearest vertex + errors car_dynamics.py:174-228, stage cost
// car_dynamics.py:230-258, potentials dpf_test.py:45-102, obstacle motion event_triggered.py:26.
// The reverse sweep is a hand-written adjoint (what CasADi's reverse AD generates for grad_f,
// main.py:54): selection indices constant, d sign/dx = 0, d wrap/da = 1.
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include "../../include/mpcb200.h"

namespace mpcb {

#define MPCB_DEV __device__ __forceinline__

// Everything the kernels need from MpcProblemDesc, passed by value (constant bank).
struct KDesc {
  int model_id, N, S, n_sub, nx, n, m, n_obs, flags, road_mode, n_roads, bounded;
  double h, Ts, v_ref, c[6];
  double kin_acc, kin_mu;
  double lane_a, lane_y_t, lane_b, lane_y_br, lane_y_bl;
  double obs_Af, obs_b, obs_sx, obs_sy, obs_adiv;
  double g_off[6], D_lo[6], D_hi[6];
};

// Vehicle constants staged once per CTA in shared memory (runtime parameter p[22], main.py:119).
struct VehP {
  double lf, lr, m, iz, inv_m, inv_iz, cm1, cm2, cr0, cr2, bf, cf, df, br, cr, dr;
  double max_steer, max_drive, acc, mu, inv_lr, lflr;
};

MPCB_DEV void load_vehicle(VehP* v, const double* __restrict__ p, const KDesc& d) {
  v->lf = p[1]; v->lr = p[2]; v->m = p[7]; v->iz = p[8];
  v->inv_m = 1.0 / p[7]; v->inv_iz = 1.0 / p[8];
  v->max_steer = p[9]; v->max_drive = p[10];
  v->bf = p[11]; v->cf = p[12]; v->df = p[13]; v->br = p[14]; v->cr = p[15]; v->dr = p[16];
  v->cm1 = p[17]; v->cm2 = p[18]; v->cr0 = p[19]; v->cr2 = p[21];
  v->acc = d.kin_acc; v->mu = d.kin_mu; v->inv_lr = 1.0 / p[2]; v->lflr = p[1] + p[2];
}

// Stage input with its hoisted trigonometry (delta is constant over the 16 model calls of a stage).
struct StageU {
  double d, dl, sdl, cdl;
  // Kinematic-4 only: beta = atan2(lf tan(dl), lf+lr) and its derivative wrt dl
  double beta, sb, cb, dbeta;
};

template <int MODEL>
MPCB_DEV void make_stage_u(StageU& su, double d, double dl, const VehP& v, bool need_grad) {
  su.d = d; su.dl = dl;
  if constexpr (MODEL == MPCB_MODEL_PACEJKA6) {
    sincos(dl, &su.sdl, &su.cdl);
  } else {
    double tl = tan(dl);
    double num = v.lf * tl;
    su.beta = atan2(num, v.lflr);
    sincos(su.beta, &su.sb, &su.cb);
    su.dbeta = need_grad ? v.lflr / (num * num + v.lflr * v.lflr) * v.lf * (1.0 + tl * tl) : 0.0;
  }
}

// Compact linearisation of one model call: what the vjp needs besides the adjoint seed.
struct Lin6 {  // Pacejka
  double sphi, cphi, vx, vy, w, ffy, Gf, qf1, qf2, Gr, qr1, qr2;
};
struct Lin4 {  // Kinematic
  double s, c, v;
};

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

// f plus the linearisation record (shares every transcendental with f).
MPCB_DEV void pacejka_lin(const double* y, const StageU& su, const VehP& v, double* out, Lin6& L) {
  double vx = y[3], vy = y[4], w = y[5];
  sincos(y[2], &L.sphi, &L.cphi);
  L.vx = vx; L.vy = vy; L.w = w;
  double a_f = fma(w, v.lf, vy), a_r = fma(w, v.lr, -vy);
  double af = su.dl - atan2(a_f, vx);
  double ar = atan2(a_r, vx);
  double sgn = vx > 0.0 ? 1.0 : (vx < 0.0 ? -1.0 : 0.0);
  double frx = (v.cm1 - v.cm2 * vx) * su.d - v.cr0 * sgn - v.cr2 * vx * vx;
  double zf = v.bf * af, zr = v.br * ar;
  double stf, ctf, str_, ctr;
  sincos(v.cf * atan(zf), &stf, &ctf);
  sincos(v.cr * atan(zr), &str_, &ctr);
  double ffy = v.df * stf, fry = v.dr * str_;
  L.ffy = ffy;
  L.Gf = v.df * ctf * v.cf * v.bf / fma(zf, zf, 1.0);
  L.Gr = v.dr * ctr * v.cr * v.br / fma(zr, zr, 1.0);
  double vx2 = vx * vx;
  double idf = 1.0 / fma(a_f, a_f, vx2), idr = 1.0 / fma(a_r, a_r, vx2);
  L.qf1 = a_f * idf; L.qf2 = vx * idf;
  L.qr1 = a_r * idr; L.qr2 = vx * idr;
  out[0] = vx * L.cphi - vy * L.sphi;
  out[1] = vx * L.sphi + vy * L.cphi;
  out[2] = w;
  out[3] = (frx - ffy * su.sdl + v.m * vy * w) * v.inv_m;
  out[4] = (fry + ffy * su.cdl - v.m * vx * w) * v.inv_m;
  out[5] = (ffy * v.lf * su.cdl - fry * v.lr) * v.inv_iz;
}

// ybar = J_y^T wv (overwrites), ubar += J_u^T wv.
MPCB_DEV void pacejka_vjp(const Lin6& L, const StageU& su, const VehP& v, const double* wv, double* ybar,
                          double* ubar) {
  double f0 = L.vx * L.cphi - L.vy * L.sphi, f1 = L.vx * L.sphi + L.vy * L.cphi;
  double w3 = wv[3] * v.inv_m, w4 = wv[4] * v.inv_m, w5 = wv[5] * v.inv_iz;
  double ffy_b = -w3 * su.sdl + w4 * su.cdl + w5 * v.lf * su.cdl;
  double fry_b = w4 - w5 * v.lr;
  double af_b = ffy_b * L.Gf, ar_b = fry_b * L.Gr;
  double a_f_b = -af_b * L.qf2, a_r_b = ar_b * L.qr2;
  ybar[0] = 0.0;
  ybar[1] = 0.0;
  ybar[2] = wv[1] * f0 - wv[0] * f1;
  ybar[3] = wv[0] * L.cphi + wv[1] * L.sphi - wv[4] * L.w + w3 * (-v.cm2 * su.d - 2.0 * v.cr2 * L.vx) +
            af_b * L.qf1 - ar_b * L.qr1;
  ybar[4] = wv[1] * L.cphi - wv[0] * L.sphi + wv[3] * L.w + a_f_b - a_r_b;
  ybar[5] = wv[2] + wv[3] * L.vy - wv[4] * L.vx + a_f_b * v.lf + a_r_b * v.lr;
  ubar[0] += w3 * (v.cm1 - v.cm2 * L.vx);
  ubar[1] += -L.ffy * (w3 * su.cdl + w4 * su.sdl + w5 * v.lf * su.sdl) + af_b;
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
MPCB_DEV void kinematic_lin(const double* y, const StageU& su, const VehP& v, double* out, Lin4& L) {
  sincos(y[2] + su.beta, &L.s, &L.c);
  L.v = y[3];
  out[0] = y[3] * L.c;
  out[1] = y[3] * L.s;
  out[2] = y[3] * su.sb * v.inv_lr;
  out[3] = v.acc * su.d - v.mu * y[3];
}
MPCB_DEV void kinematic_vjp(const Lin4& L, const StageU& su, const VehP& v, const double* wv, double* ybar,
                            double* ubar) {
  double t = L.v * (wv[1] * L.c - wv[0] * L.s);
  ybar[0] = 0.0;
  ybar[1] = 0.0;
  ybar[2] = t;
  ybar[3] = wv[0] * L.c + wv[1] * L.s + wv[2] * su.sb * v.inv_lr - wv[3] * v.mu;
  ubar[0] += wv[3] * v.acc;
  ubar[1] += (t + wv[2] * L.v * su.cb * v.inv_lr) * su.dbeta;
}

template <int MODEL> struct ModelTraits;
template <> struct ModelTraits<MPCB_MODEL_PACEJKA6> {
  static constexpr int NX = 6;
  using Lin = Lin6;
  static MPCB_DEV void f(const double* y, const StageU& su, const VehP& v, double* o) { pacejka_f(y, su, v, o); }
  static MPCB_DEV void lin(const double* y, const StageU& su, const VehP& v, double* o, Lin& L) { pacejka_lin(y, su, v, o, L); }
  static MPCB_DEV void vjp(const Lin& L, const StageU& su, const VehP& v, const double* w, double* yb, double* ub) { pacejka_vjp(L, su, v, w, yb, ub); }
};
template <> struct ModelTraits<MPCB_MODEL_KINEMATIC4> {
  static constexpr int NX = 4;
  using Lin = Lin4;
  static MPCB_DEV void f(const double* y, const StageU& su, const VehP& v, double* o) { kinematic_f(y, su, v, o); }
  static MPCB_DEV void lin(const double* y, const StageU& su, const VehP& v, double* o, Lin& L) { kinematic_lin(y, su, v, o, L); }
  static MPCB_DEV void vjp(const Lin& L, const StageU& su, const VehP& v, const double* w, double* yb, double* ub) { kinematic_vjp(L, su, v, w, yb, ub); }
};

// ---- RK4 element ----------------------------------------------------------------------------------
template <int MODEL>
MPCB_DEV void rk4_element(double* y, const StageU& su, const VehP& v, double h) {
  using MT = ModelTraits<MODEL>;
  constexpr int NX = MT::NX;
  double k[NX], t[NX], acc[NX];
  MT::f(y, su, v, k);
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] = k[i]; t[i] = fma(0.5 * h, k[i], y[i]); }
  MT::f(t, su, v, k);
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] = fma(2.0, k[i], acc[i]); t[i] = fma(0.5 * h, k[i], y[i]); }
  MT::f(t, su, v, k);
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] = fma(2.0, k[i], acc[i]); t[i] = fma(h, k[i], y[i]); }
  MT::f(t, su, v, k);
  const double h6 = h * (1.0 / 6.0);
#pragma unroll
  for (int i = 0; i < NX; ++i) y[i] = fma(h6, acc[i] + k[i], y[i]);
}

// Reverse of one element starting at state y (its checkpoint): lam <- d/dy, ubar += d/du.
template <int MODEL>
MPCB_DEV void rk4_element_adj(const double* y, const StageU& su, const VehP& v, double h, double* lam,
                              double* ubar) {
  using MT = ModelTraits<MODEL>;
  constexpr int NX = MT::NX;
  typename MT::Lin L1, L2, L3, L4;
  double k[NX], t[NX];
  MT::lin(y, su, v, k, L1);
#pragma unroll
  for (int i = 0; i < NX; ++i) t[i] = fma(0.5 * h, k[i], y[i]);
  MT::lin(t, su, v, k, L2);
#pragma unroll
  for (int i = 0; i < NX; ++i) t[i] = fma(0.5 * h, k[i], y[i]);
  MT::lin(t, su, v, k, L3);
#pragma unroll
  for (int i = 0; i < NX; ++i) t[i] = fma(h, k[i], y[i]);
  MT::lin(t, su, v, k, L4);
  const double h6 = h * (1.0 / 6.0), h3 = h * (1.0 / 3.0);
  double kb[NX], yb[NX], acc[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] = lam[i]; kb[i] = h6 * lam[i]; }
  MT::vjp(L4, su, v, kb, yb, ubar);
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] += yb[i]; kb[i] = fma(h, yb[i], h3 * lam[i]); }
  MT::vjp(L3, su, v, kb, yb, ubar);
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] += yb[i]; kb[i] = fma(0.5 * h, yb[i], h3 * lam[i]); }
  MT::vjp(L2, su, v, kb, yb, ubar);
#pragma unroll
  for (int i = 0; i < NX; ++i) { acc[i] += yb[i]; kb[i] = fma(0.5 * h, yb[i], h6 * lam[i]); }
  MT::vjp(L1, su, v, kb, yb, ubar);
#pragma unroll
  for (int i = 0; i < NX; ++i) lam[i] = acc[i] + yb[i];
}

}  // namespace mpcb
