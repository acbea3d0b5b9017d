// orc_model.hpp — CPU ORACLE (test infrastructure, NOT product code).
//
// Plain scalar restatement of the reference's maths for the MPC hot path.  Only tests/,
// __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// PARITY STATUS: "parity unpinned" against alpaqa/CasADi (neither is installable here);
// pinned against the reference's own NumPy code (dynamics.py, road.py, bezier_curves.py,
// dpf_test.py) through tests/golden/*.json, which tests/golden/gen_golden.py produced by
// importing /root/reference.
//
// Templated on the scalar so the same text runs with double (parity, CPU baseline) and with a
// counting scalar (algorithmic flop tally for the roofline).
#pragma once
#include <cmath>
#include <cstdint>
#include <vector>
#include "../include/mpcb200.h"

namespace orc {

using std::atan;
using std::atan2;
using std::cos;
using std::exp;
using std::floor;
using std::fmod;
using std::pow;
using std::sin;
using std::sqrt;
using std::tan;

static const double kPi = 3.141592653589793238462643383279502884;

template <class T>
inline T sign_of(const T& v) {  // casadi sign(): -1, 0, +1 (car_dynamics.py:103)
  return v > T(0) ? T(1) : (v < T(0) ? T(-1) : T(0));
}

template <class T>
inline T clip(const T& v, const T& lo, const T& hi) {  // np.clip (dynamics.py:56-66)
  return v < lo ? lo : (v > hi ? hi : v);
}

// np.mod(a + pi, 2 pi) - pi  (car_dynamics.py:168-172, road.py:22-34).  np.mod is floor-mod:
// fmod, then shifted by the divisor when the signs differ.
template <class T>
inline T wrap_to_pi(const T& a) {
  T two_pi = T(2.0 * kPi);
  T m = fmod(a + T(kPi), two_pi);
  if (m != T(0) && (m < T(0))) m = m + two_pi;
  return m - T(kPi);
}

// Vehicle parameter vector, order of car_dynamics.py:65-88 / main.py:88-111.
enum { P_LENGTH = 0, P_LF, P_LR, P_FRONT, P_REAR, P_WIDTH, P_HEIGHT, P_MASS, P_INERTIA, P_MAX_STEER,
       P_MAX_DRIVE, P_BF, P_CF, P_DF, P_BR, P_CR, P_DR, P_CM1, P_CM2, P_CR0, P_CR1, P_CR2 };

inline int nx_of(int model_id) { return model_id == MPCB_MODEL_KINEMATIC4 ? 4 : 6; }

// ---------------------------------------------------------------------------------------------
// Continuous-time models
// ---------------------------------------------------------------------------------------------

// Pacejka bicycle, car_dynamics.py:93-129 (twin dynamics.py:67-119).  y=[x,y,phi,vx,vy,w], u=[d,delta].
template <class T>
inline void pacejka_f(const T* y, const T* u, const T* p, T* out) {
  T d = u[0], dl = u[1];
  T phi = y[2], vx = y[3], vy = y[4], w = y[5];
  T lf = p[P_LF], lr = p[P_LR], m = p[P_MASS], iz = p[P_INERTIA];
  T cr0 = p[P_CR0] * sign_of(vx);
  T af = -atan2(w * lf + vy, vx) + dl;
  T ar = atan2(w * lr - vy, vx);
  T frx = (p[P_CM1] - p[P_CM2] * vx) * d - cr0 - p[P_CR2] * vx * vx;
  T ffy = p[P_DF] * sin(p[P_CF] * atan(p[P_BF] * af));
  T fry = p[P_DR] * sin(p[P_CR] * atan(p[P_BR] * ar));
  T sphi = sin(phi), cphi = cos(phi), sdl = sin(dl), cdl = cos(dl);
  out[0] = vx * cphi - vy * sphi;
  out[1] = vx * sphi + vy * cphi;
  out[2] = w;
  out[3] = (frx - ffy * sdl + m * vy * w) / m;
  out[4] = (fry + ffy * cdl - m * vx * w) / m;
  out[5] = (ffy * lf * cdl - fry * lr) / iz;
}

// Vector-Jacobian product of pacejka_f: ybar += J_y^T w, ubar += J_u^T w.  This is what CasADi's
// reverse AD of the graph above produces (main.py:54 -> grad_f); d sign/dvx = 0.
template <class T>
inline void pacejka_vjp(const T* y, const T* u, const T* p, const T* wv, T* ybar, T* ubar) {
  T d = u[0], dl = u[1];
  T phi = y[2], vx = y[3], vy = y[4], w = y[5];
  T lf = p[P_LF], lr = p[P_LR], m = p[P_MASS], iz = p[P_INERTIA];
  T a_f = w * lf + vy, a_r = w * lr - vy;
  T af = -atan2(a_f, vx) + dl;
  T ar = atan2(a_r, vx);
  T zf = p[P_BF] * af, zr = p[P_BR] * ar;
  T tf = p[P_CF] * atan(zf), tr = p[P_CR] * atan(zr);
  T ffy = p[P_DF] * sin(tf);
  T dffy = p[P_DF] * cos(tf) * p[P_CF] * p[P_BF] / (T(1) + zf * zf);  // d ffy / d af
  T dfry = p[P_DR] * cos(tr) * p[P_CR] * p[P_BR] / (T(1) + zr * zr);  // d fry / d ar
  T sphi = sin(phi), cphi = cos(phi), sdl = sin(dl), cdl = cos(dl);
  T f0 = vx * cphi - vy * sphi, f1 = vx * sphi + vy * cphi;
  // rows 0,1,2
  ybar[2] = ybar[2] + wv[0] * (-f1) + wv[1] * f0;
  ybar[3] = ybar[3] + wv[0] * cphi + wv[1] * sphi;
  ybar[4] = ybar[4] - wv[0] * sphi + wv[1] * cphi;
  ybar[5] = ybar[5] + wv[2];
  // force adjoints
  T frx_b = wv[3] / m;
  T ffy_b = (-wv[3] * sdl + wv[4] * cdl) / m + wv[5] * lf * cdl / iz;
  T fry_b = wv[4] / m - wv[5] * lr / iz;
  // direct m*vy*w / m*vx*w terms
  ybar[4] = ybar[4] + wv[3] * w;
  ybar[5] = ybar[5] + wv[3] * vy - wv[4] * vx;
  ybar[3] = ybar[3] - wv[4] * w;
  // direct delta terms
  ubar[1] = ubar[1] + (-wv[3] * ffy * cdl - wv[4] * ffy * sdl) / m - wv[5] * ffy * lf * sdl / iz;
  // frx
  ybar[3] = ybar[3] + frx_b * (-p[P_CM2] * d - T(2) * p[P_CR2] * vx);
  ubar[0] = ubar[0] + frx_b * (p[P_CM1] - p[P_CM2] * vx);
  // front slip angle af = dl - atan2(a_f, vx)
  T af_b = ffy_b * dffy;
  T den_f = a_f * a_f + vx * vx;
  ubar[1] = ubar[1] + af_b;
  T a_f_b = -af_b * vx / den_f;
  ybar[3] = ybar[3] + af_b * a_f / den_f;
  ybar[4] = ybar[4] + a_f_b;
  ybar[5] = ybar[5] + a_f_b * lf;
  // rear slip angle ar = atan2(a_r, vx)
  T ar_b = fry_b * dfry;
  T den_r = a_r * a_r + vx * vx;
  T a_r_b = ar_b * vx / den_r;
  ybar[3] = ybar[3] - ar_b * a_r / den_r;
  ybar[4] = ybar[4] - a_r_b;
  ybar[5] = ybar[5] + a_r_b * lr;
}

// Kinematic bicycle, dynamics.py:144-173.  y=[x,y,phi,v], u=[d,delta].
template <class T>
inline void kinematic_f(const T* y, const T* u, const T* p, const T& acc, const T& mu, T* out) {
  T d = u[0], dl = u[1];
  T phi = y[2], v = y[3];
  T lf = p[P_LF], lr = p[P_LR];
  T beta = atan2(lf * tan(dl), lf + lr);
  out[0] = v * cos(phi + beta);
  out[1] = v * sin(phi + beta);
  out[2] = v * sin(beta) / lr;
  out[3] = acc * d - mu * v;
}

template <class T>
inline void kinematic_vjp(const T* y, const T* u, const T* p, const T& acc, const T& mu, const T* wv,
                          T* ybar, T* ubar) {
  T dl = u[1];
  T phi = y[2], v = y[3];
  T lf = p[P_LF], lr = p[P_LR];
  T tl = tan(dl);
  T num = lf * tl, den = lf + lr;
  T beta = atan2(num, den);
  T c = cos(phi + beta), s = sin(phi + beta), sb = sin(beta), cb = cos(beta);
  ybar[2] = ybar[2] + wv[0] * (-v * s) + wv[1] * (v * c);
  ybar[3] = ybar[3] + wv[0] * c + wv[1] * s + wv[2] * sb / lr - wv[3] * mu;
  ubar[0] = ubar[0] + wv[3] * acc;
  T beta_b = wv[0] * (-v * s) + wv[1] * (v * c) + wv[2] * v * cb / lr;
  // d beta / d dl = den/(num^2+den^2) * lf * (1 + tan^2)
  ubar[1] = ubar[1] + beta_b * den / (num * num + den * den) * lf * (T(1) + tl * tl);
}

template <class T>
struct Model {
  int model_id;
  const T* p;
  T acc, mu;
  int nx() const { return nx_of(model_id); }
  void f(const T* y, const T* u, T* out) const {
    if (model_id == MPCB_MODEL_KINEMATIC4) kinematic_f(y, u, p, acc, mu, out);
    else pacejka_f(y, u, p, out);
  }
  void vjp(const T* y, const T* u, const T* w, T* ybar, T* ubar) const {
    if (model_id == MPCB_MODEL_KINEMATIC4) kinematic_vjp(y, u, p, acc, mu, w, ybar, ubar);
    else pacejka_vjp(y, u, p, w, ybar, ubar);
  }
};

// ---------------------------------------------------------------------------------------------
// RK4 finite element (CasADi integrator 'rk', car_dynamics.py:134-147): classical RK4, step h.
// ---------------------------------------------------------------------------------------------
template <class T>
inline void rk4_element(const Model<T>& M, const T* y, const T* u, const T& h, T* yn) {
  const int nx = M.nx();
  T k1[6], k2[6], k3[6], k4[6], t[6];
  M.f(y, u, k1);
  for (int i = 0; i < nx; ++i) t[i] = y[i] + T(0.5) * h * k1[i];
  M.f(t, u, k2);
  for (int i = 0; i < nx; ++i) t[i] = y[i] + T(0.5) * h * k2[i];
  M.f(t, u, k3);
  for (int i = 0; i < nx; ++i) t[i] = y[i] + h * k3[i];
  M.f(t, u, k4);
  for (int i = 0; i < nx; ++i) yn[i] = y[i] + h / T(6) * (k1[i] + T(2) * k2[i] + T(2) * k3[i] + k4[i]);
}

// Reverse of one element: given lam = d(cost)/d(yn), overwrite lam with d(cost)/d(y), add to ubar.
template <class T>
inline void rk4_element_adj(const Model<T>& M, const T* y, const T* u, const T& h, T* lam, T* ubar) {
  const int nx = M.nx();
  T k1[6], k2[6], k3[6], y2[6], y3[6], y4[6];
  M.f(y, u, k1);
  for (int i = 0; i < nx; ++i) y2[i] = y[i] + T(0.5) * h * k1[i];
  M.f(y2, u, k2);
  for (int i = 0; i < nx; ++i) y3[i] = y[i] + T(0.5) * h * k2[i];
  M.f(y3, u, k3);
  for (int i = 0; i < nx; ++i) y4[i] = y[i] + h * k3[i];
  T kb[6], yb[6], acc[6];
  for (int i = 0; i < nx; ++i) acc[i] = lam[i];
  // k4
  for (int i = 0; i < nx; ++i) { kb[i] = h / T(6) * lam[i]; yb[i] = T(0); }
  M.vjp(y4, u, kb, yb, ubar);
  for (int i = 0; i < nx; ++i) acc[i] = acc[i] + yb[i];
  // k3
  for (int i = 0; i < nx; ++i) { kb[i] = h / T(3) * lam[i] + h * yb[i]; yb[i] = T(0); }
  M.vjp(y3, u, kb, yb, ubar);
  for (int i = 0; i < nx; ++i) acc[i] = acc[i] + yb[i];
  // k2
  for (int i = 0; i < nx; ++i) { kb[i] = h / T(3) * lam[i] + T(0.5) * h * yb[i]; yb[i] = T(0); }
  M.vjp(y2, u, kb, yb, ubar);
  for (int i = 0; i < nx; ++i) acc[i] = acc[i] + yb[i];
  // k1
  for (int i = 0; i < nx; ++i) { kb[i] = h / T(6) * lam[i] + T(0.5) * h * yb[i]; yb[i] = T(0); }
  M.vjp(y, u, kb, yb, ubar);
  for (int i = 0; i < nx; ++i) lam[i] = acc[i] + yb[i];
}

}  // namespace orc
