// mpcb_fwd.cuh — psi / grad psi of one MPC instance on one lane, forward ("tangent") mode.
//
// What it replaces: the CasADi-generated `f` / `grad_f` (reference main.py:54) over the RK4
// single-shooting rollout (car_dynamics.py:134-166) and the stage cost (car_dynamics.py:230-258).
//
// Why forward mode.  The gradient CasADi produces is the reverse sweep through 16 N model calls.  On
// the GPU that costs a second pass that re-evaluates every transcendental (or 48 KB of stored
// linearisations per instance).  The stage map s_{n+1} = Psi(s_n, u_n) has only 5 non-trivial input
// directions (vx, vy, omega, d, delta for Pacejka-6; v, d, delta for Kinematic-4):
//   * d Psi / d(x, y) is the identity (the model never reads the position), and
//   * d Psi / d phi follows from rotational equivariance, which RK4 preserves exactly (it commutes
//     with affine maps):  d Psi/d phi = (-(y+ - y), x+ - x, 1, 0, ...).
// So each model call propagates 5 (3) tangent vectors next to the primal and shares every
// transcendental with it: +45 % fp64 instructions instead of +200 %, no checkpoints.  The horizon
// adjoint is then a 20-step recursion over the stored stage Jacobians T_n (NX x ND doubles/stage):
//   grad_n = T_n[:, u]^T lam + dL/du_n,   lam <- Phi_n^T lam + dL/ds.
// Transcendentals are the short polynomial kernels of mpcb_fastmath.cuh.
#pragma once
#include <type_traits>
#include "mpcb_eval.cuh"
#include "mpcb_fastmath.cuh"

namespace mpcb {

template <int MODEL> struct Fm;

// ---- Pacejka-6 (car_dynamics.py:93-129): q = (phi, vx, vy, omega) ------------------------------
template <> struct Fm<MPCB_MODEL_PACEJKA6> {
  static constexpr int NX = 6, NQ = 4, ND = 5;  // directions: vx, vy, omega, d, delta
  static constexpr bool PIT = true;             // thin rounds use the parallel-in-time evaluation (eval_fm_pit)
  struct Lin {
    double s, c, f0, f1;
    double a11, a12, a13, a21, a22, a23, a31, a32, a33;
    double bd, bl1, bl2, bl3;
  };
  // Forces, outputs and (GRAD) the linearisation from the transcendental results: shared by both paths.
  template <bool GRAD>
  static MPCB_DEV void finish(const double* q, const StageU& su, const VehP& v, double s, double c, double a_f,
                              double a_r, double zf, double zr, double stf, double ctf, double str_, double ctr,
                              double sgn, double* kp, double* kq, Lin& L) {
    const double vx = q[1], vy = q[2], w = q[3];
    // Every multiply-add of the primal is spelled out: left to the compiler, `frx - ffy * sdl` is fused into an FMA in
    // one instantiation and built from the product `ffy * sdl` that the linearisation shares in another, and the state
    // bits would depend on whether (and how) the caller uses L (round 2: the primal-only stage of the parallel-in-time
    // evaluation must reproduce the gradient variants' states bit for bit).
    const double frx = fma(-(v.cr2 * vx), vx, fma(fma(-v.cm2, vx, v.cm1), su.d, -(v.cr0 * sgn)));
    const double ffy = v.df * stf, fry = v.dr * str_;
    kp[0] = fma(vx, c, -(vy * s));
    kp[1] = fma(vx, s, vy * c);
    kq[0] = w;
    kq[1] = fma(v.m * vy, w, fma(-ffy, su.sdl, frx)) * v.inv_m;
    kq[2] = fma(-(v.m * vx), w, fma(ffy, su.cdl, fry)) * v.inv_m;
    kq[3] = fma(ffy * v.lf, su.cdl, -(fry * v.lr)) * v.inv_iz;
    if (GRAD) {
      L.s = s; L.c = c; L.f0 = kp[0]; L.f1 = kp[1];
      // (explicit fma() throughout: the build uses -fmad=false, see __graft_entry__.py)
      const double Gf = v.df * ctf * v.cf * v.bf * fm_rcp(fma(zf, zf, 1.0));
      const double Gr = v.dr * ctr * v.cr * v.br * fm_rcp(fma(zr, zr, 1.0));
      const double vx2 = vx * vx;
      // d atan2 at the origin (standstill, no slip) is 0/0: use 0 there — the only place it can occur is the
      // first call from a standstill start, whose state tangents are discarded (the oracle's reverse sweep
      // never needs them either)
      const double dnf = fma(a_f, a_f, vx2), dnr = fma(a_r, a_r, vx2);
      const double idf = dnf > 0x1p-1000 ? fm_rcp(dnf) : 0.0, idr = dnr > 0x1p-1000 ? fm_rcp(dnr) : 0.0;
      // P*1 = G * |d atan2/d vx|, P*2 = G * d atan2/d a
      const double Pf1 = Gf * a_f * idf, Pf2 = Gf * vx * idf;
      const double Pr1 = Gr * a_r * idr, Pr2 = Gr * vx * idr;
      const double sm = su.sdl * v.inv_m, cm = su.cdl * v.inv_m, lc = v.lf * su.cdl * v.inv_iz, lri = v.lr * v.inv_iz;
      // d ffy = Pf1 dvx - Pf2 dvy - Pf2 lf dw + Gf ddelta ;  d fry = -Pr1 dvx - Pr2 dvy + Pr2 lr dw
      L.a11 = fma(-sm, Pf1, fma(-(2.0 * v.cr2), vx, -(v.cm2 * su.d)) * v.inv_m);
      L.a12 = fma(sm, Pf2, w);
      L.a13 = fma(sm * Pf2, v.lf, vy);
      L.a21 = fma(cm, Pf1, -(Pr1 * v.inv_m)) - w;
      L.a22 = fma(-cm, Pf2, -(Pr2 * v.inv_m));
      L.a23 = fma(-(cm * Pf2), v.lf, (Pr2 * v.lr) * v.inv_m) - vx;
      L.a31 = fma(lc, Pf1, lri * Pr1);
      L.a32 = fma(-lc, Pf2, lri * Pr2);
      L.a33 = fma(-(lc * Pf2), v.lf, -((lri * Pr2) * v.lr));
      L.bd = fma(-v.cm2, vx, v.cm1) * v.inv_m;
      const double gl = fma(Gf, su.cdl, -(ffy * su.sdl));
      L.bl1 = -fma(Gf, su.sdl, ffy * su.cdl) * v.inv_m;
      L.bl2 = gl * v.inv_m;
      L.bl3 = gl * v.lf * v.inv_iz;
    }
  }
  // Fast path: one branch-free block; the front and rear tyre chains run interleaved (ILP 2) next to
  // the heading sincos.  Preconditions (checked by call): 0 < vx normal, |a_f|,|a_r| <= vx (SMALL:
  // <= vx/2), su.z2_mode allows the second-stage kernels, |phi| < 2^17.
  template <bool GRAD, bool SMALL>
  static MPCB_DEV void call_fast(const double* q, const StageU& su, const VehP& v, double a_f, double a_r, double* kp,
                                 double* kq, Lin& L) {
    double s, c;
    fm_sincos_nr(q[0], &s, &c);
    const double r = fm_rcp(q[1]);
    double thf, ar;
    fm_atan_nr_x2<SMALL>(a_f * r, a_r * r, &thf, &ar);  // atan2(a, vx) = atan(a / vx) for vx > 0
    const double af = su.dl - thf;
    const double zf = v.bf * af, zr = v.br * ar;
    double uf, ur;
    fm_atan_nr_x2<SMALL>(zf, zr, &uf, &ur);
    double stf, ctf = 0.0, str_, ctr = 0.0;
    if (GRAD) fm_sincos_nr_x2(v.cf * uf, v.cr * ur, &stf, &ctf, &str_, &ctr);
    else fm_sin_nr_x2(v.cf * uf, v.cr * ur, &stf, &str_);
    finish<GRAD>(q, su, v, s, c, a_f, a_r, zf, zr, stf, ctf, str_, ctr, 1.0, kp, kq, L);
  }
  // Wide path: any slip angle and either sign of vx, still one branch-free block (full-range atan by
  // selects, the +-pi of atan2's left half-plane by a select).  Preconditions: vx normal and non-zero,
  // a_f, a_r finite, |phi| < 2^17, VehP.trig_ok.  A lane that also meets call_fast's preconditions
  // (`is_fast`) gets call_fast's result bit for bit, so a warp in which SOME lane needs the wide path runs
  // all its lanes through it once instead of running both paths back to back — line-search candidates
  // that brake to a standstill or spin out made that the common case of the solver's evaluation rounds
  // (profiles/r01_round_mode.md).
  template <bool GRAD>
  static MPCB_DEV void call_wide(const double* q, const StageU& su, const VehP& v, double a_f, double a_r, bool is_fast,
                                 double* kp, double* kq, Lin& L) {
    double s, c;
    fm_sincos_nr(q[0], &s, &c);
    const double vx = q[1];
    const double r = fm_rcp(vx);
    double thf, ar;
    fm_atan_any_x2(a_f * r, a_r * r, is_fast, &thf, &ar);
    const bool rev = vx < 0.0;  // atan2(a, vx) = atan(a / vx) + copysign(pi, a) in the left half-plane
    const double pf = copysign(2.0 * MPCB_FM_PIO2_1, a_f), pr = copysign(2.0 * MPCB_FM_PIO2_1, a_r);
    const double pfl = copysign(2.0 * MPCB_FM_PIO2_2, a_f), prl = copysign(2.0 * MPCB_FM_PIO2_2, a_r);
    thf = rev ? (thf + pf) + pfl : thf;
    ar = rev ? (ar + pr) + prl : ar;
    const double af = su.dl - thf;
    const double zf = v.bf * af, zr = v.br * ar;
    double uf, ur;
    fm_atan_any_x2(zf, zr, is_fast, &uf, &ur);
    double stf, ctf = 0.0, str_, ctr = 0.0;
    if (GRAD) fm_sincos_nr_x2(v.cf * uf, v.cr * ur, &stf, &ctf, &str_, &ctr);
    else fm_sin_nr_x2(v.cf * uf, v.cr * ur, &stf, &str_);
    finish<GRAD>(q, su, v, s, c, a_f, a_r, zf, zr, stf, ctf, str_, ctr, rev ? -1.0 : 1.0, kp, kq, L);
  }
  // General path: reversing / standstill / large slip / non-finite states.
  template <bool GRAD>
  static MPCB_DEV void call_slow(const double* q, const StageU& su, const VehP& v, double a_f, double a_r,
                                              double* kp, double* kq, Lin& L) {
    const double vx = q[1];
    double s, c;
    fm_sincos(q[0], &s, &c);
    double thf, ar;
    if (vx > 0x1p-960 && vx < 0x1p960) {
      const double r = fm_rcp(vx);
      thf = fm_atan(a_f * r);
      ar = fm_atan(a_r * r);
    } else {
      thf = atan2(a_f, vx);
      ar = atan2(a_r, vx);
    }
    const double af = su.dl - thf;
    const double sgn = vx > 0.0 ? 1.0 : (vx < 0.0 ? -1.0 : 0.0);
    const double zf = v.bf * af, zr = v.br * ar;
    double stf, ctf, str_, ctr;
    fm_sincos(v.cf * fm_atan(zf), &stf, &ctf);
    fm_sincos(v.cr * fm_atan(zr), &str_, &ctr);
    finish<GRAD>(q, su, v, s, c, a_f, a_r, zf, zr, stf, ctf, str_, ctr, sgn, kp, kq, L);
  }
  template <bool GRAD>
  static MPCB_DEV void call(const double* q, const StageU& su, const VehP& v, double* kp, double* kq, Lin& L) {
    const double vx = q[1];
    const double a_f = fma(q[3], v.lf, q[2]), a_r = fma(q[3], v.lr, -q[2]);
    const double am = fmax(fabs(a_f), fabs(a_r));
    const bool ok = vx > 0x1p-960 && vx < 0x1p960 && fabs(q[0]) < MPCB_FM_TRIG_MAX;
    const bool fast = ok && su.z2_mode == 2 && am <= 0.5 * vx;
    if (s_desc.debug >= 2) atomicAdd(&g_dbg_calls[fast ? 0 : 1], 1ull);  // MPCB_DEBUG_TIMERS=2
    const double avx = fabs(vx);
    const bool wide = !fast && avx > 0x1p-960 && avx < 0x1p960 && fabs(q[0]) < MPCB_FM_TRIG_MAX && v.trig_ok &&
                      fabs(a_f) < 0x1p1000 && fabs(a_r) < 0x1p1000;
    // The vote only picks which of two bit-identical routes the fast lanes take, so any set of voters is fine.
    const unsigned act = __activemask();
    const bool any_wide = __ballot_sync(act, wide) != 0u;
    if (s_desc.debug >= 2) {
      const bool is_nan_d = !(q[0] == q[0] && vx == vx && q[2] == q[2] && q[3] == q[3]);
      const bool slow_d = !fast && !wide && !is_nan_d;
      if (!fast) atomicAdd(&g_dbg_calls[wide ? 2 : (is_nan_d ? 3 : 4)], 1ull);
      const unsigned sl = __ballot_sync(act, slow_d);
      if ((act & (0u - act)) == (1u << (threadIdx.x & 31))) {
        atomicAdd(&g_dbg_calls[5], 1ull);
        if (any_wide) atomicAdd(&g_dbg_calls[6], 1ull);
        if (sl) atomicAdd(&g_dbg_calls[7], 1ull);
      }
    }
    if (!any_wide) {
      if (fast) call_fast<GRAD, true>(q, su, v, a_f, a_r, kp, kq, L);
    } else {
      if (fast || wide) call_wide<GRAD>(q, su, v, a_f, a_r, fast, kp, kq, L);
    }
    if (!fast && !wide) {
      // A NaN velocity/heading stays NaN through every later RK4 element and makes the cost NaN whatever the
      // other components are, so such a lane (a line-search candidate that blew up: x + q is not box-feasible
      // and the drive force feeds back on vx exponentially) is given NaN outputs at once instead of dragging
      // its warp through the library sin/cos/atan2 of the general path for the rest of the horizon.
      const bool is_nan = !(q[0] == q[0] && vx == vx && q[2] == q[2] && q[3] == q[3]);
      if (is_nan) {
        const double nn = q[0] + vx + q[2] + q[3];
        kp[0] = nn; kp[1] = nn; kq[0] = nn; kq[1] = nn; kq[2] = nn; kq[3] = nn;
        if (GRAD) {
          L.s = nn; L.c = nn; L.f0 = nn; L.f1 = nn; L.a11 = nn; L.a12 = nn; L.a13 = nn; L.a21 = nn; L.a22 = nn; L.a23 = nn;
          L.a31 = nn; L.a32 = nn; L.a33 = nn; L.bd = nn; L.bl1 = nn; L.bl2 = nn; L.bl3 = nn;
        }
      } else {
        call_slow<GRAD>(q, su, v, a_f, a_r, kp, kq, L);  // standstill, infinite or huge states
      }
    }
  }
  // The model call of a lone thread's primal chain (phase 1 of eval_fm_pit): call_fast's branch-free block is started
  // at once and the routing test of `call` is evaluated beside it instead of in front of it; an input outside the
  // fast range is then evaluated again by `call`.  Bits: call<GRAD>'s.
  template <bool GRAD>
  static MPCB_DEV void call_primal(const double* q, const StageU& su, const VehP& v, double* kp, double* kq) {
    const double vx = q[1];
    const double a_f = fma(q[3], v.lf, q[2]), a_r = fma(q[3], v.lr, -q[2]);
    const double am = fmax(fabs(a_f), fabs(a_r));
    const bool fast = vx > 0x1p-960 && vx < 0x1p960 && fabs(q[0]) < MPCB_FM_TRIG_MAX && su.z2_mode == 2 && am <= 0.5 * vx;
    Lin L;
    call_fast<GRAD, true>(q, su, v, a_f, a_r, kp, kq, L);
    if (!fast) call<GRAD>(q, su, v, kp, kq, L);
  }
  // The same with the linearisation kept (the split evaluation's lanes: few warps per SM there too).
  template <bool GRAD>
  static MPCB_DEV void call_spec(const double* q, const StageU& su, const VehP& v, double* kp, double* kq, Lin& L) {
    const double vx = q[1];
    const double a_f = fma(q[3], v.lf, q[2]), a_r = fma(q[3], v.lr, -q[2]);
    const double am = fmax(fabs(a_f), fabs(a_r));
    const bool fast = vx > 0x1p-960 && vx < 0x1p960 && fabs(q[0]) < MPCB_FM_TRIG_MAX && su.z2_mode == 2 && am <= 0.5 * vx;
    call_fast<GRAD, true>(q, su, v, a_f, a_r, kp, kq, L);
    if (!fast) call<GRAD>(q, su, v, kp, kq, L);
  }
  // tangent of one direction J at the same stage point: tq = (phi', vx', vy', w')
  template <int J>
  static MPCB_DEV void tan(const Lin& L, const double* tq, double* tkp, double* tkq) {
    tkp[0] = fma(L.c, tq[1], fma(-L.s, tq[2], -L.f1 * tq[0]));
    tkp[1] = fma(L.s, tq[1], fma(L.c, tq[2], L.f0 * tq[0]));
    tkq[0] = tq[3];
    double k1 = fma(L.a11, tq[1], fma(L.a12, tq[2], L.a13 * tq[3]));
    double k2 = fma(L.a21, tq[1], fma(L.a22, tq[2], L.a23 * tq[3]));
    double k3 = fma(L.a31, tq[1], fma(L.a32, tq[2], L.a33 * tq[3]));
    if (J == 3) k1 += L.bd;
    if (J == 4) { k1 += L.bl1; k2 += L.bl2; k3 += L.bl3; }
    tkq[1] = k1; tkq[2] = k2; tkq[3] = k3;
  }
  // unit seed of direction J at the stage start
  template <int J>
  static MPCB_DEV void seed(double* tq) {
    tq[0] = 0.0; tq[1] = J == 0 ? 1.0 : 0.0; tq[2] = J == 1 ? 1.0 : 0.0; tq[3] = J == 2 ? 1.0 : 0.0;
  }
  // the same with the direction as a run-time value (split evaluation: one direction per lane)
  static MPCB_DEV void tan_rt(const Lin& L, const double* tq, int J, double* tkp, double* tkq) {
    tkp[0] = fma(L.c, tq[1], fma(-L.s, tq[2], -L.f1 * tq[0]));
    tkp[1] = fma(L.s, tq[1], fma(L.c, tq[2], L.f0 * tq[0]));
    tkq[0] = tq[3];
    double k1 = fma(L.a11, tq[1], fma(L.a12, tq[2], L.a13 * tq[3]));
    double k2 = fma(L.a21, tq[1], fma(L.a22, tq[2], L.a23 * tq[3]));
    double k3 = fma(L.a31, tq[1], fma(L.a32, tq[2], L.a33 * tq[3]));
    k1 += J == 3 ? L.bd : (J == 4 ? L.bl1 : 0.0);
    k2 += J == 4 ? L.bl2 : 0.0;
    k3 += J == 4 ? L.bl3 : 0.0;
    tkq[1] = k1; tkq[2] = k2; tkq[3] = k3;
  }
  static MPCB_DEV void seed_rt(double* tq, int J) {
    tq[0] = 0.0; tq[1] = J == 0 ? 1.0 : 0.0; tq[2] = J == 1 ? 1.0 : 0.0; tq[3] = J == 2 ? 1.0 : 0.0;
  }
};

// ---- Kinematic-4 (dynamics.py:144-173): q = (phi, v), directions v, d, delta ---------------------
template <> struct Fm<MPCB_MODEL_KINEMATIC4> {
  static constexpr int NX = 4, NQ = 2, ND = 3;
  // No parallel-in-time evaluation for this model.  Measured (B200, config 3: 8192 instances, N = 40, obstacles, Bezier
  // roads): merely instantiating eval_fm_pit<KINEMATIC4> inside k_eval_requests slowed the kernel's OTHER paths by 58 %
  // (evaluation launches 1.18 -> 1.86 s per batch, the psi-only bulk rounds included; same registers, same stack) — a
  // code-generation interaction, not the algorithm — while the thin rounds gained 6 %.  The model call is cheap here
  // (one sincos): the latency of a request is in the stage costs, not in a dependent chain of model calls.
  static constexpr bool PIT = false;
  struct Lin { double s, c, v, sb_lr, cb_lr_db, db, acc, mu; };
  template <bool GRAD>
  static MPCB_DEV void call(const double* q, const StageU& su, const VehP& v, double* kp, double* kq, Lin& L) {
    double s, c;
    fm_sincos(q[0] + su.beta, &s, &c);
    kp[0] = q[1] * c;
    kp[1] = q[1] * s;
    kq[0] = (q[1] * su.sb) * v.inv_lr;
    kq[1] = fma(v.acc, su.d, -(v.mu * q[1]));  // explicit, as in the Pacejka model: the same bits in every instantiation
    if (GRAD) {
      L.s = s; L.c = c; L.v = q[1]; L.sb_lr = su.sb * v.inv_lr; L.cb_lr_db = su.cb * v.inv_lr * su.dbeta;
      L.db = su.dbeta; L.acc = v.acc; L.mu = v.mu;
    }
  }
  template <int J>
  static MPCB_DEV void tan(const Lin& L, const double* tq, double* tkp, double* tkq) {
    double ta = tq[0];
    if (J == 2) ta += L.db;  // beta' = dbeta/ddelta
    tkp[0] = fma(L.c, tq[1], -L.v * L.s * ta);
    tkp[1] = fma(L.s, tq[1], L.v * L.c * ta);
    double k0 = tq[1] * L.sb_lr;
    if (J == 2) k0 = fma(L.v, L.cb_lr_db, k0);
    double k1 = -L.mu * tq[1];
    if (J == 1) k1 += L.acc;
    tkq[0] = k0; tkq[1] = k1;
  }
  template <int J>
  static MPCB_DEV void seed(double* tq) { tq[0] = 0.0; tq[1] = J == 0 ? 1.0 : 0.0; }
  static MPCB_DEV void tan_rt(const Lin& L, const double* tq, int J, double* tkp, double* tkq) {
    double ta = tq[0] + (J == 2 ? L.db : 0.0);
    tkp[0] = fma(L.c, tq[1], -L.v * L.s * ta);
    tkp[1] = fma(L.s, tq[1], L.v * L.c * ta);
    double k0 = tq[1] * L.sb_lr;
    k0 = J == 2 ? fma(L.v, L.cb_lr_db, k0) : k0;
    double k1 = -L.mu * tq[1];
    k1 += J == 1 ? L.acc : 0.0;
    tkq[0] = k0; tkq[1] = k1;
  }
  static MPCB_DEV void seed_rt(double* tq, int J) { tq[0] = 0.0; tq[1] = J == 0 ? 1.0 : 0.0; }
};

// Compile-time loop over tangent directions (J must be a constant inside Fm::tan / Fm::seed).
template <int J, int ND> struct ForDir {
  template <class F> static MPCB_DEV void run(F&& f) {
    f(std::integral_constant<int, J>{});
    ForDir<J + 1, ND>::run(f);
  }
};
template <int ND> struct ForDir<ND, ND> {
  template <class F> static MPCB_DEV void run(F&&) {}
};

// One RK4 element (car_dynamics.py:134-147) on p = (x, y), q = (phi, velocities), with tangents
// tp[j] = (x', y'), tq[j] = q' of every direction j.  Position tangents never feed back, so they are
// accumulated in place.  The four stages run as a rolled loop (stage point q + a_s k, weights b_s):
// one copy of the model call instead of four keeps the kernels inside the instruction cache (the
// unrolled form stalled 0.37 issue slots per instruction on no_instruction, profiles/r01_ncu_eval_v2b).
template <int MODEL, bool GRAD>
MPCB_DEV void rk4_fm(double* p, double* q, double (*tp)[2], double (*tq)[Fm<MODEL>::NQ], const StageU& su,
                     const VehP& v, double h) {
  using M = Fm<MODEL>;
  constexpr int NQ = M::NQ, ND = M::ND;
  const double h6 = h * (1.0 / 6.0);
  double kp[2] = {0.0, 0.0}, kq[NQ], accp[2] = {0.0, 0.0}, accq[NQ];
  double tkq[GRAD ? ND : 1][NQ], taq[GRAD ? ND : 1][NQ];
#pragma unroll
  for (int i = 0; i < NQ; ++i) { kq[i] = 0.0; accq[i] = 0.0; }
  if (GRAD) {
#pragma unroll
    for (int j = 0; j < ND; ++j)
#pragma unroll
      for (int i = 0; i < NQ; ++i) { tkq[j][i] = 0.0; taq[j][i] = 0.0; }
  }
#pragma unroll 1
  for (int s = 0; s < 4; ++s) {
    const double a = s == 0 ? 0.0 : (s == 3 ? h : 0.5 * h);  // stage point  q + a k_prev
    const double bw = (s == 0 || s == 3) ? 1.0 : 2.0;        // weight of k_s
    double sq[NQ];
#pragma unroll
    for (int i = 0; i < NQ; ++i) sq[i] = fma(a, kq[i], q[i]);
    typename M::Lin L;
    M::template call<GRAD>(sq, su, v, kp, kq, L);
    accp[0] = fma(bw, kp[0], accp[0]);
    accp[1] = fma(bw, kp[1], accp[1]);
#pragma unroll
    for (int i = 0; i < NQ; ++i) accq[i] = fma(bw, kq[i], accq[i]);
    if (GRAD) {
      const double wp = bw * h6;
      ForDir<0, ND>::run([&](auto J) {
        constexpr int j = decltype(J)::value;
        double tsq[NQ], tkp[2], tk[NQ];
#pragma unroll
        for (int i = 0; i < NQ; ++i) tsq[i] = fma(a, tkq[j][i], tq[j][i]);
        M::template tan<j>(L, tsq, tkp, tk);
        tp[j][0] = fma(wp, tkp[0], tp[j][0]);
        tp[j][1] = fma(wp, tkp[1], tp[j][1]);
#pragma unroll
        for (int i = 0; i < NQ; ++i) { tkq[j][i] = tk[i]; taq[j][i] = fma(bw, tk[i], taq[j][i]); }
      });
    }
  }
  p[0] = fma(h6, accp[0], p[0]);
  p[1] = fma(h6, accp[1], p[1]);
#pragma unroll
  for (int i = 0; i < NQ; ++i) q[i] = fma(h6, accq[i], q[i]);
  if (GRAD) {
#pragma unroll
    for (int j = 0; j < ND; ++j)
#pragma unroll
      for (int i = 0; i < NQ; ++i) tq[j][i] = fma(h6, taq[j][i], tq[j][i]);
  }
}

// Doubles of scratch one gradient evaluation needs per lane: per stage T (NX x ND), dx, dy, dL/ds (NX).
template <int MODEL>
__host__ __device__ constexpr int fm_stage_doubles() { return Fm<MODEL>::NX * Fm<MODEL>::ND + 2 + Fm<MODEL>::NX; }

// One horizon stage: n_sub RK4 elements from pq (in/out, NX doubles in local memory) under input (ud, udl);
// with GRAD and store the stage Jacobian T and (dx, dy) go straight to the scratch at offset o.  Not
// inlined: the 30 tangent components + RK4 temporaries need the whole register file to themselves
// (242 registers, no spills, tools/ubench_model.cu); inlined into the evaluation loop they spilled.
template <int MODEL, bool GRAD>
__device__ __noinline__ void stage_fm(double* pq, double ud, double udl, Vec scr, int o, bool store) {
  using M = Fm<MODEL>;
  constexpr int NX = M::NX, NQ = M::NQ, ND = M::ND;
  const VehP& v = s_veh;
  StageU su;
  make_stage_u<MODEL>(su, ud, udl, v, GRAD);
  double p[2] = {pq[0], pq[1]}, q[NQ];
#pragma unroll
  for (int i = 0; i < NQ; ++i) q[i] = pq[2 + i];
  double tp[GRAD ? ND : 1][2], tq[GRAD ? ND : 1][NQ];
  if (GRAD) {
    ForDir<0, ND>::run([&](auto J) {
      constexpr int j = decltype(J)::value;
      tp[j][0] = 0.0; tp[j][1] = 0.0;
      M::template seed<j>(tq[j]);
    });
  }
  const double px0 = p[0], py0 = p[1];
  const int n_sub = s_desc.n_sub;
  const double h = s_desc.h;
#pragma unroll 1
  for (int e = 0; e < n_sub; ++e) rk4_fm<MODEL, GRAD>(p, q, tp, tq, su, v, h);
  pq[0] = p[0]; pq[1] = p[1];
#pragma unroll
  for (int i = 0; i < NQ; ++i) pq[2 + i] = q[i];
  if (GRAD && store) {
#pragma unroll
    for (int j = 0; j < ND; ++j) {
      scr.set(o + j * NX, tp[j][0]);
      scr.set(o + j * NX + 1, tp[j][1]);
#pragma unroll
      for (int i = 0; i < NQ; ++i) scr.set(o + j * NX + 2 + i, tq[j][i]);
    }
    scr.set(o + NX * ND, p[0] - px0);
    scr.set(o + NX * ND + 1, p[1] - py0);
  }
}

// Deliberately NOT inlined: the evaluation gets its own register allocation (the solver's state machine
// around it kept ~60 values live and pushed the scan and model loops into local-memory reloads), and
// the kernels stay small enough for the instruction cache.  Reads the problem from s_desc / s_veh
// (stage_problem()).
template <int MODEL, bool GRAD>
__device__ __noinline__ double eval_fm(RoadView r, Extras ex, const double* x0, Vec U, Vec scr, Vec grad, bool store,
                                       Vec states = Vec{nullptr, 0}, IVec idx_out = IVec{nullptr, 0}, bool cut = false) {
  const KDesc& d = s_desc;
  using M = Fm<MODEL>;
  constexpr int NX = M::NX, ND = M::ND, K = fm_stage_doubles<MODEL>();
  double y[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) y[i] = x0[i];
  double cost = 0.0;
  int idx = -1;  // nearest vertex of the previous stage: where the next search starts
  // `cut` (the solver's evaluations only): a lane whose state holds a NaN, or a velocity component whose square
  // overflows (the speed term of the cost is then +Inf, so psi is +Inf or NaN), stops here with psi = NaN and a NaN
  // gradient.  Only line-search points x + (1 - tau) p + tau q can get there (they are not box-feasible) and PANOC
  // rejects a point whose psi is +Inf exactly as one whose psi is NaN (phi+ is not finite, every test on it is
  // false), so the iterates are unchanged; without the cut such a lane kept its whole warp in the slow routes of
  // the model call, the vertex scan and wrap_to_pi for the rest of the horizon (2.4x the time of a normal chunk,
  // and an evaluation launch lasts as long as its slowest chunk: profiles/r01_round_mode.md).
  bool dead = false;
  double ud_next = U(0), udl_next = U(1);  // the next stage's inputs are requested one stage ahead (DRAM latency)
#pragma unroll 1
  for (int n = 0; n < d.N; ++n) {
    if (dead) continue;
    const double ud = ud_next, udl = udl_next;
    if (n + 1 < d.N) { ud_next = U(2 * n + 2); udl_next = U(2 * n + 3); }
    const int o = n * K;
    stage_fm<MODEL, GRAD>(y, ud, udl, scr, o, store);
    if (cut) {
      bool sane = true;
#pragma unroll
      for (int i = 0; i < NX; ++i) sane = sane && y[i] == y[i];
      sane = sane && fabs(y[3]) < 1.3e154 && (MODEL == MPCB_MODEL_KINEMATIC4 || fabs(y[NX == 6 ? 4 : 3]) < 1.3e154);
      if (!sane) { dead = true; continue; }
    }
    double gy[NX], gu[2];
    cost += stage_cost<MODEL, GRAD, true>(d, r, ex, n, y, ud, udl, &idx, gy, gu);
    if (states.p) {
#pragma unroll
      for (int i = 0; i < NX; ++i) states.set(n * NX + i, y[i]);
    }
    if (idx_out.p) idx_out.set(n, idx);
    if (GRAD && store) {
#pragma unroll
      for (int i = 0; i < NX; ++i) scr.set(o + NX * ND + 2 + i, gy[i]);
    }
  }
  if (dead) {
    const double nn = nan("");
    if (GRAD && store)
      for (int i = 0; i < 2 * d.N; ++i) grad.set(i, nn);
    if (states.p)
      for (int i = 0; i < d.N * NX; ++i) states.set(i, nn);
    return nn;
  }
  if (GRAD && store) {
    // horizon adjoint over the stored stage Jacobians
    double lam[NX];
#pragma unroll
    for (int i = 0; i < NX; ++i) lam[i] = 0.0;
#pragma unroll 1
    for (int n = d.N - 1; n >= 0; --n) {
      const int o = n * K;
#pragma unroll
      for (int i = 0; i < NX; ++i) lam[i] += scr(o + NX * ND + 2 + i);
      double dots[ND];
#pragma unroll
      for (int j = 0; j < ND; ++j) {
        double s = 0.0;
#pragma unroll
        for (int i = 0; i < NX; ++i) s = fma(scr(o + j * NX + i), lam[i], s);
        dots[j] = s;
      }
      const double dx = scr(o + NX * ND), dy = scr(o + NX * ND + 1);
      const double ud = U(2 * n), udl = U(2 * n + 1);
      grad.set(2 * n, fma(2.0 * d.c[5], ud, dots[ND - 2]));
      grad.set(2 * n + 1, fma(2.0 * d.c[4], udl, dots[ND - 1]));
      lam[2] = fma(dx, lam[1], fma(-dy, lam[0], lam[2]));
#pragma unroll
      for (int j = 0; j < ND - 2; ++j) lam[3 + j] = dots[j];
    }
  }
  return cost;
}

// ---- split evaluation: one tangent direction per lane -------------------------------------------------
// A lone warp issues one DFMA per 3.75 cycles, so the psi + grad psi evaluation of ONE request is bound by its
// own instruction count (389 fp64 instructions per model call with all 5 tangents, 1 357 cycles) even when the
// GPU is otherwise idle — the state of the last rounds of every batch.  When few gradient requests are left a
// group of 8 lanes shares a request: every lane carries the primal (identical arithmetic, identical results)
// and ONE tangent direction J = lane % 8 (lanes J >= ND idle along), which cuts the instruction stream of a
// model call to primal + one tangent.  The group leader (J = 0) owns the stage-Jacobian scratch column the
// lanes write their Jacobian columns to, and runs the horizon adjoint.
template <int MODEL>
MPCB_DEV void rk4_fm_split(double* p, double* q, double* tp, double* tq, int J, const StageU& su, const VehP& v, double h) {
  using M = Fm<MODEL>;
  constexpr int NQ = M::NQ;
  const double h6 = h * (1.0 / 6.0);
  double kp[2] = {0.0, 0.0}, kq[NQ], accp[2] = {0.0, 0.0}, accq[NQ], tkq[NQ], taq[NQ];
#pragma unroll
  for (int i = 0; i < NQ; ++i) { kq[i] = 0.0; accq[i] = 0.0; tkq[i] = 0.0; taq[i] = 0.0; }
#pragma unroll 1
  for (int s = 0; s < 4; ++s) {
    const double a = s == 0 ? 0.0 : (s == 3 ? h : 0.5 * h);
    const double bw = (s == 0 || s == 3) ? 1.0 : 2.0;
    double sq[NQ];
#pragma unroll
    for (int i = 0; i < NQ; ++i) sq[i] = fma(a, kq[i], q[i]);
    typename M::Lin L;
    // (Pacejka-6: fast block before its range test, as in the lone thread's primal chain — the split lanes run with one or
    // two warps per SM sub-partition as well.  B = 4 096: evaluation launches 192 -> 177 ms per batch; unrolling the four
    // stages on top of it: 178 ms.)
    if constexpr (MODEL == MPCB_MODEL_PACEJKA6) M::template call_spec<true>(sq, su, v, kp, kq, L);
    else M::template call<true>(sq, su, v, kp, kq, L);
    accp[0] = fma(bw, kp[0], accp[0]);
    accp[1] = fma(bw, kp[1], accp[1]);
#pragma unroll
    for (int i = 0; i < NQ; ++i) accq[i] = fma(bw, kq[i], accq[i]);
    const double wp = bw * h6;
    double tsq[NQ], tkp[2], tk[NQ];
#pragma unroll
    for (int i = 0; i < NQ; ++i) tsq[i] = fma(a, tkq[i], tq[i]);
    M::tan_rt(L, tsq, J, tkp, tk);
    tp[0] = fma(wp, tkp[0], tp[0]);
    tp[1] = fma(wp, tkp[1], tp[1]);
#pragma unroll
    for (int i = 0; i < NQ; ++i) { tkq[i] = tk[i]; taq[i] = fma(bw, tk[i], taq[i]); }
  }
  p[0] = fma(h6, accp[0], p[0]);
  p[1] = fma(h6, accp[1], p[1]);
#pragma unroll
  for (int i = 0; i < NQ; ++i) q[i] = fma(h6, accq[i], q[i]);
#pragma unroll
  for (int i = 0; i < NQ; ++i) tq[i] = fma(h6, taq[i], tq[i]);
}

template <int MODEL>
__device__ __noinline__ void stage_fm_split(double* pq, double ud, double udl, Vec scr, int o, int J) {
  using M = Fm<MODEL>;
  constexpr int NX = M::NX, NQ = M::NQ, ND = M::ND;
  const VehP& v = s_veh;
  StageU su;
  make_stage_u<MODEL>(su, ud, udl, v, true);
  double p[2] = {pq[0], pq[1]}, q[NQ];
#pragma unroll
  for (int i = 0; i < NQ; ++i) q[i] = pq[2 + i];
  double tp[2] = {0.0, 0.0}, tq[NQ];
  M::seed_rt(tq, J);
  const double px0 = p[0], py0 = p[1];
  const int n_sub = s_desc.n_sub;
  const double h = s_desc.h;
#pragma unroll 1
  for (int e = 0; e < n_sub; ++e) rk4_fm_split<MODEL>(p, q, tp, tq, J, su, v, h);
  pq[0] = p[0]; pq[1] = p[1];
#pragma unroll
  for (int i = 0; i < NQ; ++i) pq[2 + i] = q[i];
  if (J < ND) {
    scr.set(o + J * NX, tp[0]);
    scr.set(o + J * NX + 1, tp[1]);
#pragma unroll
    for (int i = 0; i < NQ; ++i) scr.set(o + J * NX + 2 + i, tq[i]);
  }
  if (J == 0) {
    scr.set(o + NX * ND, p[0] - px0);
    scr.set(o + NX * ND + 1, p[1] - py0);
  }
}

// psi + grad psi of one request by a group of 8 lanes (all of them call this; `scr` is the LEADER's scratch
// column, J = lane % 8).  The gradient is written by the leader; every lane returns psi.  Same cut as eval_fm.
template <int MODEL>
__device__ __noinline__ double eval_fm_split(RoadView r, Extras ex, const double* x0, Vec U, Vec scr, Vec grad, int J,
                                             unsigned gmask) {
  const KDesc& d = s_desc;
  using M = Fm<MODEL>;
  constexpr int NX = M::NX, ND = M::ND, K = fm_stage_doubles<MODEL>();
  double y[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) y[i] = x0[i];
  double cost = 0.0;
  int idx = -1;
  bool dead = false;
  double ud_next = U(0), udl_next = U(1);
#pragma unroll 1
  for (int n = 0; n < d.N; ++n) {
    if (dead) continue;
    const double ud = ud_next, udl = udl_next;
    if (n + 1 < d.N) { ud_next = U(2 * n + 2); udl_next = U(2 * n + 3); }
    const int o = n * K;
    stage_fm_split<MODEL>(y, ud, udl, scr, o, J);
    bool sane = true;
#pragma unroll
    for (int i = 0; i < NX; ++i) sane = sane && y[i] == y[i];
    sane = sane && fabs(y[3]) < 1.3e154 && (MODEL == MPCB_MODEL_KINEMATIC4 || fabs(y[NX == 6 ? 4 : 3]) < 1.3e154);
    if (!sane && !d.exact) { dead = true; continue; }
    double gy[NX], gu[2];
    cost += stage_cost<MODEL, true, true>(d, r, ex, n, y, ud, udl, &idx, gy, gu);
    if (J == 0) {
#pragma unroll
      for (int i = 0; i < NX; ++i) scr.set(o + NX * ND + 2 + i, gy[i]);
    }
  }
  __syncwarp(gmask);  // the columns written by the other lanes of the group are visible to the leader
  if (J != 0) return dead ? nan("") : cost;
  if (dead) {
    const double nn = nan("");
    for (int i = 0; i < 2 * d.N; ++i) grad.set(i, nn);
    return nn;
  }
  double lam[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) lam[i] = 0.0;
#pragma unroll 1
  for (int n = d.N - 1; n >= 0; --n) {
    const int o = n * K;
#pragma unroll
    for (int i = 0; i < NX; ++i) lam[i] += scr(o + NX * ND + 2 + i);
    double dots[ND];
#pragma unroll
    for (int j = 0; j < ND; ++j) {
      double s = 0.0;
#pragma unroll
      for (int i = 0; i < NX; ++i) s = fma(scr(o + j * NX + i), lam[i], s);
      dots[j] = s;
    }
    const double dx = scr(o + NX * ND), dy = scr(o + NX * ND + 1);
    const double ud = U(2 * n), udl = U(2 * n + 1);
    grad.set(2 * n, fma(2.0 * d.c[5], ud, dots[ND - 2]));
    grad.set(2 * n + 1, fma(2.0 * d.c[4], udl, dots[ND - 1]));
    lam[2] = fma(dx, lam[1], fma(-dy, lam[0], lam[2]));
#pragma unroll
    for (int j = 0; j < ND - 2; ++j) lam[3 + j] = dots[j];
  }
  return cost;
}

// Primal-only stage.  GA: with the GRADIENT variant's arithmetic — the model call is Fm::call<true> (its sincos kernels, not
// the value-only variant's shorter sin kernel, whose last bits differ), inlined with the linearisation unused, so the
// compiler drops the Lin computation and every tangent: the state bits are those of stage_fm<MODEL, true> / stage_fm_split
// at about the value-only call's latency.  !GA: Fm::call<false>, the bits of stage_fm<MODEL, false>.  Phase 1 of the
// parallel-in-time evaluation of a gradient / a psi-only request.
template <int MODEL, bool GA>
MPCB_DEV void rk4_fm_primal(double* p, double* q, const StageU& su, const VehP& v, double h) {
  using M = Fm<MODEL>;
  constexpr int NQ = M::NQ;
  const double h6 = h * (1.0 / 6.0);
  double kp[2] = {0.0, 0.0}, kq[NQ], accp[2] = {0.0, 0.0}, accq[NQ];
#pragma unroll
  for (int i = 0; i < NQ; ++i) { kq[i] = 0.0; accq[i] = 0.0; }
  // Unrolled, and (Pacejka-6) with the fast block of the model call started before its range test (call_primal): the
  // end of one call overlaps the start of the next.  Measured on the lone thread of phase 1 (B200, cycles per model
  // call): rolled + routed 970, rolled + call_primal 880, unrolled + routed 1 040, unrolled + call_primal 700.
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const double a = s == 0 ? 0.0 : (s == 3 ? h : 0.5 * h);
    const double bw = (s == 0 || s == 3) ? 1.0 : 2.0;
    double sq[NQ];
#pragma unroll
    for (int i = 0; i < NQ; ++i) sq[i] = fma(a, kq[i], q[i]);
    if constexpr (MODEL == MPCB_MODEL_PACEJKA6) {
      M::template call_primal<GA>(sq, su, v, kp, kq);
    } else {
      typename M::Lin L;
      M::template call<GA>(sq, su, v, kp, kq, L);
    }
    accp[0] = fma(bw, kp[0], accp[0]);
    accp[1] = fma(bw, kp[1], accp[1]);
#pragma unroll
    for (int i = 0; i < NQ; ++i) accq[i] = fma(bw, kq[i], accq[i]);
  }
  p[0] = fma(h6, accp[0], p[0]);
  p[1] = fma(h6, accp[1], p[1]);
#pragma unroll
  for (int i = 0; i < NQ; ++i) q[i] = fma(h6, accq[i], q[i]);
}

template <int MODEL, bool GA>
__device__ __noinline__ void stage_fm_primal(double* pq, double ud, double udl) {
  using M = Fm<MODEL>;
  constexpr int NQ = M::NQ;
  const VehP& v = s_veh;
  StageU su;
  make_stage_u<MODEL>(su, ud, udl, v, true);
  double p[2] = {pq[0], pq[1]}, q[NQ];
#pragma unroll
  for (int i = 0; i < NQ; ++i) q[i] = pq[2 + i];
  const int n_sub = s_desc.n_sub;
  const double h = s_desc.h;
#pragma unroll 1
  for (int e = 0; e < n_sub; ++e) rk4_fm_primal<MODEL, GA>(p, q, su, v, h);
  pq[0] = p[0]; pq[1] = p[1];
#pragma unroll
  for (int i = 0; i < NQ; ++i) pq[2 + i] = q[i];
}

// ---- parallel-in-time evaluation: one CTA per gradient request -----------------------------------------
// Only the PRIMAL rollout is sequential.  The stage Jacobian T_n (and the stage cost and its gradient) depends on the
// stage's start state s_n and input u_n alone — not on earlier tangents; the tangents only meet in the horizon adjoint.
// When few gradient requests are left (the last rounds of every batch, a lone vehicle's tick) a whole CTA takes one:
//   phase 1  one thread rolls the primal out, s_0 .. s_N to shared memory              (16 N dependent model calls)
//   phase 2  every stage at once: group g = stage, lane J < ND = tangent direction J of that stage from s_n,
//            lane J = ND = cost and cost gradient of the stage at s_{n+1}               (16 dependent model calls)
//   phase 3  one thread sums the stage costs in order and runs the N-step adjoint over shared memory.
// The stage arithmetic is that of the split evaluation (stage_fm_primal / stage_fm_split / stage_cost: the same model
// call on the same operands), and the sums run in the same order: psi and the gradient are bit for bit those of eval_fm_split / eval_fm
// (tests/test_gpu_exactness.py).  Latency of one request: 0.33 -> 0.17 ms; no HBM scratch at all.
// Shared memory (doubles): (N + 1) NX states, 2 N inputs, N K stage rows, N stage costs, 2 flags.
template <int MODEL>
__host__ __device__ constexpr int fm_pit_doubles(int N) {
  return (N + 1) * Fm<MODEL>::NX + 2 * N + N * fm_stage_doubles<MODEL>() + N + 2;
}

// Called by every thread of the CTA; returns psi on every thread (NaN for a cut request).  `grad` is the request's
// gradient row (global, strided).
// GRAD = false: a psi-only request the same way (phase 2 = the stage costs, phase 3 = their sum in stage order; `grad`
// is then the optional row of states s_1 .. s_N, as eval_fm<MODEL, false> writes them): 20 stage costs (vertex scan,
// errors) no longer sit between the model calls of the chain.
template <int MODEL, bool GRAD>
__device__ __noinline__ double eval_fm_pit(RoadView r, Extras ex, const double* x0, Vec U, Vec grad, double* sm, bool cut) {
  const KDesc& d = s_desc;
  using M = Fm<MODEL>;
  constexpr int NX = M::NX, ND = M::ND, K = fm_stage_doubles<MODEL>();
  const int N = d.N;
  double* S = sm;                    // [(N + 1), NX]
  double* UU = S + (N + 1) * NX;     // [2 N]
  double* Jb = UU + 2 * N;           // [N, K]
  double* Lc = Jb + N * K;           // [N]
  double* flag = Lc + N;             // [0] dead, [1] psi
  for (int i = threadIdx.x; i < 2 * N; i += blockDim.x) UU[i] = U(i);
  if (threadIdx.x < NX) S[threadIdx.x] = x0[threadIdx.x];
  if (threadIdx.x == 0) flag[0] = 0.0;
  __syncthreads();
  // ---- phase 1: the primal, sequential
  if (threadIdx.x == 0) {
    double y[NX];
#pragma unroll
    for (int i = 0; i < NX; ++i) y[i] = S[i];
#pragma unroll 1
    for (int n = 0; n < N; ++n) {
      stage_fm_primal<MODEL, GRAD>(y, UU[2 * n], UU[2 * n + 1]);
      bool sane = true;
#pragma unroll
      for (int i = 0; i < NX; ++i) sane = sane && y[i] == y[i];
      sane = sane && fabs(y[3]) < 1.3e154 && (MODEL == MPCB_MODEL_KINEMATIC4 || fabs(y[NX == 6 ? 4 : 3]) < 1.3e154);
      if (!sane && cut) { flag[0] = 1.0; break; }
#pragma unroll
      for (int i = 0; i < NX; ++i) S[(n + 1) * NX + i] = y[i];
    }
  }
  __syncthreads();
  const bool dead = flag[0] != 0.0;
  // ---- phase 2: every stage at once
  if (!dead) {
    if (GRAD) {
      const int J = threadIdx.x & 7;
      if (J < ND) {
        for (int n = threadIdx.x >> 3; n < N; n += blockDim.x >> 3) {
          double y[NX];
#pragma unroll
          for (int i = 0; i < NX; ++i) y[i] = S[n * NX + i];
          stage_fm_split<MODEL>(y, UU[2 * n], UU[2 * n + 1], Vec{Jb, 1}, n * K, J);
        }
      }
    }
    // stage costs: from the LAST lanes of the CTA backwards, so that for the usual horizons they run in warps of their
    // own beside the tangent warps instead of after them in the same warp
    for (int n = (int)(blockDim.x - 1u - threadIdx.x); n < N; n += blockDim.x) {
      double y[NX], gy[NX], gu[2];
#pragma unroll
      for (int i = 0; i < NX; ++i) y[i] = S[(n + 1) * NX + i];
      int idx = -1;
      Lc[n] = stage_cost<MODEL, GRAD, true>(d, r, ex, n, y, UU[2 * n], UU[2 * n + 1], &idx, gy, gu);
      if (GRAD) {
#pragma unroll
        for (int i = 0; i < NX; ++i) Jb[n * K + NX * ND + 2 + i] = gy[i];
      }
    }
  }
  __syncthreads();
  // ---- phase 3: cost sum and horizon adjoint
  if (!GRAD) {
    if (grad.p) {
      const double nn = nan("");
      for (int i = threadIdx.x; i < N * NX; i += blockDim.x) grad.set(i, dead ? nn : S[NX + i]);
    }
    if (threadIdx.x == 0) {
      double cost = 0.0;
      for (int n = 0; n < N; ++n) cost += Lc[n];
      flag[1] = dead ? nan("") : cost;
    }
  } else if (threadIdx.x == 0) {
    if (dead) {
      const double nn = nan("");
      for (int i = 0; i < 2 * N; ++i) grad.set(i, nn);
      flag[1] = nn;
    } else {
      double cost = 0.0;
      for (int n = 0; n < N; ++n) cost += Lc[n];
      const Vec scr{Jb, 1};
      double lam[NX];
#pragma unroll
      for (int i = 0; i < NX; ++i) lam[i] = 0.0;
#pragma unroll 1
      for (int n = N - 1; n >= 0; --n) {
        const int o = n * K;
#pragma unroll
        for (int i = 0; i < NX; ++i) lam[i] += scr(o + NX * ND + 2 + i);
        double dots[ND];
#pragma unroll
        for (int j = 0; j < ND; ++j) {
          double sacc = 0.0;
#pragma unroll
          for (int i = 0; i < NX; ++i) sacc = fma(scr(o + j * NX + i), lam[i], sacc);
          dots[j] = sacc;
        }
        const double dx = scr(o + NX * ND), dy = scr(o + NX * ND + 1);
        const double ud = UU[2 * n], udl = UU[2 * n + 1];
        grad.set(2 * n, fma(2.0 * d.c[5], ud, dots[ND - 2]));
        grad.set(2 * n + 1, fma(2.0 * d.c[4], udl, dots[ND - 1]));
        lam[2] = fma(dx, lam[1], fma(-dy, lam[0], lam[2]));
#pragma unroll
        for (int j = 0; j < ND - 2; ++j) lam[3 + j] = dots[j];
      }
      flag[1] = cost;
    }
  }
  __syncthreads();
  const double psi = flag[1];
  __syncthreads();  // the buffers are reused by the CTA's next request
  return psi;
}

}  // namespace mpcb
