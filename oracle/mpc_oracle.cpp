// mpc_oracle.cpp — CPU ORACLE shared library (test infrastructure, NOT product code).
// Exposes the restatement in orc_*.hpp through a small C API for ctypes (tests/, smoke(), and
// bench.py's cpu_baseline / --impl reference legs).  Build: see oracle/Makefile.
#include <omp.h>
#include <cstdio>
#include "orc_solver.hpp"
#include "orc_game.hpp"

using namespace orc;

// ---- counting scalar: every arithmetic op / compare / transcendental = 1 (SURVEY §8d "strict") ----
struct Cnt {
  double v;
  static thread_local long long arith, transc;
  Cnt() : v(0) {}
  Cnt(double x) : v(x) {}
};
thread_local long long Cnt::arith = 0;
thread_local long long Cnt::transc = 0;
#define CNT_BIN(op) \
  inline Cnt operator op(const Cnt& a, const Cnt& b) { Cnt::arith++; return Cnt(a.v op b.v); }
CNT_BIN(+) CNT_BIN(-) CNT_BIN(*) CNT_BIN(/)
inline Cnt operator-(const Cnt& a) { return Cnt(-a.v); }
#define CNT_CMP(op) \
  inline bool operator op(const Cnt& a, const Cnt& b) { Cnt::arith++; return a.v op b.v; }
CNT_CMP(<) CNT_CMP(>) CNT_CMP(<=) CNT_CMP(>=) CNT_CMP(==) CNT_CMP(!=)
#define CNT_FN1(fn) \
  inline Cnt fn(const Cnt& a) { Cnt::transc++; return Cnt(std::fn(a.v)); }
CNT_FN1(sin) CNT_FN1(cos) CNT_FN1(tan) CNT_FN1(atan) CNT_FN1(sqrt) CNT_FN1(exp)
inline Cnt atan2(const Cnt& a, const Cnt& b) { Cnt::transc++; return Cnt(std::atan2(a.v, b.v)); }
inline Cnt fmod(const Cnt& a, const Cnt& b) { Cnt::transc++; return Cnt(std::fmod(a.v, b.v)); }
inline Cnt pow(const Cnt& a, const Cnt& b) { Cnt::transc++; return Cnt(std::pow(a.v, b.v)); }

namespace {

template <class T>
Instance<T> make_instance(const MpcProblemDesc* d, const T* p, const T* road, const T* obs, const T* lam,
                          const T* Sig) {
  Instance<T> I;
  I.d = d;
  I.M.model_id = d->model_id;
  I.M.p = p;
  I.M.acc = T(d->kin_accel);
  I.M.mu = T(d->kin_friction);
  I.road = road;
  I.obs = obs;
  I.lam = lam;
  I.Sig = Sig;
  return I;
}

// Resolve the polyline of instance b for the road_mode; Bezier roads are sampled into `tmp`.
const double* road_of(const MpcProblemDesc* d, const double* road, const int32_t* road_idx, int64_t b,
                      std::vector<double>& tmp) {
  const int S = d->S;
  switch (d->road_mode) {
    case MPCB_ROAD_SHARED: return road;
    case MPCB_ROAD_TABLE: return road + size_t(road_idx ? road_idx[b] : 0) * 2 * S;
    case MPCB_ROAD_PER_INSTANCE: return road + size_t(b) * 2 * S;
    case MPCB_ROAD_BEZIER:
      tmp.resize(2 * S);
      bezier_sample(road + size_t(b) * 12, S, tmp.data());
      return tmp.data();
  }
  return road;
}

bool is_bounded(const MpcProblemDesc* d) {
  if ((d->flags & MPCB_FLAG_SAFE_DIST) && d->n_obs > 0) return true;
  if (!(d->flags & MPCB_FLAG_BOUNDED_D)) return false;
  for (int i = 0; i < nx_of(d->model_id); ++i)
    if (!std::isinf(d->D_lower[i]) || !std::isinf(d->D_upper[i])) return true;
  return false;
}

}  // namespace

extern "C" {

int orc_num_threads() { return omp_get_max_threads(); }

void orc_dynamics(int model_id, const double* p, double acc, double mu, const double* x, const double* u,
                  int clip_inputs, double* out) {
  Model<double> M{model_id, p, acc, mu};
  double uu[2] = {u[0], u[1]};
  if (clip_inputs) {  // dynamics.py:56-66,99
    uu[0] = clip(uu[0], -p[P_MAX_DRIVE], p[P_MAX_DRIVE]);
    uu[1] = clip(uu[1], -p[P_MAX_STEER], p[P_MAX_STEER]);
  }
  M.f(x, uu, out);
}

void orc_dynamics_vjp(int model_id, const double* p, double acc, double mu, const double* x, const double* u,
                      const double* w, double* xbar, double* ubar) {
  Model<double> M{model_id, p, acc, mu};
  for (int i = 0; i < 6; ++i) xbar[i] = 0;
  ubar[0] = ubar[1] = 0;
  M.vjp(x, u, w, xbar, ubar);
}

void orc_rk4_stage(const MpcProblemDesc* d, const double* p, const double* x, const double* u, double* out) {
  Instance<double> I = make_instance<double>(d, p, nullptr, nullptr, nullptr, nullptr);
  double y[6], yn[6];
  const int nx = I.M.nx();
  for (int i = 0; i < nx; ++i) y[i] = x[i];
  double h = d->Ts / d->n_sub;
  for (int e = 0; e < d->n_sub; ++e) {
    rk4_element(I.M, y, u, h, yn);
    for (int i = 0; i < nx; ++i) y[i] = yn[i];
  }
  for (int i = 0; i < nx; ++i) out[i] = y[i];
}

double orc_wrap_to_pi(double a) { return wrap_to_pi(a); }

int orc_road_errors(int S, const double* road, double px, double py, double heading, int normalised,
                    double* err3) {
  return road_errors(S, road, px, py, heading, normalised != 0, err3);
}

void orc_bezier_sample(const double* P12, int S, double* road) { bezier_sample(P12, S, road); }
void orc_bezier_point(const double* P12, double j, double* xy) { bezier_point(P12, j, &xy[0], &xy[1]); }
void orc_bezier_control_points(double i, double h, double W, double theta, double l, double Lf, double v0,
                               double v1, double D1, double* P12, double* tca) {
  bezier_control_points(i, h, W, theta, l, Lf, v0, v1, D1, P12, tca);
}

void orc_safe_distances(const MpcProblemDesc* d, const double* Xe, const double* Xo, double* out2) {  // dpf_test.py:57-64
  safe_distances(*d, Xe[2], Xe[3], Xo[2], Xo[3], &out2[0], &out2[1]);
}
double orc_u_trgl(double y, double y_t, double a) { return u_trgl(y, y_t, a); }
double orc_u_rbd(double y, double y_br, double y_bl, double b) { return u_rbd(y, y_br, y_bl, b); }
double orc_dnf(double x, double y, double phi, double v, int n_obs, const double* xo, const double* yo,
               const double* po, const double* vo, double Af, double b, double sx, double sy) {
  double U = 0;
  for (int i = 0; i < n_obs; ++i)
    U += dnf_one(x, y, phi, v, xo[i], yo[i], po[i], vo[i], Af, b, sx, sy, 5.0, (double*)nullptr);
  return U;
}
double orc_event_state(double x, double ux, double a, double t0, double t1) {  // event_triggered.py:26
  return x + ux * (t0 + t1) + 1.0 / 2 * a * (t0 + t1) * (t0 + t1);
}

// ---- batched evaluation -------------------------------------------------------------------------
void orc_rollout_batch(const MpcProblemDesc* d, int64_t B, const double* x0, const double* U, const double* p,
                       const double* road, const int32_t* road_idx, const double* obs, double* states,
                       int32_t* idx, double* cost) {
  const int nx = nx_of(d->model_id), N = d->N;
#pragma omp parallel
  {
    std::vector<double> tmp;
    std::vector<int> ib(N);
#pragma omp for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
      const double* r = road_of(d, road, road_idx, b, tmp);
      Instance<double> I = make_instance<double>(d, p, r, obs ? obs + b * d->n_obs * MPCB_OBS_DIM : nullptr,
                                                 nullptr, nullptr);
      double c = rollout(I, x0 + b * nx, U + b * 2 * N, states ? states + b * N * nx : nullptr, ib.data(),
                         (double*)nullptr);
      if (cost) cost[b] = c;
      if (idx) for (int n = 0; n < N; ++n) idx[b * N + n] = ib[n];
    }
  }
}

void orc_value_grad_batch(const MpcProblemDesc* d, int64_t B, const double* x0, const double* U, const double* p,
                          const double* road, const int32_t* road_idx, const double* obs, const double* lam,
                          const double* Sig, double* cost, double* grad) {
  const int nx = nx_of(d->model_id), N = d->N;
#pragma omp parallel
  {
    std::vector<double> tmp;
#pragma omp for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
      const double* r = road_of(d, road, road_idx, b, tmp);
      Instance<double> I = make_instance<double>(d, p, r, obs ? obs + b * d->n_obs * MPCB_OBS_DIM : nullptr,
                                                 lam ? lam + b * rows_of(*d) * N : nullptr, Sig ? Sig + b * rows_of(*d) * N : nullptr);
      cost[b] = value_grad(I, x0 + b * nx, U + b * 2 * N, grad + b * 2 * N);
    }
  }
}

// Fixed-point residual of the forward-backward map at a GIVEN step size (Appendix A 2, 3b):
//   r = || Pi_C(U - gamma grad psi(U)) - U ||_2,  psi_out = psi(U)      (unbounded D: psi = f).
// The certificate tests/test_gpu_exactness.py issues for every GPU answer, with the GPU's own final gamma.
void orc_fb_residual_batch(const MpcProblemDesc* d, int64_t B, const double* x0, const double* U, const double* p,
                           const double* road, const int32_t* road_idx, const double* obs, const double* lb,
                           const double* ub, const double* gamma, double* res_out, double* psi_out) {
  const int nx = nx_of(d->model_id), N = d->N, n = 2 * N;
#pragma omp parallel
  {
    std::vector<double> tmp, g(n);
#pragma omp for schedule(static)
    for (int64_t b = 0; b < B; ++b) {
      const double* r = road_of(d, road, road_idx, b, tmp);
      Instance<double> I = make_instance<double>(d, p, r, obs ? obs + b * d->n_obs * MPCB_OBS_DIM : nullptr, nullptr, nullptr);
      const double* u = U + b * n;
      const double psi = value_grad(I, x0 + b * nx, u, g.data());
      double s = 0;
      for (int i = 0; i < n; ++i) {
        double v = u[i] - gamma[b] * g[i];
        v = v < lb[i] ? lb[i] : (v > ub[i] ? ub[i] : v);
        s += (v - u[i]) * (v - u[i]);
      }
      res_out[b] = std::sqrt(s);
      if (psi_out) psi_out[b] = psi;
    }
  }
}

// game_theory.py:205-244 for a batch: ego [B,3] and cars [B,M,3] as (x, v, lane); payoff [B,2] for target lanes 1, 2;
// decision [B] = 1 when lane 2 pays more (game_theory.py:377-379); ctrl [B,12] / tca [B] = get_bezier_control_points
// towards the car in front in ego's lane (NaN when there is none or the speeds are equal).
void orc_game_payoff_batch(int64_t B, int M, const double* ego, const double* cars, const double* car_par11,
                           double* payoff, int32_t* decision, double* ctrl, double* tca) {
  CarPar c = car_defaults();
  if (car_par11) c = CarPar{car_par11[0], car_par11[1], car_par11[2], car_par11[3], car_par11[4], car_par11[5],
                            car_par11[6], car_par11[7], car_par11[8], car_par11[9], car_par11[10]};
  for (int64_t b = 0; b < B; ++b) {
    CarRec e{ego[3 * b], ego[3 * b + 1], (int)ego[3 * b + 2]};
    std::vector<CarRec> cs(M);
    for (int k = 0; k < M; ++k) cs[k] = CarRec{cars[(b * M + k) * 3], cars[(b * M + k) * 3 + 1], (int)cars[(b * M + k) * 3 + 2]};
    payoff[2 * b] = total_payoff(c, e, cs.data(), M, 1);
    payoff[2 * b + 1] = total_payoff(c, e, cs.data(), M, 2);
    if (decision) decision[b] = payoff[2 * b + 1] > payoff[2 * b] ? 1 : 0;
    if (ctrl) {
      const int f = car_in_front(e, cs.data(), M, -1, nullptr, e.lane);
      if (f >= 0 && e.v != cs[f].v) lane_change_points(c, e, cs[f], 5.0, ctrl + 12 * b, tca + b);
      else { for (int i = 0; i < 12; ++i) ctrl[12 * b + i] = NAN; tca[b] = NAN; }
    }
  }
}

// threads <= 0: all cores.  Per-instance wall time (seconds) optionally written to time_out[B].
void orc_solve_batch(const MpcProblemDesc* d, const MpcSolverParams* sp, int64_t B, const double* x0,
                     const double* p, const double* road, const int32_t* road_idx, const double* obs,
                     const double* lb, const double* ub, double* U, double* lam, int32_t* status, int32_t* iters,
                     int32_t* evals, double* psi, double* res, double* time_out, int threads) {
  const int nx = nx_of(d->model_id), N = d->N;
  if (threads <= 0) threads = omp_get_max_threads();
  const bool bounded = is_bounded(d);
#pragma omp parallel num_threads(threads)
  {
    std::vector<double> tmp, lam_local(size_t(rows_of(*d)) * N);
#pragma omp for schedule(dynamic, 1)
    for (int64_t b = 0; b < B; ++b) {
      double t0 = omp_get_wtime();
      const double* r = road_of(d, road, road_idx, b, tmp);
      Problem P;
      P.I = make_instance<double>(d, p, r, obs ? obs + b * d->n_obs * MPCB_OBS_DIM : nullptr, nullptr, nullptr);
      P.x0 = x0 + b * nx;
      P.lb = lb;
      P.ub = ub;
      P.n = 2 * N;
      P.m = rows_of(*d) * N;
      P.bounded = bounded;
      double* lm = lam ? lam + b * rows_of(*d) * N : lam_local.data();
      if (!lam) std::fill(lam_local.begin(), lam_local.end(), 0.0);
      SolveResult R = alm_solve(P, *sp, U + b * 2 * N, lm);
      if (status) status[b] = R.status;
      if (iters) { iters[2 * b] = R.outer; iters[2 * b + 1] = R.inner; }
      if (evals) { evals[2 * b] = R.n_value; evals[2 * b + 1] = R.n_grad; }
      if (psi) psi[b] = R.psi;
      if (res) res[b] = R.eps;
      if (time_out) time_out[b] = omp_get_wtime() - t0;
    }
  }
}

// Algorithmic op tally of ONE psi + grad psi evaluation (counts[0] = arith+compare, [1] = transcendental)
// and of one psi-only evaluation (counts[2], counts[3]).
void orc_count_ops(const MpcProblemDesc* d, const double* x0, const double* U, const double* p, const double* road,
                   const double* obs, long long* counts) {
  const int nx = nx_of(d->model_id), N = d->N, S = d->S;
  std::vector<Cnt> cx0(nx), cU(2 * N), cp(MPCB_NP), croad(2 * S), cobs(size_t(d->n_obs) * MPCB_OBS_DIM), g(2 * N);
  for (int i = 0; i < nx; ++i) cx0[i] = x0[i];
  for (int i = 0; i < 2 * N; ++i) cU[i] = U[i];
  for (int i = 0; i < MPCB_NP; ++i) cp[i] = p[i];
  for (int i = 0; i < 2 * S; ++i) croad[i] = road[i];
  for (size_t i = 0; i < cobs.size(); ++i) cobs[i] = obs ? obs[i] : 0.0;
  Instance<Cnt> I = make_instance<Cnt>(d, cp.data(), croad.data(), obs ? cobs.data() : nullptr, nullptr, nullptr);
  Cnt::arith = Cnt::transc = 0;
  value_grad(I, cx0.data(), cU.data(), g.data());
  counts[0] = Cnt::arith;
  counts[1] = Cnt::transc;
  Cnt::arith = Cnt::transc = 0;
  rollout(I, cx0.data(), cU.data(), (Cnt*)nullptr, (int*)nullptr, (Cnt*)nullptr);
  counts[2] = Cnt::arith;
  counts[3] = Cnt::transc;
}

}  // extern "C"
