// orc_solver.hpp — CPU ORACLE (test infrastructure, NOT product code): structured PANOC with masked
// L-BFGS and the ALM outer loop, written straight from SURVEY.md Appendix A (the canonical solver
// spec: alpaqa 0.0.1 behaviour as configured by reference controller.py:27-48).
// "parity unpinned": alpaqa itself cannot be run here; this text is the spec's CPU statement.
#pragma once
#include <cfloat>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include "orc_cost.hpp"

namespace orc {

struct EvalCounter { int n_value = 0, n_grad = 0; };

struct Problem {
  Instance<double> I;
  const double* x0;
  const double *lb, *ub;  // [n] box C (main.py:55-56)
  int n;                  // 2N
  int m;                  // nx*N rows of g
  bool bounded;           // any finite D side and MPCB_FLAG_BOUNDED_D
  EvalCounter cnt;

  double psi(const double* x) {
    cnt.n_value++;
    return rollout(I, x0, x, (double*)nullptr, (int*)nullptr, (double*)nullptr);
  }
  double psi_grad(const double* x, double* g) {
    cnt.n_grad++;
    return value_grad(I, x0, x, g);
  }
  // err_z = g(x) - Pi_D(zeta(x)), yhat(x)   (Appendix A 3b); zero when D is unbounded.
  void err_z_yhat(const double* x, double* err_z, double* yhat) {
    const MpcProblemDesc& d = *I.d;
    const int nx = I.M.nx(), R = rows_of(d);
    if (!bounded) { for (int i = 0; i < m; ++i) { err_z[i] = 0; yhat[i] = 0; } return; }
    std::vector<double> st(size_t(d.N) * nx);
    rollout(I, x0, x, st.data(), (int*)nullptr, (double*)nullptr);
    for (int nn = 0; nn < d.N; ++nn) {
      for (int i = 0; i < nx; ++i) {
        double y = st[nn * nx + i];
        double g = y * y - d.g_offset[i];
        const bool on = (d.flags & MPCB_FLAG_BOUNDED_D) != 0;   // state rows are live only with MPCB_FLAG_BOUNDED_D
        alm_row(g, I.lam[nn * R + i], I.Sig[nn * R + i], on ? d.D_lower[i] : -INFINITY, on ? d.D_upper[i] : INFINITY,
                &yhat[nn * R + i], &err_z[nn * R + i]);
      }
      if (d.flags & MPCB_FLAG_SAFE_DIST) {
        const double t = double(nn + 1) * d.Ts;
        for (int o = 0; o < d.n_obs; ++o) {
          double xo, yo, po, vo;
          obstacle_at(I.obs + MPCB_OBS_DIM * o, t, &xo, &yo, &po, &vo);
          double g = safe_dist_row(d, &st[nn * nx], xo, yo, po, vo, 0.0, (double*)nullptr);
          alm_row(g, I.lam[nn * R + nx + o], I.Sig[nn * R + nx + o], -INFINITY, 0.0, &yhat[nn * R + nx + o], &err_z[nn * R + nx + o]);
        }
      }
    }
  }
};

inline double dot(int n, const double* a, const double* b) {
  double s = 0;
  for (int i = 0; i < n; ++i) s += a[i] * b[i];
  return s;
}

// Masked L-BFGS (Appendix A 3c/3e): circular buffer of (s, y) pairs.
struct LBFGS {
  int n, mem, idx = 0;
  bool full = false;
  std::vector<double> S, Y, alpha, rho;
  LBFGS(int n_, int mem_) : n(n_), mem(mem_), S(size_t(n_) * mem_), Y(size_t(n_) * mem_), alpha(mem_), rho(mem_) {}
  void reset() { idx = 0; full = false; }
  int count() const { return full ? mem : idx; }
  bool update(const double* x, const double* xn, const double* g, const double* gn) {
    // The pair is stored only if it is accepted (Appendix A 3e): with a full buffer slot idx holds the oldest live
    // pair, which a rejected pair must not destroy.
    double sy = 0, ss = 0;
    for (int i = 0; i < n; ++i) {
      const double si = xn[i] - x[i], yi = gn[i] - g[i];
      sy += si * yi;
      ss += si * si;
    }
    const double tiny = std::sqrt(DBL_MIN);
    if (!std::isfinite(sy) || !(ss >= tiny) || !(sy >= tiny)) return false;
    double* s = &S[size_t(idx) * n];
    double* y = &Y[size_t(idx) * n];
    for (int i = 0; i < n; ++i) {
      s[i] = xn[i] - x[i];
      y[i] = gn[i] - g[i];
    }
    idx++;
    if (idx == mem) { idx = 0; full = true; }
    return true;
  }
  // q <- H q restricted to the index set mask (1 = inactive/free).  gamma_fallback is used as H0
  // when no stored pair has positive curvature on the set.  Returns false when no pair was usable.
  bool apply(double* q, const unsigned char* mask, double gamma_fallback) {
    const int cnt = count();
    double H0 = -1;
    // newest -> oldest
    for (int k = 0; k < cnt; ++k) {
      int i = idx - 1 - k;
      if (i < 0) i += mem;
      const double* s = &S[size_t(i) * n];
      const double* y = &Y[size_t(i) * n];
      double sy = 0, yy = 0, sq = 0;
      for (int j = 0; j < n; ++j)
        if (mask[j]) { sy += s[j] * y[j]; yy += y[j] * y[j]; sq += s[j] * q[j]; }
      if (!(sy > 0)) { rho[i] = NAN; continue; }
      rho[i] = 1.0 / sy;
      if (H0 < 0) H0 = sy / yy;
      alpha[i] = rho[i] * sq;
      for (int j = 0; j < n; ++j)
        if (mask[j]) q[j] -= alpha[i] * y[j];
    }
    bool ok = H0 >= 0;
    if (!ok) H0 = gamma_fallback;
    for (int j = 0; j < n; ++j)
      if (mask[j]) q[j] *= H0;
    // oldest -> newest
    for (int k = cnt - 1; k >= 0; --k) {
      int i = idx - 1 - k;
      if (i < 0) i += mem;
      if (std::isnan(rho[i])) continue;
      const double* s = &S[size_t(i) * n];
      const double* y = &Y[size_t(i) * n];
      double yq = 0;
      for (int j = 0; j < n; ++j)
        if (mask[j]) yq += y[j] * q[j];
      double beta = rho[i] * yq;
      for (int j = 0; j < n; ++j)
        if (mask[j]) q[j] += (alpha[i] - beta) * s[j];
    }
    return ok;
  }
};

struct PanocResult {
  int status = MPCB_STATUS_UNKNOWN;
  int iterations = 0;
  double eps = 0;  // achieved ||p||_2
  double psi = 0;  // psi at the returned point
};

// x is overwritten with xhat_k on every exit except NotFinite (builder's decision).
inline PanocResult panoc(Problem& P, const MpcSolverParams& sp, double* x, double eps_tol) {
  const int n = P.n;
  std::vector<double> g(n), xh(n), gh(n), xhat(n), p(n), q(n), xn(n), gn(n), xhatn(n), pn(n), work(n), gw(n);
  std::vector<unsigned char> mask(n);
  LBFGS lb(n, sp.lbfgs_mem);
  PanocResult R;

  auto prox_step = [&](const double* xx, const double* gg, double gamma, double* xhat_, double* p_) {
    for (int i = 0; i < n; ++i) {
      double v = xx[i] - gamma * gg[i];
      v = v < P.lb[i] ? P.lb[i] : (v > P.ub[i] ? P.ub[i] : v);
      xhat_[i] = v;
      p_[i] = v - xx[i];
    }
  };
  // descent lemma (Appendix A 3a); returns true if gamma changed
  auto lemma = [&](const double* xx, const double* gg, double psi_x, double* xhat_, double* p_, double& psi_hat,
                   double& gTp, double& pTp, double& L, double& gamma) {
    bool changed = false;
    double margin = (1 + std::fabs(psi_x)) * sp.quad_tol;
    while (psi_hat - psi_x > gTp + 0.5 * L * pTp + margin) {
      if (!(L * 2 <= sp.L_max)) break;
      L *= 2;
      gamma /= 2;
      changed = true;
      prox_step(xx, gg, gamma, xhat_, p_);
      psi_hat = P.psi(xhat_);
      gTp = dot(n, gg, p_);
      pTp = dot(n, p_, p_);
    }
    return changed;
  };

  double psi = P.psi_grad(x, g.data());
  // Lipschitz estimate (Appendix A 1)
  double hh = 0;
  for (int i = 0; i < n; ++i) {
    double h = std::fmax(std::fabs(x[i]) * sp.lip_eps, sp.lip_delta);
    xh[i] = x[i] + h;
    hh += h * h;
  }
  P.psi_grad(xh.data(), gh.data());
  double dd = 0;
  for (int i = 0; i < n; ++i) dd += (gh[i] - g[i]) * (gh[i] - g[i]);
  double L = std::sqrt(dd) / std::sqrt(hh);
  if (!std::isfinite(L)) { R.status = MPCB_STATUS_NOT_FINITE; R.eps = NAN; R.psi = psi; return R; }
  L = L < sp.L_min ? sp.L_min : (L > sp.L_max ? sp.L_max : L);
  double gamma = sp.Lgamma_factor / L;

  prox_step(x, g.data(), gamma, xhat.data(), p.data());
  double psi_hat = P.psi(xhat.data());
  double gTp = dot(n, g.data(), p.data()), pTp = dot(n, p.data(), p.data());
  double phi = 0;
  int no_progress = 0;

  for (int k = 0;; ++k) {
    if (k == 0) {
      lemma(x, g.data(), psi, xhat.data(), p.data(), psi_hat, gTp, pTp, L, gamma);
      phi = psi + pTp / (2 * gamma) + gTp;
    }
    // stop test (ProjGradNorm2, controller.py:29)
    double eps_k = std::sqrt(pTp);
    int stop = 0;
    if (!std::isfinite(eps_k)) stop = MPCB_STATUS_NOT_FINITE;
    else if (eps_k <= eps_tol) stop = MPCB_STATUS_CONVERGED;
    else if (k == sp.max_iter) stop = MPCB_STATUS_MAX_ITER;
    else if (no_progress > sp.max_no_progress) stop = MPCB_STATUS_NO_PROGRESS;
    if (stop) {
      R.status = stop;
      R.iterations = k;
      R.eps = eps_k;
      R.psi = psi_hat;
      if (stop != MPCB_STATUS_NOT_FINITE) std::memcpy(x, xhat.data(), sizeof(double) * n);
      return R;
    }
    // direction (Appendix A 3c)
    double tau = 0;
    if (k > 0) {
      int nJ = 0;
      for (int i = 0; i < n; ++i) {
        double gd = x[i] - gamma * g[i];
        bool active = gd < P.lb[i] || P.ub[i] < gd;
        mask[i] = !active;
        nJ += !active;
        q[i] = active ? p[i] : 0.0;
      }
      if (nJ > 0) {
        if (nJ < n && sp.hess_vec_fd) {
          double xnorm = std::sqrt(dot(n, x, x));
          double h = std::cbrt(DBL_EPSILON) * (1 + xnorm);
          for (int i = 0; i < n; ++i) work[i] = x[i] + h * q[i];
          P.psi_grad(work.data(), gw.data());
          for (int i = 0; i < n; ++i)
            if (mask[i]) q[i] = -g[i] - (gw[i] - g[i]) / h;
        } else {
          for (int i = 0; i < n; ++i)
            if (mask[i]) q[i] = -g[i];
        }
        lb.apply(q.data(), mask.data(), gamma);
      }
      tau = 1;
      bool fin = true;
      for (int i = 0; i < n; ++i) fin = fin && std::isfinite(q[i]);
      if (!fin) { tau = 0; lb.reset(); }
    }
    // line search (Appendix A 3d)
    double sigma = (1 - gamma * L) * pTp / (2 * gamma);
    double psi_n, psi_hat_n, gTp_n, pTp_n, L_n, gamma_n, phi_n, ls_cond;
    do {
      L_n = L;
      gamma_n = gamma;
      if (tau / 2 < sp.tau_min) std::memcpy(xn.data(), xhat.data(), sizeof(double) * n);
      else
        for (int i = 0; i < n; ++i) xn[i] = x[i] + (1 - tau) * p[i] + tau * q[i];
      psi_n = P.psi_grad(xn.data(), gn.data());
      prox_step(xn.data(), gn.data(), gamma_n, xhatn.data(), pn.data());
      psi_hat_n = P.psi(xhatn.data());
      gTp_n = dot(n, gn.data(), pn.data());
      pTp_n = dot(n, pn.data(), pn.data());
      lemma(xn.data(), gn.data(), psi_n, xhatn.data(), pn.data(), psi_hat_n, gTp_n, pTp_n, L_n, gamma_n);
      phi_n = psi_n + pTp_n / (2 * gamma_n) + gTp_n;
      ls_cond = phi_n - (phi - sigma);
      static const bool dbg = getenv("ORC_DEBUG") != nullptr;
      if (dbg)
        fprintf(stderr, "k=%d tau=%g psi+=%g psihat+=%g phi+=%g phi=%g sigma=%g L+=%g gamma+=%g pTp+=%g\n", k, tau,
                psi_n, psi_hat_n, phi_n, phi, sigma, L_n, gamma_n, pTp_n);
      tau /= 2;
      // NaN-safe: a non-finite phi+ (the L-BFGS step left the model's domain) is a FAILED trial, so
      // back-tracking continues down to the always-finite proximal-gradient step.  (Appendix A's
      // "accept if phi+ <= ..." read literally; `ls_cond > margin` would accept NaN.)
    } while (!(ls_cond <= (1 + std::fabs(phi)) * sp.ls_tol) && tau >= sp.tau_min);
    // L-BFGS update (Appendix A 3e) and no-progress counter
    lb.update(x, xn.data(), g.data(), gn.data());
    if (no_progress > 0 || k % (sp.max_no_progress > 0 ? sp.max_no_progress : 1) == 0) {
      bool same = true;
      for (int i = 0; i < n; ++i) same = same && (x[i] == xn[i]);
      no_progress = same ? no_progress + 1 : 0;
    }
    // advance
    std::memcpy(x, xn.data(), sizeof(double) * n);
    g.swap(gn);
    xhat.swap(xhatn);
    p.swap(pn);
    psi = psi_n;
    psi_hat = psi_hat_n;
    gTp = gTp_n;
    pTp = pTp_n;
    L = L_n;
    gamma = gamma_n;
    phi = phi_n;
  }
}

struct SolveResult {
  int status, outer, inner, n_value, n_grad, inner_failures;
  double psi, eps;
};

// ALM outer loop (Appendix A; controller.py:39-48).  lam [m] in/out, U [n] in/out.
inline SolveResult alm_solve(Problem& P, const MpcSolverParams& sp, double* U, double* lam) {
  const int n = P.n, m = P.m;
  const MpcProblemDesc& d = *P.I.d;
  const int nx = P.I.M.nx();
  std::vector<double> Sig(m, sp.Sigma0), e1(m, 0.0), e2(m, 0.0), yhat(m), x_bak(n), lam_in(m);
  SolveResult R{};
  double eps = sp.eps0, Delta = sp.Delta;
  double norm_e1 = 0;
  std::memcpy(lam_in.data(), lam, sizeof(double) * m);
  for (int i = 0;; ++i) {
    if (P.bounded)
      for (int r = 0; r < m; ++r) {  // clamp y to +-M respecting infinite D sides
        const int R = rows_of(d), i = r % R;
        const bool state_on = i < nx && (d.flags & MPCB_FLAG_BOUNDED_D);
        double lo = i < nx ? ((!state_on || std::isinf(d.D_lower[i])) ? 0.0 : -sp.M) : 0.0;   // ellipse rows: D = (-inf, 0]
        double hi = i < nx ? ((!state_on || std::isinf(d.D_upper[i])) ? 0.0 : sp.M) : sp.M;
        lam[r] = lam[r] < lo ? lo : (lam[r] > hi ? hi : lam[r]);
      }
    P.I.lam = P.bounded ? lam : nullptr;
    P.I.Sig = P.bounded ? Sig.data() : nullptr;
    std::memcpy(x_bak.data(), U, sizeof(double) * n);
    PanocResult pr = panoc(P, sp, U, eps);
    R.inner += pr.iterations;
    R.outer = i + 1;
    R.psi = pr.psi;
    R.eps = pr.eps;
    if (pr.status == MPCB_STATUS_NOT_FINITE) {
      std::memcpy(U, x_bak.data(), sizeof(double) * n);  // failed instances keep their warm start
      R.status = MPCB_STATUS_NOT_FINITE;
      break;
    }
    bool inner_conv = pr.status == MPCB_STATUS_CONVERGED;
    if (!inner_conv) R.inner_failures++;
    P.err_z_yhat(U, e2.data(), yhat.data());
    if (P.bounded) std::memcpy(lam, yhat.data(), sizeof(double) * m);
    double norm_e2 = 0;
    for (int r = 0; r < m; ++r) norm_e2 = std::fmax(norm_e2, std::fabs(e2[r]));
    bool conv = inner_conv && pr.eps <= sp.eps_target && norm_e2 <= sp.delta_target;
    if (conv) { R.status = MPCB_STATUS_CONVERGED; break; }
    if (i + 1 == sp.alm_max_iter) { R.status = MPCB_STATUS_MAX_ITER; break; }
    if (!inner_conv && eps <= sp.eps_target) { R.status = pr.status; break; }  // final tolerance failed
    // penalty update
    if (norm_e2 > sp.delta_target)
      for (int r = 0; r < m; ++r)
        if (i == 0 || std::fabs(e2[r]) > sp.theta * std::fabs(e1[r]))
          Sig[r] = std::fmin(sp.Sigma_max, std::fmax(Delta * std::fabs(e2[r]) / norm_e2, 1.0) * Sig[r]);
    eps = std::fmax(sp.rho * eps, sp.eps_target);
    e1.swap(e2);
    norm_e1 = norm_e2;
  }
  (void)norm_e1;
  R.n_value = P.cnt.n_value;
  R.n_grad = P.cnt.n_grad;
  return R;
}

}  // namespace orc
