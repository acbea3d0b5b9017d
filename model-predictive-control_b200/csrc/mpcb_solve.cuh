// mpcb_solve.cuh — the fused, persistent ALM + structured-PANOC + masked-L-BFGS kernel.
//
// Replaces `self.solver(problem, U, λ)` (reference controller.py:57): alpaqa's ALMSolver over
// StructuredPANOCLBFGSSolver calling CasADi-generated f/grad_f.  Canonical algorithm: SURVEY.md
// Appendix A (restated on the CPU in oracle/orc_solver.hpp).
//
// Mapping ("instance-per-lane"): every lane owns one MPC instance and walks a small state machine.
// All lanes of a warp meet at ONE evaluation site per loop trip (forward rollout, plus the reverse
// sweep for the lanes that asked for a gradient), so the dominant fp64 work never diverges even
// though each lane is at a different PANOC iteration / line-search trial.  Finished lanes pull the
// next instance from a global queue (iteration counts vary 10x between instances).  Per-lane
// vectors (iterates, L-BFGS history, checkpoints) live in a lane-interleaved workspace so a warp's
// accesses to element i of 32 instances are one 256-byte coalesced transaction.
#pragma once
#include <float.h>
#include "mpcb_fwd.cuh"

namespace mpcb {

struct SolveArgs {
  int64_t B;
  const double *x0, *p, *obs, *lb, *ub;
  RoadArg ra;
  double *U, *lam;
  int32_t *status, *iters, *evals;
  double *psi, *res;
  double* ws;
  int32_t* wsi;
  int64_t ns;
  unsigned long long* queue;
};

enum SolveState {
  ST_IDLE = 0, ST_START_PANOC, ST_INIT_G, ST_LIP_G, ST_XHAT_V, ST_ITER_TOP, ST_FD_G, ST_LS_START, ST_LS_TRIAL,
  ST_LS_G, ST_LS_V, ST_PANOC_EXIT, ST_ALM_V, ST_ALM_UPDATE
};

// Workspace row map (units: doubles per lane).
struct WsMap {
  int n, mem;
  int Q, XE, GE, XB, LS, LY, AL, RH, CHK, ST, SIG, E1, E2, total;
  __host__ __device__ WsMap(int n_, int mem_, int N, int n_sub, int nx, int m) : n(n_), mem(mem_) {
    Q = 8 * n; XE = 9 * n; GE = 10 * n; XB = 11 * n;
    LS = 12 * n; LY = LS + mem * n; AL = LY + mem * n; RH = AL + mem;
    CHK = RH + mem; ST = CHK + N * (nx * (nx - 1) + 2 + nx); SIG = ST + N * nx; E1 = SIG + m; E2 = E1 + m; total = E2 + m;
  }
  __host__ __device__ int X(int bank) const { return bank * 4 * n; }
  __host__ __device__ int G(int bank) const { return bank * 4 * n + n; }
  __host__ __device__ int XH(int bank) const { return bank * 4 * n + 2 * n; }
  __host__ __device__ int P(int bank) const { return bank * 4 * n + 3 * n; }
};

template <int MODEL>
__global__ void __launch_bounds__(128)
k_solve(KDesc d, MpcSolverParams sp, SolveArgs a) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  extern __shared__ double smem[];
  __shared__ VehP v;
  const int n = d.n, m = d.m, mem = sp.lbfgs_mem;
  double* sroad = smem;
  double* sbox = smem + (a.ra.in_smem ? a.ra.R * d.S * 2 : 0);  // lb[n], ub[n]
  if (threadIdx.x == 0) load_vehicle(&v, a.p, d);
  if (a.ra.in_smem) stage_roads(sroad, a.ra.road, a.ra.R, d.S);
  for (int i = threadIdx.x; i < n; i += blockDim.x) { sbox[i] = a.lb[i]; sbox[n + i] = a.ub[i]; }
  __syncthreads();
  const double* lbv = sbox;
  const double* ubv = sbox + n;

  const int64_t ns = a.ns;
  const int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const WsMap M(n, mem, d.N, d.n_sub, NX, m);
  auto W = [&](int row) { return Vec{a.ws + (int64_t)row * ns + slot, ns}; };
  const Vec scr = W(M.CHK);  // stage Jacobians of the forward-mode evaluation

  // ---- per-lane registers ----
  int state = ST_IDLE, want = 0, ep = 0, eg = 0;
  bool exhausted = false;
  int64_t b = -1;
  int bank = 0, k = 0, no_prog = 0, lb_idx = 0, lb_full = 0;
  double psi = 0, psi_hat = 0, gTp = 0, pTp = 0, L = 0, gam = 0, phi = 0, tau = 0, sigma = 0, hfd = 0, hh = 0;
  double psi_n = 0, gTp_n = 0, pTp_n = 0, L_n = 0, gam_n = 0;
  double eps_tol = 0, peps = 0, ppsi = 0, norm_e2 = 0, val = 0;
  int alm_i = 0, inner_total = 0, n_val = 0, n_grad = 0, pstatus = 0;
  unsigned long long mask0 = 0, mask1 = 0;  // inactive-set bitmask (n <= 128)
  auto mbit = [&](int j) { return j < 64 ? (mask0 >> j) & 1ull : (mask1 >> (j - 64)) & 1ull; };

  // x̂ = Π_C(x − γ g), p = x̂ − x; also gᵀp and pᵀp (sequential sums, same order as the oracle)
  auto prox = [&](int xr, int gr, double gamma, int xhr, int pr, double& gtp, double& ptp) {
    Vec X = W(xr), G = W(gr), XH = W(xhr), P = W(pr);
    double s1 = 0.0, s2 = 0.0;
    for (int i = 0; i < n; ++i) {
      double xi = X(i), gi = G(i);
      double t = xi - gamma * gi;
      t = t < lbv[i] ? lbv[i] : (t > ubv[i] ? ubv[i] : t);
      double pi = t - xi;
      XH.set(i, t);
      P.set(i, pi);
      s1 += gi * pi;
      s2 += pi * pi;
    }
    gtp = s1;
    ptp = s2;
  };

  // q <- H q on the inactive set (masked two-loop recursion, Appendix A 3c)
  auto lbfgs_apply = [&]() {
    Vec Q = W(M.Q), AL = W(M.AL), RH = W(M.RH);
    const int cnt = lb_full ? mem : lb_idx;
    double H0 = -1.0;
    for (int kk = 0; kk < cnt; ++kk) {
      int i = lb_idx - 1 - kk;
      if (i < 0) i += mem;
      Vec S = W(M.LS + i * n), Y = W(M.LY + i * n);
      double sy = 0.0, yy = 0.0, sq = 0.0;
      for (int j = 0; j < n; ++j)
        if (mbit(j)) { double s = S(j), y = Y(j); sy += s * y; yy += y * y; sq += s * Q(j); }
      if (!(sy > 0.0)) { RH.set(i, nan("")); continue; }
      double rho = 1.0 / sy;
      RH.set(i, rho);
      if (H0 < 0.0) H0 = sy / yy;
      double al = rho * sq;
      AL.set(i, al);
      for (int j = 0; j < n; ++j)
        if (mbit(j)) Q.set(j, Q(j) - al * Y(j));
    }
    if (!(H0 >= 0.0)) H0 = gam;
    for (int j = 0; j < n; ++j)
      if (mbit(j)) Q.set(j, Q(j) * H0);
    for (int kk = cnt - 1; kk >= 0; --kk) {
      int i = lb_idx - 1 - kk;
      if (i < 0) i += mem;
      double rho = RH(i);
      if (isnan(rho)) continue;
      Vec S = W(M.LS + i * n), Y = W(M.LY + i * n);
      double yq = 0.0;
      for (int j = 0; j < n; ++j)
        if (mbit(j)) yq += Y(j) * Q(j);
      double coef = AL(i) - rho * yq;
      for (int j = 0; j < n; ++j)
        if (mbit(j)) Q.set(j, Q(j) + coef * S(j));
    }
  };

  for (;;) {
    // ================= transition phase: run the lane's state machine up to its next evaluation ==
    want = 0;
    for (;;) {
      if (state == ST_IDLE) {
        if (exhausted) break;
        b = (int64_t)atomicAdd(a.queue, 1ull);
        if (b >= a.B) { exhausted = true; break; }
        // controller.py:20-21,57: warm start (U, λ) comes in through U_inout / lambda_inout
        Vec X = W(M.X(0));
        for (int i = 0; i < n; ++i) X.set(i, a.U[b * n + i]);
        bank = 0;
        eps_tol = sp.eps0;
        alm_i = 0; inner_total = 0; n_val = 0; n_grad = 0;
        if (d.bounded) {
          Vec SG = W(M.SIG), E1 = W(M.E1);
          for (int r = 0; r < m; ++r) { SG.set(r, sp.Sigma0); E1.set(r, 0.0); }
        }
        state = ST_START_PANOC;
      } else if (state == ST_START_PANOC) {
        if (d.bounded) {  // clamp multipliers to ±M, 0 on infinite sides
          for (int r = 0; r < m; ++r) {
            double lo = isinf(d.D_lo[r % NX]) ? 0.0 : -sp.M, hi = isinf(d.D_hi[r % NX]) ? 0.0 : sp.M;
            double lv = a.lam[b * m + r];
            a.lam[b * m + r] = lv < lo ? lo : (lv > hi ? hi : lv);
          }
        }
        Vec X = W(M.X(bank)), XB = W(M.XB);
        for (int i = 0; i < n; ++i) XB.set(i, X(i));
        k = 0; no_prog = 0; lb_idx = 0; lb_full = 0;
        ep = M.X(bank); eg = M.G(bank); want = 2;
        state = ST_INIT_G;
        break;
      } else if (state == ST_INIT_G) {
        psi = val;
        Vec X = W(M.X(bank)), XE = W(M.XE);
        hh = 0.0;
        for (int i = 0; i < n; ++i) {
          double xi = X(i);
          double h = fmax(fabs(xi) * sp.lip_eps, sp.lip_delta);
          XE.set(i, xi + h);
          hh += h * h;
        }
        ep = M.XE; eg = M.GE; want = 2;
        state = ST_LIP_G;
        break;
      } else if (state == ST_LIP_G) {
        Vec G = W(M.G(bank)), GE = W(M.GE);
        double dd = 0.0;
        for (int i = 0; i < n; ++i) { double t = GE(i) - G(i); dd += t * t; }
        L = sqrt(dd) / sqrt(hh);
        if (!isfinite(L)) { pstatus = MPCB_STATUS_NOT_FINITE; peps = nan(""); ppsi = psi; state = ST_PANOC_EXIT; continue; }
        L = L < sp.L_min ? sp.L_min : (L > sp.L_max ? sp.L_max : L);
        gam = sp.Lgamma_factor / L;
        prox(M.X(bank), M.G(bank), gam, M.XH(bank), M.P(bank), gTp, pTp);
        ep = M.XH(bank); want = 1;
        state = ST_XHAT_V;
        break;
      } else if (state == ST_XHAT_V) {
        psi_hat = val;
        double margin = (1.0 + fabs(psi)) * sp.quad_tol;
        if (psi_hat - psi > gTp + 0.5 * L * pTp + margin && L * 2.0 <= sp.L_max) {  // descent lemma, k = 0
          L *= 2.0; gam *= 0.5;
          prox(M.X(bank), M.G(bank), gam, M.XH(bank), M.P(bank), gTp, pTp);
          ep = M.XH(bank); want = 1;
          break;
        }
        phi = psi + pTp / (2.0 * gam) + gTp;
        state = ST_ITER_TOP;
      } else if (state == ST_ITER_TOP) {
        double eps_k = sqrt(pTp);
        int stop = 0;
        if (!isfinite(eps_k)) stop = MPCB_STATUS_NOT_FINITE;
        else if (eps_k <= eps_tol) stop = MPCB_STATUS_CONVERGED;
        else if (k == sp.max_iter) stop = MPCB_STATUS_MAX_ITER;
        else if (no_prog > sp.max_no_progress) stop = MPCB_STATUS_NO_PROGRESS;
        if (stop) { pstatus = stop; peps = eps_k; ppsi = psi_hat; state = ST_PANOC_EXIT; continue; }
        if (k == 0) { state = ST_LS_START; continue; }
        // direction (Appendix A 3c)
        Vec X = W(M.X(bank)), G = W(M.G(bank)), P = W(M.P(bank)), Q = W(M.Q);
        int nJ = 0;
        mask0 = 0; mask1 = 0;
        for (int i = 0; i < n; ++i) {
          double gd = X(i) - gam * G(i);
          bool active = gd < lbv[i] || ubv[i] < gd;
          if (!active) { ++nJ; if (i < 64) mask0 |= 1ull << i; else mask1 |= 1ull << (i - 64); }
        }
        bool use_fd = sp.hess_vec_fd && nJ > 0 && nJ < n;
        double xx = 0.0;
        for (int i = 0; i < n; ++i) {
          double xi = X(i);
          xx += xi * xi;
          Q.set(i, mbit(i) ? (use_fd ? 0.0 : -G(i)) : P(i));
        }
        if (use_fd) {
          hfd = 6.0554544523933395e-06 /* cbrt(DBL_EPSILON) */ * (1.0 + sqrt(xx));
          Vec XE = W(M.XE);
          for (int i = 0; i < n; ++i) XE.set(i, X(i) + hfd * Q(i));
          ep = M.XE; eg = M.GE; want = 2;
          state = ST_FD_G;
          break;
        }
        if (nJ > 0) lbfgs_apply();
        state = ST_LS_START;
      } else if (state == ST_FD_G) {
        Vec G = W(M.G(bank)), GE = W(M.GE), Q = W(M.Q);
        for (int i = 0; i < n; ++i)
          if (mbit(i)) { double gi = G(i); Q.set(i, -gi - (GE(i) - gi) / hfd); }
        lbfgs_apply();
        state = ST_LS_START;
      } else if (state == ST_LS_START) {
        tau = 0.0;
        if (k > 0) {
          tau = 1.0;
          Vec Q = W(M.Q);
          bool fin = true;
          for (int i = 0; i < n; ++i) fin = fin && isfinite(Q(i));
          if (!fin) { tau = 0.0; lb_idx = 0; lb_full = 0; }
        }
        sigma = (1.0 - gam * L) * pTp / (2.0 * gam);
        state = ST_LS_TRIAL;
      } else if (state == ST_LS_TRIAL) {
        L_n = L; gam_n = gam;
        const int nb = bank ^ 1;
        Vec X = W(M.X(bank)), XN = W(M.X(nb));
        if (tau * 0.5 < sp.tau_min) {
          Vec XH = W(M.XH(bank));
          for (int i = 0; i < n; ++i) XN.set(i, XH(i));
        } else {
          Vec P = W(M.P(bank)), Q = W(M.Q);
          for (int i = 0; i < n; ++i) XN.set(i, X(i) + (1.0 - tau) * P(i) + tau * Q(i));
        }
        ep = M.X(nb); eg = M.G(nb); want = 2;
        state = ST_LS_G;
        break;
      } else if (state == ST_LS_G) {
        psi_n = val;
        const int nb = bank ^ 1;
        prox(M.X(nb), M.G(nb), gam_n, M.XH(nb), M.P(nb), gTp_n, pTp_n);
        ep = M.XH(nb); want = 1;
        state = ST_LS_V;
        break;
      } else if (state == ST_LS_V) {
        const int nb = bank ^ 1;
        double psi_hat_n = val;
        double margin = (1.0 + fabs(psi_n)) * sp.quad_tol;
        if (psi_hat_n - psi_n > gTp_n + 0.5 * L_n * pTp_n + margin && L_n * 2.0 <= sp.L_max) {
          L_n *= 2.0; gam_n *= 0.5;
          prox(M.X(nb), M.G(nb), gam_n, M.XH(nb), M.P(nb), gTp_n, pTp_n);
          ep = M.XH(nb); want = 1;
          break;
        }
        double phi_n = psi_n + pTp_n / (2.0 * gam_n) + gTp_n;
        double ls_cond = phi_n - (phi - sigma);
        tau *= 0.5;
        // NaN-safe: a non-finite phi+ is a failed trial (keeps back-tracking to the prox step)
        if (!(ls_cond <= (1.0 + fabs(phi)) * sp.ls_tol) && tau >= sp.tau_min) { state = ST_LS_TRIAL; continue; }
        // accept: L-BFGS update with s = x⁺ − x, y = ∇ψ⁺ − ∇ψ (Appendix A 3e)
        {
          Vec X = W(M.X(bank)), XN = W(M.X(nb)), G = W(M.G(bank)), GN = W(M.G(nb));
          Vec S = W(M.LS + lb_idx * n), Y = W(M.LY + lb_idx * n);
          double sy = 0.0, ss = 0.0;
          bool same = true;
          for (int i = 0; i < n; ++i) {
            double xi = X(i), xni = XN(i);
            double s = xni - xi, y = GN(i) - G(i);
            S.set(i, s);
            Y.set(i, y);
            sy += s * y;
            ss += s * s;
            same = same && (xi == xni);
          }
          const double tiny = 1.4916681462400413e-154;  // sqrt(DBL_MIN)
          if (isfinite(sy) && ss >= tiny && sy >= tiny) {
            if (++lb_idx == mem) { lb_idx = 0; lb_full = 1; }
          }
          if (no_prog > 0 || k % sp.max_no_progress == 0) no_prog = same ? no_prog + 1 : 0;
        }
        bank = nb;
        psi = psi_n; psi_hat = psi_hat_n; gTp = gTp_n; pTp = pTp_n; L = L_n; gam = gam_n; phi = phi_n;
        ++k;
        state = ST_ITER_TOP;
      } else if (state == ST_PANOC_EXIT) {
        inner_total += k;
        Vec X = W(M.X(bank));
        if (pstatus == MPCB_STATUS_NOT_FINITE) {  // failed instances keep their warm start
          Vec XB = W(M.XB);
          for (int i = 0; i < n; ++i) a.U[b * n + i] = XB(i);
          a.status[b] = MPCB_STATUS_NOT_FINITE;
          a.iters[2 * b] = alm_i + 1; a.iters[2 * b + 1] = inner_total;
          a.evals[2 * b] = n_val; a.evals[2 * b + 1] = n_grad;
          a.psi[b] = ppsi; a.res[b] = peps;
          state = ST_IDLE;
          continue;
        }
        Vec XH = W(M.XH(bank));
        for (int i = 0; i < n; ++i) X.set(i, XH(i));  // return x̂_k
        norm_e2 = 0.0;
        if (d.bounded) { ep = M.X(bank); want = 3; state = ST_ALM_V; break; }
        state = ST_ALM_UPDATE;
      } else if (state == ST_ALM_V) {
        // err_z = g(x̂) − Π_D(ζ), λ ← ŷ(x̂)
        Vec STt = W(M.ST), SG = W(M.SIG), E2 = W(M.E2);
        norm_e2 = 0.0;
        for (int r = 0; r < m; ++r) {
          int i = r % NX;
          double y = STt(r);
          double yh, ez;
          alm_row(y * y - d.g_off[i], a.lam[b * m + r], SG(r), d.D_lo[i], d.D_hi[i], &yh, &ez);
          E2.set(r, ez);
          norm_e2 = fmax(norm_e2, fabs(ez));
        }
        for (int r = 0; r < m; ++r) {
          int i = r % NX;
          double y = STt(r);
          double yh, ez;
          alm_row(y * y - d.g_off[i], a.lam[b * m + r], SG(r), d.D_lo[i], d.D_hi[i], &yh, &ez);
          a.lam[b * m + r] = yh;
        }
        state = ST_ALM_UPDATE;
      } else /* ST_ALM_UPDATE */ {
        bool inner_conv = pstatus == MPCB_STATUS_CONVERGED;
        int fin = 0;
        if (inner_conv && peps <= sp.eps_target && norm_e2 <= sp.delta_target) fin = MPCB_STATUS_CONVERGED;
        else if (alm_i + 1 == sp.alm_max_iter) fin = MPCB_STATUS_MAX_ITER;
        else if (!inner_conv && eps_tol <= sp.eps_target) fin = pstatus;
        if (fin) {
          Vec X = W(M.X(bank));
          for (int i = 0; i < n; ++i) a.U[b * n + i] = X(i);
          a.status[b] = fin;
          a.iters[2 * b] = alm_i + 1; a.iters[2 * b + 1] = inner_total;
          a.evals[2 * b] = n_val; a.evals[2 * b + 1] = n_grad;
          a.psi[b] = ppsi; a.res[b] = peps;
          state = ST_IDLE;
          continue;
        }
        if (d.bounded) {
          Vec SG = W(M.SIG), E1 = W(M.E1), E2 = W(M.E2);
          for (int r = 0; r < m; ++r) {
            double e2 = E2(r);
            if (norm_e2 > sp.delta_target && (alm_i == 0 || fabs(e2) > sp.theta * fabs(E1(r))))
              SG.set(r, fmin(sp.Sigma_max, fmax(sp.Delta * fabs(e2) / norm_e2, 1.0) * SG(r)));
            E1.set(r, e2);
          }
        }
        eps_tol = fmax(sp.rho * eps_tol, sp.eps_target);
        ++alm_i;
        state = ST_START_PANOC;
      }
    }
    // ================= evaluation phase: the one convergent site ================================
    if (!__any_sync(0xffffffffu, want != 0)) break;
    if (want) {
      RoadView r = road_view(a.ra, sroad, d.S, b);
      Extras ex{a.obs ? a.obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, d.bounded ? a.lam + b * m : nullptr,
                d.bounded ? a.ws + (int64_t)M.SIG * ns + slot : nullptr, 1, ns};
      double y0[NX];
#pragma unroll
      for (int i = 0; i < NX; ++i) y0[i] = a.x0[b * NX + i];
      Vec Uv = W(ep);
      const Vec stv = want == 3 ? W(M.ST) : Vec{nullptr, 0};
      if (__any_sync(__activemask(), want == 2)) val = eval_fm<MODEL, true>(d, v, r, ex, y0, Uv, scr, W(eg), want == 2, stv);
      else val = eval_fm<MODEL, false>(d, v, r, ex, y0, Uv, scr, W(eg), false, stv);
      if (want == 2) ++n_grad;
      else if (want == 1) ++n_val;
    }
  }
}

}  // namespace mpcb
