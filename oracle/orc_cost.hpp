// orc_cost.hpp — CPU ORACLE (test infrastructure, NOT product code): road errors, potentials,
// Bezier road, stage cost, horizon rollout and reverse sweep.  See orc_model.hpp for the status.
#pragma once
#include <limits>
#include "orc_model.hpp"

namespace orc {

// car_dynamics.py:174-192: distances WITH sqrt, scan i = 1..S-2, strict '<', initial guess vertex 0
// with previous = vertex 0 and next = vertex 1.  Returns the winning index; prev/next indices out.
template <class T>
inline int nearest_casadi(int S, const T* road, const T& px, const T& py, int* iprev, int* inext) {
  auto dist = [&](int i) {
    T dx = road[i] - px, dy = road[S + i] - py;
    return sqrt(dx * dx + dy * dy);
  };
  T min_val = dist(0);
  int best = 0;
  *iprev = 0;
  *inext = 1;
  for (int i = 1; i < S - 1; ++i) {
    T di = dist(i);
    if (di < min_val) {
      min_val = di;
      best = i;
      *iprev = i - 1;
      *inext = i + 1;
    }
  }
  return best;
}

// road.py:36-50: plain argmin over every vertex (np.linalg.norm + np.argmin: first minimum).
template <class T>
inline int nearest_numpy(int S, const T* road, const T& px, const T& py) {
  int best = 0;
  T min_val = T(0);
  for (int i = 0; i < S; ++i) {
    T dx = road[i] - px, dy = road[S + i] - py;
    T di = sqrt(dx * dx + dy * dy);
    if (i == 0 || di < min_val) { min_val = di; best = i; }
  }
  return best;
}

// car_dynamics.py:194-228 (normalised = false) and road.py:52-81 (normalised = true).
// err = [cte, heading_error, pos_error].
template <class T>
inline int road_errors(int S, const T* road, const T& px, const T& py, const T& heading, bool normalised,
                       T* err, int* iprev_out = nullptr, int* inext_out = nullptr) {
  int ip, in, idx;
  if (normalised) {
    idx = nearest_numpy(S, road, px, py);
    ip = idx == 0 ? S - 1 : idx - 1;        // python negative index (road.py:66)
    in = idx + 1 < S ? idx + 1 : S - 1;     // road.py:76 (the heading line would raise at idx = S-1)
  } else {
    idx = nearest_casadi(S, road, px, py, &ip, &in);
  }
  T nx_ = road[idx], ny_ = road[S + idx];
  T vx = px - road[ip], vy = py - road[S + ip];
  T wx = nx_ - road[ip], wy = ny_ - road[S + ip];
  T cte = vx * wy - vy * wx;
  T w2x = road[in] - nx_, w2y = road[S + in] - ny_;
  T desired = atan2(w2y, w2x);
  T he = wrap_to_pi(desired - heading);
  T v2x = px - nx_, v2y = py - ny_;
  T pe = v2x * w2y - v2y * w2x;
  if (normalised) {
    cte = cte / sqrt(wx * wx + wy * wy);
    pe = pe / sqrt(w2x * w2x + w2y * w2y);
  }
  err[0] = cte;
  err[1] = he;
  err[2] = pe;
  if (iprev_out) *iprev_out = ip;
  if (inext_out) *inext_out = in;
  return idx;
}

// bezier_curves.py:15-25: quintic Bernstein evaluation with math.factorial binomials.
template <class T>
inline void bezier_point(const T* P /*[12]: Px0..5, Py0..5*/, const T& j, T* x, T* y) {
  static const double C5[6] = {1, 5, 10, 10, 5, 1};
  T sx = T(0), sy = T(0);
  for (int i = 0; i < 6; ++i) {
    T a = pow(T(1) - j, T(double(5 - i)));
    T b = pow(j, T(double(i)));
    T w = T(C5[i]) * a * b;
    sx = sx + w * P[i];
    sy = sy + w * P[6 + i];
  }
  *x = sx;
  *y = sy;
}

template <class T>
inline void bezier_sample(const T* P, int S, T* road /*[2S]*/) {
  for (int i = 0; i < S; ++i) {
    T j = T(double(i)) / T(double(S - 1));  // np.linspace(0, 1, S)
    bezier_point(P, j, &road[i], &road[S + i]);
  }
}

// bezier_curves.py:28-48 / game_theory.py:92-113.
inline void bezier_control_points(double i, double h, double W, double theta, double l, double Lf, double v0,
                                  double v1, double D1, double* P /*[12]*/, double* tca) {
  double Li = Lf + l;
  double Di = Li * std::cos(std::atan2(W, 2 * Lf) - theta);
  double tc1 = D1 / (v0 - v1);
  double Px2 = v0 * tc1 - Di, Px3 = Px2;
  double Px5 = 2 * Px2;
  double Px1 = (Px2 - 0.0) / i;
  double Px4 = Px5 - (Px5 - Px3) / i;
  P[0] = 0; P[1] = Px1; P[2] = Px2; P[3] = Px3; P[4] = Px4; P[5] = Px5;
  P[6] = 0; P[7] = 0; P[8] = 0; P[9] = h; P[10] = h; P[11] = h;
  *tca = Px2 / (v0 - v1);
}

// dpf_test.py:45-54
template <class T>
inline T u_trgl(const T& y, const T& y_t, const T& a) { return a * (y - y_t) * (y - y_t); }
template <class T>
inline T u_rbd(const T& y, const T& y_br, const T& y_bl, const T& b) {
  if (y >= y_bl) return b * (y - y_bl) * (y - y_bl);
  if (y <= y_br) return b * (y - y_br) * (y - y_br);
  return T(0);
}

// dpf_test.py:81-102, one obstacle; optional gradient wrt (x, y, phi, v) ADDED to g4.
template <class T>
inline T dnf_one(const T& x, const T& y, const T& phi, const T& v, const T& xo, const T& yo, const T& phio,
                 const T& vo, const T& Af, const T& b, const T& sx, const T& sy, const T& adiv, T* g4) {
  T th = phi - phio;
  T c = cos(th), s = sin(th);
  T xi = c * xo - s * yo, yi = s * xo + c * yo;  // rotate(x_obs, y_obs, theta)
  T xr = c * x - s * y, yr = s * x + c * y;      // rotate(x, y, theta)
  T a = xr - xi, bb = yr - yi;
  T al = (v - vo) / adiv;
  T q = a * a / (T(2) * sx * sx) + bb * bb / (T(2) * sy * sy);
  bool b1 = (b == T(1));
  T qb = b1 ? q : pow(q, b);
  T U = Af * exp(-qb) * exp(-al * a);
  if (g4) {
    T dq = b1 ? T(1) : b * pow(q, b - T(1));
    T Ua = U * (-dq * a / (sx * sx) - al);
    T Ub = U * (-dq * bb / (sy * sy));
    g4[0] = g4[0] + Ua * c + Ub * s;
    g4[1] = g4[1] - Ua * s + Ub * c;
    g4[2] = g4[2] + Ua * (-bb) + Ub * a;
    g4[3] = g4[3] - a * U / adiv;
  }
  return U;
}

// Obstacle motion over the horizon, event_triggered.py:18-28: SE = x + u_x t + a t^2 / 2 along the
// obstacle's heading; t = time of the stage.  obs = [x, y, phi, v, a].
template <class T>
inline void obstacle_at(const T* obs, const T& t, T* xo, T* yo, T* phio, T* vo) {
  T s = obs[3] * t + T(0.5) * obs[4] * t * t;
  *xo = obs[0] + s * cos(obs[2]);
  *yo = obs[1] + s * sin(obs[2]);
  *phio = obs[2];
  *vo = obs[3] + obs[4] * t;
}

// One instance's problem data.
template <class T>
struct Instance {
  const MpcProblemDesc* d;
  Model<T> M;
  const T* road;  // [2S] polyline of THIS instance (already sampled if Bezier)
  const T* obs;   // [n_obs,5] or null
  const T* lam;   // [nx*N] ALM multipliers or null
  const T* Sig;   // [nx*N] ALM penalties or null
};

// Augmented-Lagrangian pieces for one constraint row (Appendix A): zeta = g + lam/Sigma,
// yhat = Sigma (zeta - Pi_D zeta); returns the psi contribution Sigma/2 dist^2.
template <class T>
inline T alm_row(const T& g, const T& lam, const T& Sig, double lo, double hi, T* yhat, T* err_z) {
  T zeta = g + lam / Sig;
  T pr = zeta < T(lo) ? T(lo) : (zeta > T(hi) ? T(hi) : zeta);
  T dz = zeta - pr;
  *yhat = Sig * dz;
  if (err_z) *err_z = g - pr;
  return T(0.5) * Sig * dz * dz;
}

// Rows of g per stage: the nx state rows of main.py:46-51, plus one safe-distance ellipse per obstacle with
// MPCB_FLAG_SAFE_DIST.
inline int rows_of(const MpcProblemDesc& d) {
  return nx_of(d.model_id) + ((d.flags & MPCB_FLAG_SAFE_DIST) ? d.n_obs : 0);
}

// dpf_test.py:57-64 get_safe_distances: X = (x, y, v_x, v_y) in the world frame.
template <class T>
inline void safe_distances(const MpcProblemDesc& d, const T& vxe, const T& vye, const T& vxo, const T& vyo, T* xs, T* ys) {
  *xs = T(d.safe_x0) / T(2) + (vxe - vxo) * (vxe - vxo) / (T(2) * T(d.safe_ax_max));
  *ys = T(d.safe_y0) / T(2) + (vye - vyo) * (vye - vyo) / (T(2) * T(d.safe_ay_max));
}

// The ellipse of semi-axes (x_s, y_s) around an obstacle as a constraint row: g = 1 - (dx/x_s)^2 - (dy/y_s)^2 <= 0.
// The ego's world-frame velocity is (v cos phi, v sin phi) for Kinematic-4 and R(phi) (vx, vy) for Pacejka-6; gy
// (nullable, nx entries, ACCUMULATED with weight w) receives w * dg/dy.
template <class T>
inline T safe_dist_row(const MpcProblemDesc& d, const T* y, const T& xo, const T& yo, const T& po, const T& vo, const T& w, T* gy) {
  const bool kin = d.model_id == MPCB_MODEL_KINEMATIC4;
  T s = sin(y[2]), c = cos(y[2]);
  T vx = kin ? y[3] * c : y[3] * c - y[4] * s, vy = kin ? y[3] * s : y[3] * s + y[4] * c;
  T vxo = vo * cos(po), vyo = vo * sin(po);
  T xs, ys;
  safe_distances(d, vx, vy, vxo, vyo, &xs, &ys);
  T dx = y[0] - xo, dy = y[1] - yo;
  T g = T(1) - dx * dx / (xs * xs) - dy * dy / (ys * ys);
  if (gy) {
    T gvx = T(2) * dx * dx / (xs * xs * xs) * (vx - vxo) / T(d.safe_ax_max);   // dg/dv_x through x_s
    T gvy = T(2) * dy * dy / (ys * ys * ys) * (vy - vyo) / T(d.safe_ay_max);
    gy[0] = gy[0] - w * T(2) * dx / (xs * xs);
    gy[1] = gy[1] - w * T(2) * dy / (ys * ys);
    gy[2] = gy[2] + w * (gvx * (-vy) + gvy * vx);                               // d(v_x, v_y)/dphi = (-v_y, v_x)
    if (kin) {
      gy[3] = gy[3] + w * (gvx * c + gvy * s);
    } else {
      gy[3] = gy[3] + w * (gvx * c + gvy * s);
      gy[4] = gy[4] + w * (-gvx * s + gvy * c);
    }
  }
  return g;
}

// Stage cost L(y_{n+1}, u_n) of car_dynamics.py:230-258 plus the optional terms; gy/gu (nullable)
// receive dL/dy, dL/du with the selection indices held constant and d wrap/da = 1.
template <class T>
inline T stage_cost(const Instance<T>& I, int n, const T* y, const T* u, int* idx_out, T* gy, T* gu) {
  const MpcProblemDesc& d = *I.d;
  const int S = d.S, nx = I.M.nx();
  const bool kin = d.model_id == MPCB_MODEL_KINEMATIC4;
  T err[3];
  int ip, in;
  int idx = road_errors(S, I.road, y[0], y[1], y[2], false, err, &ip, &in);
  if (idx_out) *idx_out = idx;
  T speed = kin ? y[3] : sqrt(y[3] * y[3] + y[4] * y[4]);
  T dv = speed - T(d.v_ref);
  T L = T(d.c[0]) * dv * dv + T(d.c[1]) * err[0] * err[0] + T(d.c[2]) * err[2] * err[2] +
        T(d.c[3]) * err[1] * err[1] + T(d.c[4]) * u[1] * u[1] + T(d.c[5]) * u[0] * u[0];
  if (gy) {
    for (int i = 0; i < nx; ++i) gy[i] = T(0);
    const T* r = I.road;
    T wx = r[idx] - r[ip], wy = r[S + idx] - r[S + ip];
    T w2x = r[in] - r[idx], w2y = r[S + in] - r[S + idx];
    T a1 = T(2) * T(d.c[1]) * err[0], a2 = T(2) * T(d.c[2]) * err[2];
    gy[0] = a1 * wy + a2 * w2y;
    gy[1] = -a1 * wx - a2 * w2x;
    gy[2] = -T(2) * T(d.c[3]) * err[1];
    if (kin) {
      gy[3] = T(2) * T(d.c[0]) * dv;
    } else {
      T k = T(2) * T(d.c[0]) * dv / speed;  // NaN at standstill, as the reference's AD would give
      gy[3] = k * y[3];
      gy[4] = k * y[4];
    }
    gu[0] = T(2) * T(d.c[5]) * u[0];
    gu[1] = T(2) * T(d.c[4]) * u[1];
  }
  if (d.flags & MPCB_FLAG_LANE) {
    L = L + u_trgl(y[1], T(d.lane_y_t), T(d.lane_a)) + u_rbd(y[1], T(d.lane_y_br), T(d.lane_y_bl), T(d.lane_b));
    if (gy) {
      gy[1] = gy[1] + T(2) * T(d.lane_a) * (y[1] - T(d.lane_y_t));
      if (y[1] >= T(d.lane_y_bl)) gy[1] = gy[1] + T(2) * T(d.lane_b) * (y[1] - T(d.lane_y_bl));
      else if (y[1] <= T(d.lane_y_br)) gy[1] = gy[1] + T(2) * T(d.lane_b) * (y[1] - T(d.lane_y_br));
    }
  }
  if ((d.flags & MPCB_FLAG_OBST) && I.obs) {
    T t = T(double(n + 1)) * T(d.Ts);
    for (int o = 0; o < d.n_obs; ++o) {
      T xo, yo, po, vo;
      obstacle_at(I.obs + MPCB_OBS_DIM * o, t, &xo, &yo, &po, &vo);
      T g4[4] = {T(0), T(0), T(0), T(0)};
      L = L + dnf_one(y[0], y[1], y[2], y[3], xo, yo, po, vo, T(d.obs_Af), T(d.obs_b), T(d.obs_sigma_x),
                      T(d.obs_sigma_y), T(d.obs_alpha_div), gy ? g4 : nullptr);
      if (gy) for (int i = 0; i < 4; ++i) gy[i] = gy[i] + g4[i];
    }
  }
  const int R = rows_of(d);
  if ((d.flags & MPCB_FLAG_BOUNDED_D) && I.lam && I.Sig) {
    for (int i = 0; i < nx; ++i) {
      T g = y[i] * y[i] - T(d.g_offset[i]);  // main.py:46-51
      T yh;
      L = L + alm_row(g, I.lam[n * R + i], I.Sig[n * R + i], d.D_lower[i], d.D_upper[i], &yh, (T*)nullptr);
      if (gy) gy[i] = gy[i] + yh * T(2) * y[i];
    }
  }
  if ((d.flags & MPCB_FLAG_SAFE_DIST) && I.obs && I.lam && I.Sig) {
    T t = T(double(n + 1)) * T(d.Ts);
    for (int o = 0; o < d.n_obs; ++o) {
      T xo, yo, po, vo;
      obstacle_at(I.obs + MPCB_OBS_DIM * o, t, &xo, &yo, &po, &vo);
      T g = safe_dist_row(d, y, xo, yo, po, vo, T(0), (T*)nullptr);
      T yh;
      L = L + alm_row(g, I.lam[n * R + nx + o], I.Sig[n * R + nx + o], -INFINITY, 0.0, &yh, (T*)nullptr);
      if (gy) safe_dist_row(d, y, xo, yo, po, vo, yh, gy);
    }
  }
  return L;
}

// Forward rollout (car_dynamics.py:158-166 mapaccum + main.py:34-40 cost sum).
//   states [N*nx] (nullable): state after each input;  chk [N*n_sub*nx] (nullable): element starts.
template <class T>
inline T rollout(const Instance<T>& I, const T* x0, const T* U, T* states, int* idx, T* chk) {
  const MpcProblemDesc& d = *I.d;
  const int nx = I.M.nx();
  T h = T(d.Ts) / T(double(d.n_sub));
  T y[6], yn[6];
  for (int i = 0; i < nx; ++i) y[i] = x0[i];
  T cost = T(0);
  for (int n = 0; n < d.N; ++n) {
    const T* u = U + 2 * n;
    for (int e = 0; e < d.n_sub; ++e) {
      if (chk) for (int i = 0; i < nx; ++i) chk[(n * d.n_sub + e) * nx + i] = y[i];
      rk4_element(I.M, y, u, h, yn);
      for (int i = 0; i < nx; ++i) y[i] = yn[i];
    }
    if (states) for (int i = 0; i < nx; ++i) states[n * nx + i] = y[i];
    cost = cost + stage_cost(I, n, y, u, idx ? idx + n : nullptr, (T*)nullptr, (T*)nullptr);
  }
  return cost;
}

// psi and grad_U psi: forward with element checkpoints, then the reverse sweep.
template <class T>
inline T value_grad(const Instance<T>& I, const T* x0, const T* U, T* grad) {
  const MpcProblemDesc& d = *I.d;
  const int nx = I.M.nx(), N = d.N;
  std::vector<T> chk(size_t(N) * d.n_sub * nx), states(size_t(N) * nx);
  T cost = rollout(I, x0, U, states.data(), (int*)nullptr, chk.data());
  T h = T(d.Ts) / T(double(d.n_sub));
  T lam[6];
  for (int i = 0; i < nx; ++i) lam[i] = T(0);
  for (int n = N - 1; n >= 0; --n) {
    const T* u = U + 2 * n;
    T gy[6], gu[2];
    stage_cost(I, n, &states[n * nx], u, (int*)nullptr, gy, gu);
    for (int i = 0; i < nx; ++i) lam[i] = lam[i] + gy[i];
    T ub[2] = {gu[0], gu[1]};
    for (int e = d.n_sub - 1; e >= 0; --e) rk4_element_adj(I.M, &chk[(n * d.n_sub + e) * nx], u, h, lam, ub);
    grad[2 * n] = ub[0];
    grad[2 * n + 1] = ub[1];
  }
  return cost;
}

}  // namespace orc
