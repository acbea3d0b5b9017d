// mpcb_cost.cuh — nearest vertex, tracking errors, potentials and the stage cost (+ its gradient).
// Reference: car_dynamics.py:174-258, road.py:22-81, dpf_test.py:45-102, event_triggered.py:26,
// main.py:43-52 (state constraints through the augmented Lagrangian).
#pragma once
#include "mpcb_device.cuh"

namespace mpcb {

// A road as the kernels see it.
//   shared-memory road (table modes): interleaved (x, y) pairs at shared address sbase + 16 i, read with
//     one explicit ld.shared.v2.f64 (through a generic pointer the compiler emitted LD.E + spilled
//     address arithmetic in the scan loop);
//   SoA scratch road (per-instance modes): rx = roadT + b, ry = roadT + S*ld + b, stride ld.
struct RoadView {
  const double* rx;
  const double* ry;
  int64_t stride;
  unsigned sbase;
  int smem;
  // shared-memory roads carry two derived tables behind the vertices (stage_roads, mpcb_eval.cuh):
  // the heading of every segment, and an (xmin, xmax, ymin, ymax) box per chunk of MPCB_SCAN_CHUNK candidates
  unsigned hbase, bbase;
  MPCB_DEV void xy(int i, double& x, double& y) const {
    if (smem) {
      asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(sbase + 16u * (unsigned)i));
    } else {
      x = rx[(int64_t)i * stride];
      y = ry[(int64_t)i * stride];
    }
  }
  MPCB_DEV double x(int i) const { double a, b; xy(i, a, b); return a; }
  MPCB_DEV double y(int i) const { double a, b; xy(i, a, b); return b; }
};

#define MPCB_PI 3.141592653589793238462643383279502884
#define MPCB_SCAN_CHUNK 8
// doubles one staged road occupies in shared memory: S vertices (x, y), S headings, 4 per chunk box
__host__ __device__ inline int road_box_offset(int S) { return (3 * S + 1) & ~1; }  // 16-byte aligned for any S
__host__ __device__ inline int road_smem_doubles(int S) {
  return road_box_offset(S) + 4 * ((S - 1 + MPCB_SCAN_CHUNK - 1) / MPCB_SCAN_CHUNK);
}

// np.mod(a + pi, 2 pi) - pi: floor-mod via fmod + sign fix (car_dynamics.py:168-172).  fmod is exact, so
// within one period either side it is the identity or one exact subtraction (Sterbenz); only angles
// further out take the library loop.
MPCB_DEV double wrap_to_pi(double a) {
  const double two_pi = 2.0 * MPCB_PI;
  const double t = a + MPCB_PI;
  double m;
  if (t >= 0.0 && t < two_pi) m = t;
  else if (t >= two_pi && t < 2.0 * two_pi) m = t - two_pi;
  else if (t < 0.0 && t > -two_pi) m = t + two_pi;
  else {
    m = fmod(t, two_pi);
    if (m < 0.0) m += two_pi;
  }
  return m - MPCB_PI;
}

MPCB_DEV double d2_of(double x, double y, double px, double py) {
  double dx = x - px, dy = y - py;
  return __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));  // rounded like diff**2 summed (no FMA)
}
MPCB_DEV double road_d2(const RoadView& r, int i, double px, double py) {
  double x, y;
  r.xy(i, x, y);
  return d2_of(x, y, px, py);
}

// Exact restatement of car_dynamics.py:174-192 (sqrt'ed distances, strict '<', i = 1..S-2 from vertex 0):
// only runs when two vertices are within rounding of the minimum.
__device__ __noinline__ int nearest_scan_exact(RoadView r, int S, double px, double py) {
  double m = sqrt(road_d2(r, 0, px, py));
  int best = 0;
  for (int i = 1; i < S - 1; ++i) {
    double di = sqrt(road_d2(r, i, px, py));
    if (di < m) { m = di; best = i; }
  }
  return best;
}

// car_dynamics.py:174-192.  The reference compares sqrt'ed distances with a strict '<', so its answer is
// the lowest index attaining the minimal sqrt.  sqrt is monotone: that is the first index of the minimal
// squared distance unless another vertex lies within rounding of it.  The scan therefore tracks the
// minimum, its first index and the second-smallest entry with selects only (two independent chains
// over the two halves of the range, no branches, no sqrt), and falls back to the exact scan when the
// runner-up is within 4.5e-16 of the minimum — same winner, bit for bit.  Not inlined: its own
// register allocation (inside the gradient evaluation it was reloading spilled addresses per vertex).
struct ScanLoadShared {
  unsigned a;
  MPCB_DEV void at(int i, double& x, double& y) const {
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(a + 16u * (unsigned)i));
  }
};
struct ScanLoadStrided {
  const double* rx;
  const double* ry;
  int64_t stride;
  MPCB_DEV void at(int i, double& x, double& y) const { x = rx[(int64_t)i * stride]; y = ry[(int64_t)i * stride]; }
};
template <class LD>
MPCB_DEV bool nearest_scan_core(const LD& ld, int S, double px, double py, int* best_out) {
  const int n = S - 1;         // candidates 0 .. S-2
  const int h = (n + 1) >> 1;  // chain A: [0, h), chain B: [h, n)
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  double x, y;
  ld.at(0, x, y);
  double ma = d2_of(x, y, px, py), sa = inf, mb = inf, sb = inf;
  int ba = 0, bb = n;
#pragma unroll 4
  for (int i = 1; i < h; ++i) {
    const int j = h + i - 1;  // < n for every i < h
    double xa, ya, xb, yb;
    ld.at(i, xa, ya);
    ld.at(j, xb, yb);
    const double da = d2_of(xa, ya, px, py), db = d2_of(xb, yb, px, py);
    const bool la = da < ma, lb = db < mb;
    sa = la ? ma : fmin(sa, da);
    sb = lb ? mb : fmin(sb, db);
    ba = la ? i : ba;
    bb = lb ? j : bb;
    ma = la ? da : ma;
    mb = lb ? db : mb;
  }
  if (2 * h - 1 < n) {  // n even: chain B holds one more element than the loop visited
    const int j = 2 * h - 1;
    ld.at(j, x, y);
    const double db = d2_of(x, y, px, py);
    const bool lb = db < mb;
    sb = lb ? mb : fmin(sb, db);
    bb = lb ? j : bb;
    mb = lb ? db : mb;
  }
  const bool lb = mb < ma;  // merge: the lower-index chain wins ties
  const double m = lb ? mb : ma;
  const double second = fmin(fmin(sa, sb), lb ? ma : mb);
  *best_out = lb ? bb : ba;
  return second > m * (1.0 + 4.5e-16);  // false: near-tie (or NaN) -> the exact rule decides
}
__device__ __noinline__ int nearest_scan_shared(unsigned sbase, int S, double px, double py) {
  int best;
  if (nearest_scan_core(ScanLoadShared{sbase}, S, px, py, &best)) return best;
  return -1;
}
__device__ __noinline__ int nearest_scan_strided(const double* rx, const double* ry, int64_t stride, int S, double px,
                                                 double py) {
  int best;
  if (nearest_scan_core(ScanLoadStrided{rx, ry, stride}, S, px, py, &best)) return best;
  return -1;
}
// Shared-memory roads: the same search, pruned.  The chunk holding `hint` (the previous stage's winner; the
// vehicle moves a fraction of a segment per stage) is scanned first; every other chunk is scanned only if
// the distance from (px, py) to its bounding box does not exceed the running minimum.  The box distance is
// built from the same rounded operations as d2_of and is monotone in |dx|, |dy|, so it never exceeds the
// d2 of a vertex inside the box; the 1e-15 slack keeps every vertex that could be a near-tie in the
// scan, so the runner-up test (and the exact fallback behind it) sees what the full scan would see.
// Returns -1 on a near-tie / NaN, like nearest_scan_shared.
__device__ __noinline__ int nearest_scan_pruned(unsigned sbase, unsigned bbase, int S, double px, double py, int hint) {
  const int n = S - 1;  // candidates 0 .. S-2
  const int nc = (n + MPCB_SCAN_CHUNK - 1) / MPCB_SCAN_CHUNK;
  const double inf = __longlong_as_double(0x7ff0000000000000LL);
  double m = inf, second = inf;
  int best = 0;
  auto scan_chunk = [&](int c) {
    const int i0 = c * MPCB_SCAN_CHUNK;
#pragma unroll
    for (int k = 0; k < MPCB_SCAN_CHUNK; ++k) {
      const int i = i0 + k;
      if (i < n) {
        double x, y;
        asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x), "=d"(y) : "r"(sbase + 16u * (unsigned)i));
        const double dd = d2_of(x, y, px, py);
        const bool lt = dd < m;
        second = lt ? m : fmin(second, dd);
        best = lt ? i : best;
        m = lt ? dd : m;
      }
    }
  };
  const int hc = hint >= 0 && hint < n ? hint / MPCB_SCAN_CHUNK : 0;
  scan_chunk(hc);
#pragma unroll 1
  for (int c = 0; c < nc; ++c) {
    if (c == hc) continue;
    double x0, x1, y0, y1;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(x0), "=d"(x1) : "r"(bbase + 32u * (unsigned)c));
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(y0), "=d"(y1) : "r"(bbase + 32u * (unsigned)c + 16u));
    const double dx = fmax(fmax(x0 - px, px - x1), 0.0), dy = fmax(fmax(y0 - py, py - y1), 0.0);
    const double lb = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
    if (!(lb > m * (1.0 + 1e-15))) scan_chunk(c);  // NaN: scan
  }
  return second > m * (1.0 + 4.5e-16) ? best : -1;
}
MPCB_DEV int nearest_scan(const RoadView& r, int S, double px, double py, int hint = -1) {
  int best = r.smem ? nearest_scan_pruned(r.sbase, r.bbase, S, px, py, hint)
                    : nearest_scan_strided(r.rx, r.ry, r.stride, S, px, py);
  if (best < 0) best = nearest_scan_exact(r, S, px, py);
  return best;
}

// road.py:36-50: argmin over every vertex (first minimum of the sqrt'ed norms).
MPCB_DEV int nearest_scan_numpy(const RoadView& r, int S, double px, double py) {
  double dx = r.x(0) - px, dy = r.y(0) - py;
  double m = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
  int best = 0;
  for (int i = 1; i < S; ++i) {
    dx = r.x(i) - px;
    dy = r.y(i) - py;
    double di = sqrt(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
    if (di < m) { m = di; best = i; }
  }
  return best;
}

struct RoadGeom {  // the three vertices the errors are built from, as difference vectors
  double nx, ny;    // nearest
  double wx, wy;    // nearest - previous
  double w2x, w2y;  // next - nearest
  double ax, ay;    // previous
  double hd;        // atan2(w2y, w2x)
};

MPCB_DEV void road_geom(const RoadView& r, int idx, RoadGeom& g) {
  int ip = idx > 0 ? idx - 1 : 0;  // vertex 0 wins: previous = vertex 0 (car_dynamics.py:183)
  int in = idx + 1;
  double qx, qy;
  r.xy(idx, g.nx, g.ny);
  r.xy(ip, g.ax, g.ay);
  r.xy(in, qx, qy);
  g.wx = g.nx - g.ax; g.wy = g.ny - g.ay;
  g.w2x = qx - g.nx; g.w2y = qy - g.ny;
  // staged roads hold this very atan2 of these very operands per segment (stage_roads)
  if (r.smem) asm volatile("ld.shared.f64 %0, [%1];" : "=d"(g.hd) : "r"(r.hbase + 8u * (unsigned)idx));
  else g.hd = atan2(g.w2y, g.w2x);
}

// car_dynamics.py:194-228: cte, heading error, positional error (raw cross products).
MPCB_DEV void road_errors(const RoadGeom& g, double px, double py, double phi, double* cte, double* he,
                          double* pe) {
  double vx = px - g.ax, vy = py - g.ay;
  *cte = fma(vx, g.wy, -(vy * g.wx));
  *he = wrap_to_pi(g.hd - phi);
  double v2x = px - g.nx, v2y = py - g.ny;
  *pe = fma(v2x, g.w2y, -(v2y * g.w2x));
}

// dpf_test.py:81-102, one obstacle; g4 (nullable) accumulates d/d(x, y, phi, v).
MPCB_DEV double dnf_one(double x, double y, double phi, double v, double xo, double yo, double phio, double vo,
                        const KDesc& d, double* g4) {
  double s, c;
  sincos(phi - phio, &s, &c);
  double xi = c * xo - s * yo, yi = s * xo + c * yo;
  double xr = c * x - s * y, yr = s * x + c * y;
  double a = xr - xi, bb = yr - yi;
  double al = (v - vo) / d.obs_adiv;
  double isx2 = 1.0 / (d.obs_sx * d.obs_sx), isy2 = 1.0 / (d.obs_sy * d.obs_sy);
  double q = 0.5 * a * a * isx2 + 0.5 * bb * bb * isy2;
  bool b1 = d.obs_b == 1.0;
  double qb = b1 ? q : pow(q, d.obs_b);
  double U = d.obs_Af * exp(-qb) * exp(-al * a);
  if (g4) {
    double dq = b1 ? 1.0 : d.obs_b * pow(q, d.obs_b - 1.0);
    double Ua = U * (-dq * a * isx2 - al);
    double Ub = U * (-dq * bb * isy2);
    g4[0] += Ua * c + Ub * s;
    g4[1] += -Ua * s + Ub * c;
    g4[2] += -Ua * bb + Ub * a;
    g4[3] += -a * U / d.obs_adiv;
  }
  return U;
}

// Augmented-Lagrangian row: returns Sigma/2 dist^2, yhat out, err_z optional.
MPCB_DEV double alm_row(double g, double lam, double Sig, double lo, double hi, double* yhat, double* err_z) {
  double zeta = g + lam / Sig;
  double pr = zeta < lo ? lo : (zeta > hi ? hi : zeta);
  double dz = zeta - pr;
  *yhat = Sig * dz;
  if (err_z) *err_z = g - pr;
  return 0.5 * Sig * dz * dz;
}

// dpf_test.py:57-64 get_safe_distances + the ellipse of those semi-axes around an obstacle as a constraint row:
// g = 1 - (dx / x_s)^2 - (dy / y_s)^2 <= 0, x_s = x_0/2 + (v_x - v_x,obs)^2 / (2 a_x_max) (same in y), velocities in
// the world frame: (v cos phi, v sin phi) for Kinematic-4, R(phi) (vx, vy) for Pacejka-6.  gy (nullable) is
// ACCUMULATED with weight w: gy += w dg/dy.
template <int MODEL>
MPCB_DEV double safe_dist_row(const KDesc& d, const double* y, double xo, double yo, double po, double vo, double w, double* gy) {
  double s, c, so, co;
  sincos(y[2], &s, &c);
  sincos(po, &so, &co);
  const bool kin = MODEL == MPCB_MODEL_KINEMATIC4;
  const double vx = kin ? y[3] * c : y[3] * c - y[MODEL == MPCB_MODEL_KINEMATIC4 ? 3 : 4] * s;
  const double vy = kin ? y[3] * s : y[3] * s + y[MODEL == MPCB_MODEL_KINEMATIC4 ? 3 : 4] * c;
  const double dvx = vx - vo * co, dvy = vy - vo * so;
  const double xs = d.safe_x0 / 2.0 + dvx * dvx / (2.0 * d.safe_ax);
  const double ys = d.safe_y0 / 2.0 + dvy * dvy / (2.0 * d.safe_ay);
  const double dx = y[0] - xo, dy = y[1] - yo;
  const double g = 1.0 - dx * dx / (xs * xs) - dy * dy / (ys * ys);
  if (gy) {
    const double gvx = 2.0 * dx * dx / (xs * xs * xs) * dvx / d.safe_ax;
    const double gvy = 2.0 * dy * dy / (ys * ys * ys) * dvy / d.safe_ay;
    gy[0] -= w * 2.0 * dx / (xs * xs);
    gy[1] -= w * 2.0 * dy / (ys * ys);
    gy[2] += w * (gvx * (-vy) + gvy * vx);
    gy[3] += w * (gvx * c + gvy * s);
    if constexpr (MODEL != MPCB_MODEL_KINEMATIC4) gy[4] += w * (-gvx * s + gvy * c);
  }
  return g;
}

// Per-instance extras for the optional cost terms.
struct Extras {
  const double* obs;         // [n_obs,5] of this instance or null
  const double* lam;         // ALM multipliers of this instance: lam[(n*nx+i)*lam_stride]
  const double* Sig;         // ALM penalties, same indexing with sig_stride
  int64_t lam_stride, sig_stride;
};

// Stage cost L(y_{n+1}, u_n) (car_dynamics.py:249-258) + optional terms.  With GRAD, gy[NX]/gu[2]
// receive dL/dy, dL/du.  idx_io: forward pass writes the nearest index, reverse pass reuses it.
template <int MODEL, bool GRAD, bool SCAN>
MPCB_DEV double stage_cost(const KDesc& d, const RoadView& r, const Extras& ex, int n, const double* y, double ud,
                           double udl, int* idx_io, double* gy, double* gu) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  int idx = SCAN ? nearest_scan(r, d.S, y[0], y[1], idx_io ? *idx_io : -1) : *idx_io;  // *idx_io in: search hint
  if (SCAN && idx_io) *idx_io = idx;
  RoadGeom g;
  road_geom(r, idx, g);
  double cte, he, pe;
  road_errors(g, y[0], y[1], y[2], &cte, &he, &pe);
  double speed;
  if constexpr (MODEL == MPCB_MODEL_KINEMATIC4) speed = y[3];
  else speed = sqrt(fma(y[3], y[3], y[4] * y[4]));
  double dv = speed - d.v_ref;
  // (explicit fma() chain: the build uses -fmad=false)
  double L = fma(d.c[5] * ud, ud, fma(d.c[4] * udl, udl, fma(d.c[3] * he, he, fma(d.c[2] * pe, pe, fma(d.c[1] * cte, cte, d.c[0] * dv * dv)))));
  if (GRAD) {
    double a1 = 2.0 * d.c[1] * cte, a2 = 2.0 * d.c[2] * pe;
    gy[0] = fma(a1, g.wy, a2 * g.w2y);
    gy[1] = -fma(a1, g.wx, a2 * g.w2x);
    gy[2] = -2.0 * d.c[3] * he;
    if constexpr (MODEL == MPCB_MODEL_KINEMATIC4) {
      gy[3] = 2.0 * d.c[0] * dv;
    } else {
      double k = 2.0 * d.c[0] * dv / speed;  // NaN at standstill, like the reference's AD
      gy[3] = k * y[3];
      gy[4] = k * y[4];
      gy[5] = 0.0;
    }
    gu[0] = 2.0 * d.c[5] * ud;
    gu[1] = 2.0 * d.c[4] * udl;
  }
  if (d.flags & MPCB_FLAG_LANE) {
    double dy = y[1] - d.lane_y_t;
    L += d.lane_a * dy * dy;
    double e = 0.0;
    if (y[1] >= d.lane_y_bl) e = y[1] - d.lane_y_bl;
    else if (y[1] <= d.lane_y_br) e = y[1] - d.lane_y_br;
    L += d.lane_b * e * e;
    if (GRAD) gy[1] += 2.0 * d.lane_a * dy + 2.0 * d.lane_b * e;
  }
  if ((d.flags & MPCB_FLAG_OBST) && ex.obs) {
    double t = (double)(n + 1) * d.Ts;
    for (int o = 0; o < d.n_obs; ++o) {
      const double* ob = ex.obs + MPCB_OBS_DIM * o;
      double so, co;
      sincos(ob[2], &so, &co);
      double s = ob[3] * t + 0.5 * ob[4] * t * t;  // event_triggered.py:26
      double g4[4] = {0.0, 0.0, 0.0, 0.0};
      L += dnf_one(y[0], y[1], y[2], y[3], ob[0] + s * co, ob[1] + s * so, ob[2], ob[3] + ob[4] * t, d,
                   GRAD ? g4 : nullptr);
      if (GRAD) { gy[0] += g4[0]; gy[1] += g4[1]; gy[2] += g4[2]; gy[3] += g4[3]; }
    }
  }
  if (d.bounded && ex.lam && ex.Sig) {
    const int R = d.rows;
#pragma unroll
    for (int i = 0; i < NX; ++i) {
      double gi = y[i] * y[i] - d.g_off[i];
      double yh;
      L += alm_row(gi, ex.lam[(int64_t)(n * R + i) * ex.lam_stride], ex.Sig[(int64_t)(n * R + i) * ex.sig_stride],
                   d.D_lo[i], d.D_hi[i], &yh, nullptr);
      if (GRAD) gy[i] += 2.0 * yh * y[i];
    }
    if ((d.flags & MPCB_FLAG_SAFE_DIST) && ex.obs) {  // one ellipse row per obstacle, D = (-inf, 0]
      const double t = (double)(n + 1) * d.Ts;
      const double ninf = __longlong_as_double(0xfff0000000000000LL);
      for (int o = 0; o < d.n_obs; ++o) {
        const double* ob = ex.obs + MPCB_OBS_DIM * o;
        double so, co;
        sincos(ob[2], &so, &co);
        const double sdist = ob[3] * t + 0.5 * ob[4] * t * t;  // event_triggered.py:26
        const double xo = ob[0] + sdist * co, yo = ob[1] + sdist * so, vo = ob[3] + ob[4] * t;
        const double gi = safe_dist_row<MODEL>(d, y, xo, yo, ob[2], vo, 0.0, nullptr);
        double yh;
        L += alm_row(gi, ex.lam[(int64_t)(n * R + NX + o) * ex.lam_stride], ex.Sig[(int64_t)(n * R + NX + o) * ex.sig_stride],
                     ninf, 0.0, &yh, nullptr);
        if (GRAD) safe_dist_row<MODEL>(d, y, xo, yo, ob[2], vo, yh, gy);
      }
    }
  }
  return L;
}

// Quintic Bernstein point (bezier_curves.py:19-25).
MPCB_DEV void bezier_point(const double* P /*[12]*/, double j, double* x, double* y) {
  double u = 1.0 - j;
  double u2 = u * u, u3 = u2 * u, u4 = u2 * u2, u5 = u4 * u;
  double j2 = j * j, j3 = j2 * j, j4 = j2 * j2, j5 = j4 * j;
  double w0 = u5, w1 = 5.0 * u4 * j, w2 = 10.0 * u3 * j2, w3 = 10.0 * u2 * j3, w4 = 5.0 * u * j4, w5 = j5;
  *x = w0 * P[0] + w1 * P[1] + w2 * P[2] + w3 * P[3] + w4 * P[4] + w5 * P[5];
  *y = w0 * P[6] + w1 * P[7] + w2 * P[8] + w3 * P[9] + w4 * P[10] + w5 * P[11];
}

}  // namespace mpcb
