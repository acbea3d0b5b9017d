// mpcb_kernels.cuh — utility, rollout and value+gradient kernels (instance-per-lane).
#pragma once
#include "mpcb_fwd.cuh"

namespace mpcb {

// ---- dynamics.py:67 / :144 __call__ and CasADi `f` (car_dynamics.py:132) ------------------------
template <int MODEL>
__global__ void k_dynamics(KDesc d, int64_t B, const double* __restrict__ x, const double* __restrict__ u,
                           const double* __restrict__ p, int clip, double* __restrict__ out) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  __shared__ VehP v;
  if (threadIdx.x == 0) load_vehicle(&v, p, d);
  __syncthreads();
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double y[NX], o[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) y[i] = x[b * NX + i];
  double ud = u[2 * b], udl = u[2 * b + 1];
  if (clip) {  // dynamics.py:56-66,99
    ud = fmin(fmax(ud, -v.max_drive), v.max_drive);
    udl = fmin(fmax(udl, -v.max_steer), v.max_steer);
  }
  StageU su;
  make_stage_u<MODEL>(su, ud, udl, v, false);
  ModelTraits<MODEL>::f(y, su, v, o);
#pragma unroll
  for (int i = 0; i < NX; ++i) out[b * NX + i] = o[i];
}

// ---- main.py:145 plant step: one stage of n_sub RK4 elements ------------------------------------
template <int MODEL>
__global__ void k_plant_step(KDesc d, int64_t B, const double* __restrict__ x, const double* __restrict__ u,
                             const double* __restrict__ p, double* __restrict__ out) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  fm_load_smem();
  stage_problem(d, p);
  const VehP& v = s_veh;
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  double pq[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) pq[i] = x[b * NX + i];
  StageU su;
  make_stage_u<MODEL>(su, u[2 * b], u[2 * b + 1], v, false);
#pragma unroll 1
  for (int e = 0; e < d.n_sub; ++e) rk4_fm<MODEL, false>(pq, pq + 2, nullptr, nullptr, su, v, d.h);
#pragma unroll
  for (int i = 0; i < NX; ++i) out[b * NX + i] = pq[i];
}

// ---- car_dynamics.py:194-228 / road.py:52-81 as a batched query ---------------------------------
__global__ void k_road_errors(KDesc d, int64_t B, const double* __restrict__ pos, const double* __restrict__ heading,
                              const double* __restrict__ road, const int32_t* __restrict__ road_idx, int normalised,
                              int32_t* __restrict__ idx_out, double* __restrict__ err_out) {
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int S = d.S;
  const double* base = road;
  if (d.road_mode == MPCB_ROAD_TABLE) base = road + (int64_t)(road_idx ? road_idx[b] : 0) * 2 * S;
  else if (d.road_mode == MPCB_ROAD_PER_INSTANCE) base = road + b * 2 * S;
  RoadView r{base, base + S, 1, 0u, 0, 0u, 0u};
  double px = pos[2 * b], py = pos[2 * b + 1], phi = heading[b];
  double cte, he, pe;
  int idx;
  if (!normalised) {
    idx = nearest_scan(r, S, px, py);
    RoadGeom g;
    road_geom(r, idx, g);
    road_errors(g, px, py, phi, &cte, &he, &pe);
  } else {  // road.py: argmin over all vertices, python index wrap for idx-1, errors / segment length
    idx = nearest_scan_numpy(r, S, px, py);
    int ip = idx == 0 ? S - 1 : idx - 1;
    int in = idx + 1 < S ? idx + 1 : S - 1;
    RoadGeom g;
    g.nx = r.x(idx); g.ny = r.y(idx);
    g.ax = r.x(ip); g.ay = r.y(ip);
    g.wx = g.nx - g.ax; g.wy = g.ny - g.ay;
    g.w2x = r.x(in) - g.nx; g.w2y = r.y(in) - g.ny;
    g.hd = atan2(g.w2y, g.w2x);
    road_errors(g, px, py, phi, &cte, &he, &pe);
    cte = cte / sqrt(g.wx * g.wx + g.wy * g.wy);
    pe = pe / sqrt(g.w2x * g.w2x + g.w2y * g.w2y);
  }
  idx_out[b] = idx;
  err_out[3 * b] = cte;
  err_out[3 * b + 1] = he;
  err_out[3 * b + 2] = pe;
}

// ---- bezier_curves.py:19-25: sample every instance's quintic to S vertices -----------------------
// transposed = 0: out[B,2S] row-major (the API's road_out); 1: out[2S, ld] scratch for the solvers.
__global__ void k_bezier_sample(int S, int64_t B, const double* __restrict__ ctrl, double* __restrict__ out,
                                int transposed, int64_t ld) {
  int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= B * S) return;
  int64_t b;
  int i;
  if (transposed) { i = (int)(t / B); b = t - (int64_t)i * B; }  // consecutive threads -> consecutive b
  else { b = t / S; i = (int)(t - b * S); }
  double P[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) P[k] = ctrl[b * 12 + k];
  double j = (double)i / (double)(S - 1);  // np.linspace(0, 1, S)
  double x, y;
  bezier_point(P, j, &x, &y);
  if (transposed) { out[(int64_t)i * ld + b] = x; out[(int64_t)(S + i) * ld + b] = y; }
  else { out[b * 2 * S + i] = x; out[b * 2 * S + S + i] = y; }
}

// [B,2S] row-major -> [2S, ld] (32x32 shared-memory tile transpose, both sides coalesced).
__global__ void k_transpose_road(int64_t B, int C /*=2S*/, const double* __restrict__ in, double* __restrict__ out,
                                 int64_t ld) {
  __shared__ double tile[32][33];
  int64_t b0 = (int64_t)blockIdx.x * 32;
  int c0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int64_t b = b0 + r;
    int c = c0 + threadIdx.x;
    if (b < B && c < C) tile[r][threadIdx.x] = in[b * C + c];
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    int c = c0 + r;
    int64_t b = b0 + threadIdx.x;
    if (b < B && c < C) out[(int64_t)c * ld + b] = tile[threadIdx.x][r];
  }
}

// ---- rollout + cost (model.simulate + cost sum) ---------------------------------------------------
template <int MODEL>
__global__ void __launch_bounds__(128)
k_rollout(KDesc d, int64_t B, const double* __restrict__ x0, const double* __restrict__ U,
          const double* __restrict__ p, RoadArg ra, const double* __restrict__ obs, double* states_out,
          int32_t* idx_out, double* cost_out) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  extern __shared__ double sroad[];
  fm_load_smem();
  if (ra.in_smem) stage_roads(sroad, ra.road, ra.R, d.S);
  stage_problem(d, p);
  int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  RoadView r = road_view(ra, sroad, d.S, b);
  Extras ex{obs ? obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, nullptr, nullptr, 0, 0};
  Vec Uv{const_cast<double*>(U) + b * d.n, 1};
  // stage-end states and nearest indices go straight to the caller's row-major outputs
  double y0[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) y0[i] = x0[b * NX + i];
  double cost = eval_fm<MODEL, false>(r, ex, y0, Uv, Vec{nullptr, 0}, Vec{nullptr, 0}, false,
                                      Vec{states_out ? states_out + b * d.N * NX : nullptr, 1},
                                      IVec{idx_out ? idx_out + b * d.N : nullptr, 1});
  if (cost_out) cost_out[b] = cost;
}

// ---- psi + grad psi: forward-mode rollout + horizon adjoint (mpcb_fwd.cuh) --------------------------
// Persistent grid-stride over instances; the stage-Jacobian scratch is lane-interleaved by slot.
template <int MODEL>
__global__ void __launch_bounds__(128)
k_value_grad(KDesc d, int64_t B, const double* __restrict__ x0, const double* __restrict__ U,
             const double* __restrict__ p, RoadArg ra, const double* __restrict__ obs,
             const double* __restrict__ lam, const double* __restrict__ Sig, double* __restrict__ cost_out,
             double* __restrict__ grad_out, double* scratch, int64_t ns) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  extern __shared__ double sroad[];
  fm_load_smem();
  if (ra.in_smem) stage_roads(sroad, ra.road, ra.R, d.S);
  stage_problem(d, p);
  int64_t slot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const Vec scr{scratch + slot, ns};
  for (int64_t b = slot; b < B; b += ns) {
    RoadView r = road_view(ra, sroad, d.S, b);
    Extras ex{obs ? obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, lam ? lam + b * d.m : nullptr,
              Sig ? Sig + b * d.m : nullptr, 1, 1};
    double y0[NX];
#pragma unroll
    for (int i = 0; i < NX; ++i) y0[i] = x0[b * NX + i];
    Vec Uv{const_cast<double*>(U) + b * d.n, 1};
    Vec Gv{grad_out + b * d.n, 1};
    cost_out[b] = eval_fm<MODEL, true>(r, ex, y0, Uv, scr, Gv, true);
  }
}

// ---- dependent-chain latency probe (one warp): cycles per DFMA / div / sqrt / sincos / atan / atan2 ----
__global__ void k_latency_probe(double* out, double seed, int iters) {
  double x = seed + threadIdx.x * 1e-3;
  long long t0, t1;
  auto tic = [&]() { __syncwarp(); t0 = clock64(); };
  auto toc = [&](int slot) { t1 = clock64(); if (threadIdx.x == 0) out[slot] = (double)(t1 - t0) / iters; __syncwarp(); };
  tic(); for (int i = 0; i < iters; ++i) x = fma(x, 0.999999, 1e-7); toc(0);
  tic(); for (int i = 0; i < iters; ++i) x = 1.0 / (x + 1.5); toc(1);
  tic(); for (int i = 0; i < iters; ++i) x = sqrt(x + 1.5); toc(2);
  tic(); for (int i = 0; i < iters; ++i) { double s, c; sincos(x, &s, &c); x = s + 0.5 * c; } toc(3);
  tic(); for (int i = 0; i < iters; ++i) x = atan(x * 1.3 + 0.2); toc(4);
  tic(); for (int i = 0; i < iters; ++i) x = atan2(x + 0.1, 1.0 + 0.3 * x); toc(5);
  tic(); for (int i = 0; i < iters; ++i) x = sin(x * 1.1); toc(6);
  // one Pacejka model call in a dependent chain
  {
    __shared__ VehP v;
    if (threadIdx.x == 0) {
      v.lf = 4.7e-2; v.lr = 5e-2; v.m = 0.1735; v.iz = 18.3e-5; v.inv_m = 1 / 0.1735; v.inv_iz = 1 / 18.3e-5;
      v.cm1 = 0.266; v.cm2 = 0.1; v.cr0 = 0.1025; v.cr2 = 0.0011; v.bf = 0.268; v.cf = 2.165; v.df = 3.47;
      v.br = 0.242; v.cr = 2.38; v.dr = 2.84;
    }
    __syncwarp();
    StageU su; su.d = 0.5; su.dl = 0.1; sincos(0.1, &su.sdl, &su.cdl);
    double y[6] = {0, 0, 0.1 * x, 1.0, 0.01, 0.1}, o[6];
    tic();
    for (int i = 0; i < iters; ++i) {
      pacejka_f(y, su, v, o);
#pragma unroll
      for (int k = 0; k < 6; ++k) y[k] = fma(1e-4, o[k], y[k]);
    }
    toc(7);
    x += y[0] + y[3] + y[5];
  }
  out[8 + threadIdx.x] = x;
}

// ---- FP64 FMA-chain peak probe ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int iters) {
  double a0 = 1.0 + threadIdx.x * 1e-9, a1 = a0 + 0.1, a2 = a0 + 0.2, a3 = a0 + 0.3;
  double a4 = a0 + 0.4, a5 = a0 + 0.5, a6 = a0 + 0.6, a7 = a0 + 0.7;
  const double m = 0.999999, c = 1e-7;
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
  }
  out[(int64_t)blockIdx.x * blockDim.x + threadIdx.x] = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
}

}  // namespace mpcb
