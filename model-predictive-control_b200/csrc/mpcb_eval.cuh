// mpcb_eval.cuh — one instance's psi / grad psi evaluation on one thread ("instance-per-lane").
//
// Forward: RK4 single-shooting rollout over the horizon (car_dynamics.py:158-166, main.py:34-40),
// element-start states checkpointed to a lane-interleaved scratch so the warp's 32 instances store
// 32 consecutive doubles (coalesced).  Reverse: costate sweep re-linearising every model call from
// its checkpoint (recompute beats storing the linearisation: 30 flop/B vs a 5.7 flop/B ridge).
#pragma once
#include "mpcb_cost.cuh"

namespace mpcb {

// Strided view of a per-instance vector: row-major rows (stride 1) or lane-interleaved scratch.
struct Vec {
  double* p;
  int64_t s;
  MPCB_DEV double operator()(int i) const { return p[(int64_t)i * s]; }
  MPCB_DEV void set(int i, double v) const { p[(int64_t)i * s] = v; }
};
struct IVec {
  int32_t* p;
  int64_t s;
  MPCB_DEV int operator()(int i) const { return p[(int64_t)i * s]; }
  MPCB_DEV void set(int i, int v) const { p[(int64_t)i * s] = v; }
};

// Scratch of one evaluation: element checkpoints, stage-end states, nearest indices.
struct EvalScratch {
  Vec chk;   // [(n*n_sub+e)*NX + i]
  Vec st;    // [n*NX + i]
  IVec idx;  // [n]
};

template <int MODEL>
MPCB_DEV double eval_forward(const KDesc& d, const VehP& v, const RoadView& r, const Extras& ex, const double* x0,
                             const Vec& U, const EvalScratch& sc, bool STORE) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  double y[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) y[i] = x0[i];
  double cost = 0.0;
#pragma unroll 1
  for (int n = 0; n < d.N; ++n) {
    double ud = U(2 * n), udl = U(2 * n + 1);
    StageU su;
    make_stage_u<MODEL>(su, ud, udl, v, false);
#pragma unroll 1
    for (int e = 0; e < d.n_sub; ++e) {
      if (STORE) {
#pragma unroll
        for (int i = 0; i < NX; ++i) sc.chk.set((n * d.n_sub + e) * NX + i, y[i]);
      }
      rk4_element<MODEL>(y, su, v, d.h);
    }
    int idx;
    cost += stage_cost<MODEL, false, true>(d, r, ex, n, y, ud, udl, &idx, nullptr, nullptr);
    if (STORE) {
#pragma unroll
      for (int i = 0; i < NX; ++i) sc.st.set(n * NX + i, y[i]);
      sc.idx.set(n, idx);
    }
  }
  return cost;
}

template <int MODEL>
MPCB_DEV void eval_reverse(const KDesc& d, const VehP& v, const RoadView& r, const Extras& ex, const Vec& U,
                           const EvalScratch& sc, const Vec& grad) {
  constexpr int NX = ModelTraits<MODEL>::NX;
  double lam[NX];
#pragma unroll
  for (int i = 0; i < NX; ++i) lam[i] = 0.0;
#pragma unroll 1
  for (int n = d.N - 1; n >= 0; --n) {
    double ud = U(2 * n), udl = U(2 * n + 1);
    StageU su;
    make_stage_u<MODEL>(su, ud, udl, v, true);
    double y[NX], gy[NX], ub[2];
#pragma unroll
    for (int i = 0; i < NX; ++i) y[i] = sc.st(n * NX + i);
    int idx = sc.idx(n);
    stage_cost<MODEL, true, false>(d, r, ex, n, y, ud, udl, &idx, gy, ub);
#pragma unroll
    for (int i = 0; i < NX; ++i) lam[i] += gy[i];
#pragma unroll 1
    for (int e = d.n_sub - 1; e >= 0; --e) {
#pragma unroll
      for (int i = 0; i < NX; ++i) y[i] = sc.chk((n * d.n_sub + e) * NX + i);
      rk4_element_adj<MODEL>(y, su, v, d.h, lam, ub);
    }
    grad.set(2 * n, ub[0]);
    grad.set(2 * n + 1, ub[1]);
  }
}

// Stage the shared road table into shared memory as interleaved (x, y) pairs: one 16-byte broadcast
// load per vertex in the scan.  Coalesced global reads, done once per CTA.
MPCB_DEV void stage_roads(double* sroad, const double* __restrict__ road, int R, int S) {
  for (int t = threadIdx.x; t < R * S; t += blockDim.x) {
    int rr = t / S, i = t - rr * S;
    sroad[2 * t] = road[(int64_t)rr * 2 * S + i];
    sroad[2 * t + 1] = road[(int64_t)rr * 2 * S + S + i];
  }
}

// Road source as passed to the evaluation kernels.
struct RoadArg {
  const double* road;       // table [R,2S] (modes SHARED/TABLE) or transposed scratch [2S, ld] (other modes)
  const int32_t* road_idx;  // [B] or null
  int64_t ld;               // leading dimension of the transposed scratch
  int in_smem;              // 1: stage `road` into shared memory
  int R;
};

MPCB_DEV RoadView road_view(const RoadArg& ra, const double* sroad, int S, int64_t b) {
  RoadView r;
  if (ra.in_smem) {
    int ridx = ra.road_idx ? ra.road_idx[b] : 0;
    r.rx = sroad + (int64_t)ridx * 2 * S;
    r.ry = r.rx + 1;
    r.stride = 2;
  } else {
    r.rx = ra.road + b;
    r.ry = ra.road + (int64_t)S * ra.ld + b;
    r.stride = ra.ld;
  }
  return r;
}

}  // namespace mpcb
