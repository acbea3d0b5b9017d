// mpcb_eval.cuh — strided per-lane vector views and road staging shared by the evaluation kernels.
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
    r.rx = nullptr; r.ry = nullptr; r.stride = 0;
    r.sbase = (unsigned)__cvta_generic_to_shared(sroad + (int64_t)ridx * 2 * S);
    r.smem = 1;
  } else {
    r.rx = ra.road + b;
    r.ry = ra.road + (int64_t)S * ra.ld + b;
    r.stride = ra.ld;
    r.sbase = 0;
    r.smem = 0;
  }
  return r;
}

}  // namespace mpcb
