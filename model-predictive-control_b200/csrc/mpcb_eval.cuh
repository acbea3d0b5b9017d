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

// Stage the shared road table into shared memory, per road (road_smem_doubles(S) doubles each):
//   [0, 2S)      interleaved (x, y) pairs — one 16-byte broadcast load per vertex in the scan;
//   [2S, 3S)     heading of segment i -> i+1, the atan2 the heading error needs (car_dynamics.py:214),
//                evaluated once per CTA instead of once per stage of every evaluation;
//   [3S (even), ...)  (xmin, xmax, ymin, ymax) of every chunk of MPCB_SCAN_CHUNK scan candidates (nearest_scan_pruned).
// Coalesced global reads, done once per CTA; ends with a block barrier.
MPCB_DEV void stage_roads(double* sroad, const double* __restrict__ road, int R, int S) {
  const int RS = road_smem_doubles(S);
  const int n = S - 1, nc = (n + MPCB_SCAN_CHUNK - 1) / MPCB_SCAN_CHUNK;
  for (int t = threadIdx.x; t < R * S; t += blockDim.x) {
    int rr = t / S, i = t - rr * S;
    sroad[rr * RS + 2 * i] = road[(int64_t)rr * 2 * S + i];
    sroad[rr * RS + 2 * i + 1] = road[(int64_t)rr * 2 * S + S + i];
  }
  __syncthreads();
  for (int t = threadIdx.x; t < R * S; t += blockDim.x) {
    int rr = t / S, i = t - rr * S;
    const double* v = sroad + rr * RS;
    double hd = 0.0;
    if (i + 1 < S) hd = atan2(v[2 * i + 3] - v[2 * i + 1], v[2 * i + 2] - v[2 * i]);
    sroad[rr * RS + 2 * S + i] = hd;
  }
  for (int t = threadIdx.x; t < R * nc; t += blockDim.x) {
    int rr = t / nc, c = t - rr * nc;
    const double* v = sroad + rr * RS;
    double x0 = v[2 * c * MPCB_SCAN_CHUNK], x1 = x0, y0 = v[2 * c * MPCB_SCAN_CHUNK + 1], y1 = y0;
    for (int i = c * MPCB_SCAN_CHUNK + 1; i < (c + 1) * MPCB_SCAN_CHUNK && i < n; ++i) {
      x0 = fmin(x0, v[2 * i]); x1 = fmax(x1, v[2 * i]);
      y0 = fmin(y0, v[2 * i + 1]); y1 = fmax(y1, v[2 * i + 1]);
    }
    double* o = sroad + rr * RS + road_box_offset(S) + 4 * c;
    o[0] = x0; o[1] = x1; o[2] = y0; o[3] = y1;
  }
  __syncthreads();
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
    ridx = ridx < 0 ? 0 : (ridx >= ra.R ? ra.R - 1 : ridx);  // an out-of-range index must not read past the staged table
    r.rx = nullptr; r.ry = nullptr; r.stride = 0;
    r.sbase = (unsigned)__cvta_generic_to_shared(sroad + (int64_t)ridx * road_smem_doubles(S));
    r.hbase = r.sbase + 16u * (unsigned)S;
    r.bbase = r.sbase + 8u * (unsigned)road_box_offset(S);
    r.smem = 1;
  } else {
    r.rx = ra.road + b;
    r.ry = ra.road + (int64_t)S * ra.ld + b;
    r.stride = ra.ld;
    r.sbase = 0; r.hbase = 0; r.bbase = 0;
    r.smem = 0;
  }
  return r;
}

}  // namespace mpcb
