// mpcb_api.cu — extern "C" boundary of libmpcb200.so (see include/mpcb200.h).
// Host side only: argument checks, scratch management, kernel launches.  No CPU compute fallback:
// every entry point fails with MPCB_E_NODEVICE / MPCB_E_CUDA when the GPU path cannot run.
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <string>
#include <algorithm>
#include <utility>
#include <vector>
#include "mpcb_kernels.cuh"
#include "mpcb_solve.cuh"
#include "mpcb_game.cuh"

using namespace mpcb;

struct MpcHandle {
  MpcProblemDesc desc;
  KDesc kd;
  int device = 0;
  int num_sms = 148;
  std::string err;
  int64_t launches = 0;
  int64_t h2d_bytes = 0, d2h_bytes = 0;
  // grow-only device scratch
  double* road_t = nullptr; size_t road_t_cap = 0;     // transposed per-instance roads [2S, ld]
  double* ws = nullptr; size_t ws_cap = 0;             // solver / value_grad workspace
  double* wst = nullptr; size_t wst_cap = 0;           // per-tile solver workspace
  long long* dbg = nullptr; size_t dbg_cap = 0; int64_t dbg_warps = 0; int debug_timers = 0;  // MPCB_DEBUG_TIMERS=1
  int warps_per_sm = 8;                                // MPCB_WARPS_PER_SM env (experiments)
  int tile_override = 0;                               // MPCB_TILE env: force the lanes-per-instance width
  int last_tile = 0;
  int mode_override = -1;                              // MPCB_MODE env: 0 persistent kernel, 1 rounds
  // MPCB_GUARDS=1: every grow-only scratch buffer sits between two 4 KiB canary regions (0xA5); mpcb_check_guards counts
  // the canary bytes that changed.  (compute-sanitizer is closed on the B200 pool: this is the bounds evidence.)
  int guards = 0;
  int poison = 0;  // MPCB_POISON=1: every workspace buffer is filled with 0xFF (NaN / -1) before a solve — a read of memory the
                   // solve did not write shows up as a changed result (tests: results must not depend on it)
  struct GuardedAlloc { void* user; void* base; size_t bytes; };
  std::vector<GuardedAlloc> guarded;
  int spec_w = 0;                                      // MPCB_SPECW env: fixed speculation window (0 = device policy)
  int spec_live_wide = 4096, spec_live_mid = 12288;    // MPCB_SPEC_LIVE="wide,mid": live-tile thresholds of the policy
  int split_ok = 1;                                    // MPCB_SPLIT env (0: never use the split evaluation)
  int pit_ok = 1;                                      // MPCB_PIT env (0: never use the parallel-in-time evaluation, 2: gradient requests only)
  // Round mode's loop: driven on the device by a CUDA graph (WHILE node; the default: mpcb_solve returns at once), or by
  // the host (MPCB_LOOP=host; also taken when a per-round diagnostic is on: MPCB_ROUND_LOG, MPCB_EVAL_TRACE)
  int loop_graph = 1;
  struct GraphRun { cudaGraphExec_t exec; cudaGraph_t graph; cudaEvent_t done; };
  std::vector<GraphRun> graph_runs;                    // kept alive until their completion events have fired
  bool graph_profile_pending = false; cudaStream_t graph_stream = nullptr; int64_t graph_ens = 0;
  unsigned long long* rs_stamps = nullptr;             // [kStatRounds][2] %globaltimer stamps of a graph-driven solve
  long long eval_trace_round = -1;                     // MPCB_EVAL_TRACE env
  int last_mode = 0;
  int64_t last_rounds = 0;
  double last_eval_ms = 0.0, last_trans_ms = 0.0;
  // bulk phase of the last round-mode solve (rounds with more requests than the evaluation launch has lanes)
  double bulk_eval_ms = 0.0, bulk_trans_ms = 0.0, bulk_grad = 0.0, bulk_val = 0.0; int64_t bulk_rounds = 0;
  unsigned* rs_stats = nullptr; std::vector<unsigned> rs_stats_host; std::vector<float> rs_t_ms, rs_e_ms;
  cudaEvent_t rs_ev[3 * 64] = {};                      // per-round timing events of one chunk
  // round mode (k_solve<..., true> + k_eval_requests): saved lane state, requests, lists, counters
  double* rs_d = nullptr; size_t rs_d_cap = 0;
  int32_t* rs_i = nullptr; size_t rs_i_cap = 0;
  void* rs_req = nullptr; size_t rs_req_cap = 0;
  double* rs_escr = nullptr; size_t rs_escr_cap = 0;
  int32_t* rs_keys = nullptr; size_t rs_keys_cap = 0;  // transition buckets: counts [2][NKEY] + lists [2][NKEY][nt]
  unsigned* rs_counters = nullptr;                     // [8] device
  unsigned* rs_counters_host = nullptr;                // [64][4] pinned: every round's request counters of one chunk
  const char* round_log = nullptr;                     // MPCB_ROUND_LOG=path: per-round [grad requests, psi requests, T ms, E ms]
  unsigned long long* queue = nullptr;
  double* zero_lam = nullptr; size_t zero_lam_cap = 0;
  double* gam = nullptr; size_t gam_cap = 0; int64_t gam_B = 0;  // final step sizes of the last solve
  unsigned long long* fin = nullptr; size_t fin_cap = 0;          // finish stamps of the last solve [B + 1]
  int32_t* ifail = nullptr; size_t ifail_cap = 0;                // inner convergence failures of the last solve
  double* trace = nullptr; size_t trace_buf_cap = 0; int trace_cap = 0; int64_t trace_B = 0;  // mpcb_set_trace
  // staging for *_host entry points
  void* pin = nullptr; size_t pin_cap = 0;
  void* dev = nullptr; size_t dev_cap = 0;
  cudaStream_t own_stream = nullptr;
  cudaEvent_t last_ev = nullptr;  // completion of the last call that used the shared scratch
  cudaStream_t last_stream = nullptr;
  bool has_last = false;
};

namespace {

int fail(MpcHandle* h, int code, const std::string& msg) {
  if (h) h->err = msg;
  return code;
}
int cuda_fail(MpcHandle* h, cudaError_t e, const char* where) {
  return fail(h, MPCB_E_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CK(call)                                                  \
  do {                                                            \
    cudaError_t e__ = (call);                                     \
    if (e__ != cudaSuccess) return cuda_fail(h, e__, #call);      \
  } while (0)

constexpr size_t kGuardBytes = 4096;

// cudaFree of a scratch buffer of the handle (its allocation starts kGuardBytes earlier when it is guarded).
void free_scratch(MpcHandle* h, void* user) {
  if (!user) return;
  for (size_t i = 0; i < h->guarded.size(); ++i)
    if (h->guarded[i].user == user) {
      cudaFree(h->guarded[i].base);
      h->guarded.erase(h->guarded.begin() + (long)i);
      return;
    }
  cudaFree(user);
}

template <class T>
int grow(MpcHandle* h, T** ptr, size_t* cap, size_t need_bytes, cudaStream_t st) {
  if (need_bytes <= *cap) return 0;
  if (*ptr) { CK(cudaStreamSynchronize(st)); free_scratch(h, *ptr); *ptr = nullptr; *cap = 0; }
  size_t sz = need_bytes + need_bytes / 8;
  if (!h->guards) {
    CK(cudaMalloc((void**)ptr, sz));
  } else {
    sz = (sz + 255) & ~size_t(255);
    char* base = nullptr;
    CK(cudaMalloc((void**)&base, sz + 2 * kGuardBytes));
    CK(cudaMemsetAsync(base, 0xA5, kGuardBytes, st));
    CK(cudaMemsetAsync(base + kGuardBytes + sz, 0xA5, kGuardBytes, st));
    *ptr = (T*)(base + kGuardBytes);
    h->guarded.push_back({(void*)*ptr, (void*)base, sz});
  }
  *cap = sz;
  return 0;
}

KDesc make_kdesc(const MpcProblemDesc& d) {
  KDesc k;
  std::memset(&k, 0, sizeof(k));
  k.model_id = d.model_id; k.N = d.N; k.S = d.S; k.n_sub = d.n_sub;
  k.nx = d.model_id == MPCB_MODEL_KINEMATIC4 ? 4 : 6;
  k.rows = k.nx + ((d.flags & MPCB_FLAG_SAFE_DIST) ? d.n_obs : 0);
  k.n = 2 * d.N; k.m = k.rows * d.N; k.n_obs = d.n_obs; k.flags = d.flags; k.road_mode = d.road_mode;
  k.n_roads = d.road_mode == MPCB_ROAD_TABLE ? d.n_roads : 1;
  k.h = d.Ts / d.n_sub; k.Ts = d.Ts; k.v_ref = d.v_ref;
  for (int i = 0; i < 6; ++i) { k.c[i] = d.c[i]; k.g_off[i] = d.g_offset[i]; k.D_lo[i] = d.D_lower[i]; k.D_hi[i] = d.D_upper[i]; }
  k.kin_acc = d.kin_accel; k.kin_mu = d.kin_friction;
  k.lane_a = d.lane_a; k.lane_y_t = d.lane_y_t; k.lane_b = d.lane_b; k.lane_y_br = d.lane_y_br; k.lane_y_bl = d.lane_y_bl;
  k.obs_Af = d.obs_Af; k.obs_b = d.obs_b; k.obs_sx = d.obs_sigma_x; k.obs_sy = d.obs_sigma_y; k.obs_adiv = d.obs_alpha_div;
  k.bounded = 0;
  k.debug = 0;
  k.exact = (d.flags & MPCB_FLAG_EXACT) ? 1 : 0;
  // psi = sum of c_i * (error)^2 + lane potentials + obstacle Gaussians + Sigma/2 dist^2 (mpcb_cost.cuh)
  k.psi_nonneg = 1;
  for (int i = 0; i < 6; ++i) if (!(d.c[i] >= 0.0)) k.psi_nonneg = 0;
  if ((d.flags & MPCB_FLAG_LANE) && !(d.lane_a >= 0.0 && d.lane_b >= 0.0)) k.psi_nonneg = 0;
  if ((d.flags & MPCB_FLAG_OBST) && !(d.obs_Af >= 0.0)) k.psi_nonneg = 0;
  k.safe_x0 = d.safe_x0; k.safe_y0 = d.safe_y0; k.safe_ax = d.safe_ax_max; k.safe_ay = d.safe_ay_max;
  if (d.flags & MPCB_FLAG_BOUNDED_D) {
    for (int i = 0; i < k.nx; ++i)
      if (!std::isinf(d.D_lower[i]) || !std::isinf(d.D_upper[i])) k.bounded = 1;
  } else {  // the state rows are live only with MPCB_FLAG_BOUNDED_D: an infinite box makes them vacuous
    const double inf = std::numeric_limits<double>::infinity();
    for (int i = 0; i < 6; ++i) { k.D_lo[i] = -inf; k.D_hi[i] = inf; }
  }
  if ((d.flags & MPCB_FLAG_SAFE_DIST) && d.n_obs > 0) k.bounded = 1;
  return k;
}

// Resolve the road argument for the evaluation kernels (stage in smem, or transpose to scratch).
int prepare_road(MpcHandle* h, int64_t B, const double* road, const int32_t* road_idx, cudaStream_t st,
                 RoadArg* ra, size_t* smem_bytes) {
  const int S = h->kd.S;
  ra->road_idx = nullptr; ra->ld = 0; ra->in_smem = 0; ra->R = 1; ra->road = road;
  *smem_bytes = 0;
  switch (h->kd.road_mode) {
    case MPCB_ROAD_SHARED:
      ra->in_smem = 1; ra->R = 1; *smem_bytes = (size_t)road_smem_doubles(S) * sizeof(double);
      return 0;
    case MPCB_ROAD_TABLE:
      if (!road_idx) return fail(h, MPCB_E_ARG, "road_idx is required for MPCB_ROAD_TABLE");
      ra->in_smem = 1; ra->R = h->kd.n_roads; ra->road_idx = road_idx;
      *smem_bytes = (size_t)ra->R * road_smem_doubles(S) * sizeof(double);
      if (*smem_bytes > 160 * 1024) return fail(h, MPCB_E_UNSUPPORTED, "road table exceeds 160 KB of shared memory");
      return 0;
    case MPCB_ROAD_PER_INSTANCE: {
      int rc = grow(h, &h->road_t, &h->road_t_cap, (size_t)B * 2 * S * sizeof(double), st);
      if (rc) return rc;
      dim3 grid((unsigned)((B + 31) / 32), (unsigned)((2 * S + 31) / 32)), block(32, 8);
      k_transpose_road<<<grid, block, 0, st>>>(B, 2 * S, road, h->road_t, B);
      h->launches++;
      ra->road = h->road_t; ra->ld = B;
      return 0;
    }
    case MPCB_ROAD_BEZIER: {
      int rc = grow(h, &h->road_t, &h->road_t_cap, (size_t)B * 2 * S * sizeof(double), st);
      if (rc) return rc;
      int64_t tot = B * S;
      k_bezier_sample<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(S, B, road, h->road_t, 1, B);
      h->launches++;
      ra->road = h->road_t; ra->ld = B;
      return 0;
    }
  }
  return fail(h, MPCB_E_ARG, "bad road_mode");
}

// The handle's scratch (workspace, transposed roads, work queue) is shared by every call on it.
// Calls may arrive on different streams (torch's current stream, the host entry point's own stream):
// order a call after the previous one whenever the stream changes.
int enter(MpcHandle* h, cudaStream_t st) {
  if (h->has_last && h->last_stream != st) CK(cudaStreamWaitEvent(st, h->last_ev, 0));
  return 0;
}
int leave(MpcHandle* h, cudaStream_t st) {
  CK(cudaGetLastError());
  CK(cudaEventRecord(h->last_ev, st));
  h->last_stream = st;
  h->has_last = true;
  return 0;
}

template <class K>
int set_smem(MpcHandle* h, K kernel, size_t bytes) {
  if (bytes > 48 * 1024) CK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

// Shared-memory carve-out of a kernel whose residency is bounded by registers, not by shared memory: ask for what
// `ctas_per_sm` resident CTAs use and leave the rest of the 256 KB to L1.  (Left to the driver, a kernel that requests a
// few KB of dynamic shared memory can get a carve-out sized for many more CTAs than its registers allow: the evaluation
// launch of the Bezier-road problems, whose vertex scans live in L1, ran 1.6x slower once it asked for 8 KB.)
template <class K>
int set_carveout(MpcHandle* h, K kernel, size_t dyn_bytes, int ctas_per_sm) {
  cudaFuncAttributes fa;
  CK(cudaFuncGetAttributes(&fa, kernel));
  const size_t need = (dyn_bytes + fa.sharedSizeBytes + 1024) * (size_t)ctas_per_sm;  // + 1 KB the system reserves per CTA
  int pct = (int)((need * 100 + 228 * 1024 - 1) / (228 * 1024));
  if (pct > 100) pct = 100;
  CK(cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct));
  return 0;
}

constexpr int64_t kStatRounds = 16384;  // rounds whose request counts / time stamps are kept for the solve profile

// Release graph executables whose solves have completed (all of them when `all`).
void reap_graph_runs(MpcHandle* h, bool all) {
  for (size_t i = 0; i < h->graph_runs.size();) {
    auto& gr = h->graph_runs[i];
    if (all) cudaEventSynchronize(gr.done);
    if (all || cudaEventQuery(gr.done) == cudaSuccess) {
      cudaGraphExecDestroy(gr.exec); cudaGraphDestroy(gr.graph); cudaEventDestroy(gr.done);
      h->graph_runs.erase(h->graph_runs.begin() + (long)i);
    } else {
      ++i;
    }
  }
}

// MPCB_LOOP=graph: one graph launch per solve.  WHILE(more) { IF(small) T<small grid> ELSE T<full grid>; E; advance }.
int launch_graph_loop(MpcHandle* h, cudaStream_t st, void* t_func, void* e_func, KDesc& k, MpcSolverParams& sp, SolveArgs& a,
                      EvalArgs& e, int64_t t_blocks_full, int block, size_t box_smem, int64_t eblocks, size_t eval_smem, int G,
                      unsigned round_cap) {
  reap_graph_runs(h, false);
  if (h->graph_runs.size() >= 8) reap_graph_runs(h, true);
  if (!h->rs_stamps) CK(cudaMalloc((void**)&h->rs_stamps, kStatRounds * 2 * sizeof(unsigned long long)));
  unsigned* ctl = h->rs_counters + 8;
  a.round = 0; a.round_ptr = ctl + 1;
  e.round_ptr = ctl + 1; e.counters_base = h->rs_counters; e.key_cnt_base = a.key_cnt; e.round_stats_base = h->rs_stats;
  e.stamps = h->rs_stamps; e.stat_rounds = (int)kStatRounds; e.trace = nullptr;
  cudaGraph_t g = nullptr;
  CK(cudaGraphCreate(&g, 0));
  cudaGraphConditionalHandle more_h, small_h;
  CK(cudaGraphConditionalHandleCreate(&more_h, g, 1, cudaGraphCondAssignDefault));
  CK(cudaGraphConditionalHandleCreate(&small_h, g, 0, cudaGraphCondAssignDefault));
  cudaGraphNodeParams wp = {};
  wp.type = cudaGraphNodeTypeConditional;
  wp.conditional.handle = more_h; wp.conditional.type = cudaGraphCondTypeWhile; wp.conditional.size = 1;
  cudaGraphNode_t wnode;
  CK(cudaGraphAddNode(&wnode, g, nullptr, 0, &wp));
  cudaGraph_t body = wp.conditional.phGraph_out[0];
  cudaGraphNodeParams ip = {};
  ip.type = cudaGraphNodeTypeConditional;
  ip.conditional.handle = small_h; ip.conditional.type = cudaGraphCondTypeIf; ip.conditional.size = 2;
  cudaGraphNode_t inode;
  CK(cudaGraphAddNode(&inode, body, nullptr, 0, &ip));
  cudaGraph_t g_small = ip.conditional.phGraph_out[0], g_full = ip.conditional.phGraph_out[1];
  // the transition launch: once at most `small_tiles` tiles are alive a grid for that many (plus bucket padding)
  const unsigned small_tiles = 1024;
  int64_t t_blocks_small = (((int64_t)small_tiles + (int64_t)MPCB_NKEY * (32 / G)) * G + block - 1) / block;
  if (t_blocks_small > t_blocks_full) t_blocks_small = t_blocks_full;
  void* t_args[] = {&k, &sp, &a};
  cudaKernelNodeParams kp = {};
  kp.func = t_func; kp.blockDim = dim3((unsigned)block); kp.sharedMemBytes = (unsigned)box_smem; kp.kernelParams = t_args;
  cudaGraphNode_t tn;
  kp.gridDim = dim3((unsigned)t_blocks_small);
  CK(cudaGraphAddKernelNode(&tn, g_small, nullptr, 0, &kp));
  kp.gridDim = dim3((unsigned)t_blocks_full);
  CK(cudaGraphAddKernelNode(&tn, g_full, nullptr, 0, &kp));
  void* e_args[] = {&k, &e};
  cudaKernelNodeParams ke = {};
  ke.func = e_func; ke.gridDim = dim3((unsigned)eblocks); ke.blockDim = dim3(256); ke.sharedMemBytes = (unsigned)eval_smem;
  ke.kernelParams = e_args;
  cudaGraphNode_t en;
  CK(cudaGraphAddKernelNode(&en, body, &inode, 1, &ke));
  unsigned* counters_base = h->rs_counters;
  unsigned long long* stamps = h->rs_stamps;
  int stat_rounds = (int)kStatRounds;
  unsigned st_small = small_tiles, cap = round_cap;
  void* a_args[] = {&ctl, &counters_base, &more_h, &small_h, &st_small, &cap, &stamps, &stat_rounds};
  cudaKernelNodeParams ka = {};
  ka.func = (void*)k_round_advance; ka.gridDim = dim3(1); ka.blockDim = dim3(32); ka.kernelParams = a_args;
  cudaGraphNode_t an;
  CK(cudaGraphAddKernelNode(&an, body, &en, 1, &ka));
  cudaGraphExec_t exec = nullptr;
  CK(cudaGraphInstantiate(&exec, g, 0));
  CK(cudaGraphLaunch(exec, st));
  cudaEvent_t done;
  CK(cudaEventCreateWithFlags(&done, cudaEventDisableTiming));
  CK(cudaEventRecord(done, st));
  h->graph_runs.push_back({exec, g, done});
  h->graph_profile_pending = true; h->graph_stream = st; h->graph_ens = e.ens;
  return 0;
}

// The profile of a graph-driven solve is read from the device once somebody asks for it (synchronises the solve's stream).
void fetch_graph_profile(MpcHandle* h) {
  if (!h->graph_profile_pending) return;
  h->graph_profile_pending = false;
  if (cudaStreamSynchronize(h->graph_stream) != cudaSuccess) return;
  unsigned ctl[4] = {0, 0, 0, 0};
  if (cudaMemcpy(ctl, h->rs_counters + 8, sizeof(ctl), cudaMemcpyDeviceToHost) != cudaSuccess) return;
  const int64_t rounds = ctl[1], nr = std::min<int64_t>(rounds, kStatRounds);
  std::vector<unsigned long long> stamps((size_t)nr * 2);
  h->rs_stats_host.resize((size_t)nr * 2);
  if (nr > 0) {
    if (cudaMemcpy(stamps.data(), h->rs_stamps, stamps.size() * sizeof(unsigned long long), cudaMemcpyDeviceToHost) != cudaSuccess) return;
    if (cudaMemcpy(h->rs_stats_host.data(), h->rs_stats, (size_t)nr * 2 * sizeof(unsigned), cudaMemcpyDeviceToHost) != cudaSuccess) return;
  }
  h->last_rounds = rounds; h->launches += 3 * rounds;
  h->last_eval_ms = h->last_trans_ms = 0.0;
  h->bulk_eval_ms = h->bulk_trans_ms = h->bulk_grad = h->bulk_val = 0.0; h->bulk_rounds = 0;
  for (int64_t r = 0; r < nr; ++r) {
    const double e_ms = (double)(stamps[2 * r + 1] - stamps[2 * r]) * 1e-6;            // evaluation launch (+ the advance node)
    const double t_ms = r > 0 ? (double)(stamps[2 * r] - stamps[2 * r - 1]) * 1e-6 : 0.0;  // transition launch (+ node scheduling)
    h->last_eval_ms += e_ms; h->last_trans_ms += t_ms;
    const double g = h->rs_stats_host[2 * r], v = h->rs_stats_host[2 * r + 1];
    if (g + v > (double)h->graph_ens) { h->bulk_eval_ms += e_ms; h->bulk_trans_ms += t_ms; h->bulk_grad += g; h->bulk_val += v; h->bulk_rounds++; }
  }
}

}  // namespace

extern "C" {

int mpcb_version(void) { return MPCB_VERSION; }

void mpcb_default_desc(MpcProblemDesc* d, int32_t model_id, int32_t N, int32_t S) {
  std::memset(d, 0, sizeof(*d));
  d->model_id = model_id; d->N = N; d->S = S; d->n_sub = 4;   // car_dynamics.py:136
  d->road_mode = MPCB_ROAD_SHARED; d->n_roads = 1; d->flags = 0; d->n_obs = 0;
  d->Ts = 0.05;                                                // car_dynamics.py:93
  d->v_ref = 1.0;                                              // main.py:65
  const double c[6] = {0.5, 1, 1, 0.5, 0.1, 0.01};             // car_dynamics.py:230
  for (int i = 0; i < 6; ++i) d->c[i] = c[i];
  d->kin_friction = 1.0; d->kin_accel = 2.0;                   // dynamics.py:34-35
  d->lane_a = 0.5; d->lane_y_t = 1.75; d->lane_b = 100.0; d->lane_y_br = 1.0; d->lane_y_bl = 6.0;  // dpf_test.py:5,45-54
  d->obs_Af = 1000.0; d->obs_b = 1.0; d->obs_sigma_x = 2.0; d->obs_sigma_y = 0.5; d->obs_alpha_div = 5.0;  // dpf_test.py:81,91
  const double off[6] = {20, 1, 1, 2, 1, 0.1};                 // main.py:46-51
  const double inf = std::numeric_limits<double>::infinity();
  for (int i = 0; i < 6; ++i) { d->g_offset[i] = off[i]; d->D_lower[i] = -inf; d->D_upper[i] = inf; }  // main.py:57
  d->safe_x0 = 5.0; d->safe_y0 = 3.0; d->safe_ax_max = 3.0; d->safe_ay_max = 1.0;                      // dpf_test.py:6
}

void mpcb_default_solver_params(MpcSolverParams* sp, int32_t N) {
  std::memset(sp, 0, sizeof(*sp));
  sp->max_iter = 1000; sp->lbfgs_mem = N; sp->max_no_progress = 10; sp->hess_vec_fd = 1;  // controller.py:31,36
  sp->Lgamma_factor = 0.95; sp->L_min = 1e-5; sp->L_max = 1e20; sp->lip_eps = 1e-6; sp->lip_delta = 1e-12;
  sp->tau_min = 1.0 / 256; sp->quad_tol = 10 * 2.220446049250313e-16; sp->ls_tol = 10 * 2.220446049250313e-16;
  sp->alm_max_iter = 1000; sp->eps_target = 1e-6; sp->delta_target = 1e-4; sp->Sigma0 = 1e5;  // controller.py:41-45
  sp->eps0 = 1.0; sp->rho = 0.1; sp->theta = 0.1; sp->Delta = 10.0; sp->M = 1e9; sp->Sigma_max = 1e9;
}

void mpcb_default_vehicle_params(double* p) {  // main.py:82-111
  const double v[MPCB_NP] = {9.7e-2, 4.7e-2, 5e-2, 0.09, 0.07, 8e-2, 5.5e-2, 0.1735, 18.3e-5, 0.32, 1.0,
                             0.268, 2.165, 3.47, 0.242, 2.38, 2.84, 0.266, 0.1, 0.1025, 0.1629, 0.0011};
  for (int i = 0; i < MPCB_NP; ++i) p[i] = v[i];
}

int mpcb_create(const MpcProblemDesc* desc, MpcHandle** out) {
  if (!desc || !out) return MPCB_E_ARG;
  *out = nullptr;
  if (desc->N < 1 || desc->N > MPCB_MAX_N || desc->S < 3 || desc->n_sub < 1 || desc->n_obs < 0 ||
      desc->n_obs > MPCB_MAX_OBS || desc->model_id < 0 || desc->model_id > 1 || desc->road_mode < 0 ||
      desc->road_mode > 3)
    return MPCB_E_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MPCB_E_NODEVICE;
  MpcHandle* h = new MpcHandle();
  h->desc = *desc;
  h->kd = make_kdesc(*desc);
  cudaGetDevice(&h->device);
  cudaDeviceGetAttribute(&h->num_sms, cudaDevAttrMultiProcessorCount, h->device);
  if (cudaMalloc((void**)&h->queue, sizeof(unsigned long long)) != cudaSuccess) { delete h; return MPCB_E_CUDA; }
  if (cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking) != cudaSuccess) { cudaFree(h->queue); delete h; return MPCB_E_CUDA; }
  if (cudaEventCreateWithFlags(&h->last_ev, cudaEventDisableTiming) != cudaSuccess) { cudaFree(h->queue); delete h; return MPCB_E_CUDA; }
  if (const char* e = std::getenv("MPCB_DEBUG_TIMERS")) h->debug_timers = std::atoi(e);
  if (const char* e = std::getenv("MPCB_WARPS_PER_SM")) { int w = std::atoi(e); if (w >= 1 && w <= 8) h->warps_per_sm = w; }
  h->round_log = std::getenv("MPCB_ROUND_LOG");
  if (const char* e = std::getenv("MPCB_MODE")) {
    if (!std::strcmp(e, "rounds") || !std::strcmp(e, "1")) h->mode_override = 1;
    else if (!std::strcmp(e, "persistent") || !std::strcmp(e, "0")) h->mode_override = 0;
  }
  if (const char* e = std::getenv("MPCB_TILE")) {
    int g = std::atoi(e);
    if (g == 1 || g == 8) h->tile_override = g;
  }
  if (const char* e = std::getenv("MPCB_GUARDS")) h->guards = std::atoi(e) != 0;
  if (const char* e = std::getenv("MPCB_POISON")) h->poison = std::atoi(e) != 0;
  if (const char* e = std::getenv("MPCB_EXACT")) { if (std::atoi(e) != 0) h->kd.exact = 1; }
  if (const char* e = std::getenv("MPCB_SPECW")) { int w = std::atoi(e); if (w >= 1 && w <= 7) h->spec_w = w; }
  if (const char* e = std::getenv("MPCB_SPEC_LIVE")) {
    int a = 0, b = 0;
    if (std::sscanf(e, "%d,%d", &a, &b) == 2 && a >= 0 && b >= a) { h->spec_live_wide = a; h->spec_live_mid = b; }
  }
  if (const char* e = std::getenv("MPCB_SPLIT")) h->split_ok = std::atoi(e) != 0;
  if (const char* e = std::getenv("MPCB_PIT")) h->pit_ok = std::atoi(e);
  if (const char* e = std::getenv("MPCB_LOOP")) h->loop_graph = std::strcmp(e, "host") != 0;
  if (const char* e = std::getenv("MPCB_EVAL_TRACE")) h->eval_trace_round = std::atoll(e);
  *out = h;
  return MPCB_OK;
}

void mpcb_destroy(MpcHandle* h) {
  if (!h) return;
  for (void* q : {(void*)h->road_t, (void*)h->ws, (void*)h->wst, (void*)h->dbg, (void*)h->zero_lam, (void*)h->gam, (void*)h->fin,
                  (void*)h->ifail, (void*)h->trace, (void*)h->dev, (void*)h->rs_d, (void*)h->rs_i, (void*)h->rs_req,
                  (void*)h->rs_escr, (void*)h->rs_keys})
    free_scratch(h, q);
  cudaFree(h->queue); cudaFree(h->rs_stats); cudaFree(h->rs_counters); cudaFree(h->rs_stamps);
  for (auto& gr : h->graph_runs) { cudaEventSynchronize(gr.done); cudaGraphExecDestroy(gr.exec); cudaGraphDestroy(gr.graph); cudaEventDestroy(gr.done); }
  if (h->rs_counters_host) cudaFreeHost(h->rs_counters_host);
  for (cudaEvent_t ev : h->rs_ev) if (ev) cudaEventDestroy(ev);
  if (h->pin) cudaFreeHost(h->pin);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  if (h->last_ev) cudaEventDestroy(h->last_ev);
  delete h;
}

const char* mpcb_last_error(MpcHandle* h) { return h ? h->err.c_str() : "null handle"; }
int64_t mpcb_launch_count(MpcHandle* h) { if (h) fetch_graph_profile(h); return h ? h->launches : 0; }
void mpcb_last_solve_profile(MpcHandle* h, double* out) {
  if (!h || !out) return;
  fetch_graph_profile(h);
  out[0] = h->last_mode; out[1] = h->last_tile; out[2] = (double)h->last_rounds;
  out[3] = h->last_eval_ms; out[4] = h->last_trans_ms; out[5] = h->last_mode ? (double)h->last_rounds : 1.0;
}
void mpcb_last_solve_bulk_profile(MpcHandle* h, double* out) {
  if (!h || !out) return;
  fetch_graph_profile(h);
  out[0] = (double)h->bulk_rounds; out[1] = h->bulk_eval_ms; out[2] = h->bulk_trans_ms; out[3] = h->bulk_grad; out[4] = h->bulk_val;
}

int mpcb_dynamics(MpcHandle* h, int64_t B, const double* x, const double* u, const double* p, int32_t clip,
                  double* out, void* stream) {
  if (!h || !x || !u || !p || !out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_dynamics: bad argument");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned grid = (unsigned)((B + 127) / 128);
  if (h->kd.model_id == MPCB_MODEL_PACEJKA6) k_dynamics<MPCB_MODEL_PACEJKA6><<<grid, 128, 0, st>>>(h->kd, B, x, u, p, clip, out);
  else k_dynamics<MPCB_MODEL_KINEMATIC4><<<grid, 128, 0, st>>>(h->kd, B, x, u, p, clip, out);
  h->launches++;
  CK(cudaGetLastError());
  return MPCB_OK;
}

int mpcb_plant_step(MpcHandle* h, int64_t B, const double* x, const double* u, const double* p, double* out,
                    void* stream) {
  if (!h || !x || !u || !p || !out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_plant_step: bad argument");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned grid = (unsigned)((B + 127) / 128);
  if (h->kd.model_id == MPCB_MODEL_PACEJKA6) k_plant_step<MPCB_MODEL_PACEJKA6><<<grid, 128, 0, st>>>(h->kd, B, x, u, p, out);
  else k_plant_step<MPCB_MODEL_KINEMATIC4><<<grid, 128, 0, st>>>(h->kd, B, x, u, p, out);
  h->launches++;
  CK(cudaGetLastError());
  return MPCB_OK;
}

int mpcb_road_errors(MpcHandle* h, int64_t B, const double* pos, const double* heading, const double* road,
                     const int32_t* road_idx, int32_t normalised, int32_t* idx_out, double* err_out, void* stream) {
  if (!h || !pos || !heading || !road || !idx_out || !err_out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_road_errors: bad argument");
  if (h->kd.road_mode == MPCB_ROAD_BEZIER) return fail(h, MPCB_E_UNSUPPORTED, "sample the Bezier road first (mpcb_bezier_sample)");
  if (h->kd.road_mode == MPCB_ROAD_TABLE && !road_idx) return fail(h, MPCB_E_ARG, "road_idx required");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  k_road_errors<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(h->kd, B, pos, heading, road, road_idx, normalised, idx_out, err_out);
  h->launches++;
  CK(cudaGetLastError());
  return MPCB_OK;
}

int mpcb_bezier_sample(MpcHandle* h, int64_t B, const double* ctrl, double* road_out, void* stream) {
  if (!h || !ctrl || !road_out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_bezier_sample: bad argument");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int64_t tot = B * h->kd.S;
  k_bezier_sample<<<(unsigned)((tot + 255) / 256), 256, 0, st>>>(h->kd.S, B, ctrl, road_out, 0, 0);
  h->launches++;
  CK(cudaGetLastError());
  return MPCB_OK;
}

int mpcb_rollout(MpcHandle* h, int64_t B, const double* x0, const double* U, const double* p, const double* road,
                 const int32_t* road_idx, const double* obs, double* states_out, int32_t* idx_out, double* cost_out,
                 void* stream) {
  if (!h || B < 0) return fail(h, MPCB_E_ARG, "mpcb_rollout: bad argument");
  if (B == 0) return MPCB_OK;
  if (!x0 || !U || !p || !road) return fail(h, MPCB_E_ARG, "mpcb_rollout: null argument");
  if ((h->kd.flags & (MPCB_FLAG_OBST | MPCB_FLAG_SAFE_DIST)) && h->kd.n_obs > 0 && !obs) return fail(h, MPCB_E_ARG, "obs required with MPCB_FLAG_OBST / MPCB_FLAG_SAFE_DIST");
  cudaStream_t st = (cudaStream_t)stream;
  RoadArg ra; size_t smem;
  int rc;
  if ((rc = enter(h, st))) return rc;
  if ((rc = prepare_road(h, B, road, road_idx, st, &ra, &smem))) return rc;
  unsigned grid = (unsigned)((B + 127) / 128);
  if (h->kd.model_id == MPCB_MODEL_PACEJKA6) {
    if ((rc = set_smem(h, k_rollout<MPCB_MODEL_PACEJKA6>, smem))) return rc;
    k_rollout<MPCB_MODEL_PACEJKA6><<<grid, 128, smem, st>>>(h->kd, B, x0, U, p, ra, obs, states_out, idx_out, cost_out);
  } else {
    if ((rc = set_smem(h, k_rollout<MPCB_MODEL_KINEMATIC4>, smem))) return rc;
    k_rollout<MPCB_MODEL_KINEMATIC4><<<grid, 128, smem, st>>>(h->kd, B, x0, U, p, ra, obs, states_out, idx_out, cost_out);
  }
  h->launches++;
  return leave(h, st);
}

int mpcb_value_grad(MpcHandle* h, int64_t B, const double* x0, const double* U, const double* p, const double* road,
                    const int32_t* road_idx, const double* obs, const double* lambda, const double* Sigma,
                    double* cost_out, double* grad_out, void* stream) {
  if (!h || B < 0) return fail(h, MPCB_E_ARG, "mpcb_value_grad: bad argument");
  if (B == 0) return MPCB_OK;
  if (!x0 || !U || !p || !road || !cost_out || !grad_out) return fail(h, MPCB_E_ARG, "mpcb_value_grad: null argument");
  if ((h->kd.flags & (MPCB_FLAG_OBST | MPCB_FLAG_SAFE_DIST)) && h->kd.n_obs > 0 && !obs) return fail(h, MPCB_E_ARG, "obs required with MPCB_FLAG_OBST / MPCB_FLAG_SAFE_DIST");
  cudaStream_t st = (cudaStream_t)stream;
  RoadArg ra; size_t smem;
  int rc;
  if ((rc = enter(h, st))) return rc;
  if ((rc = prepare_road(h, B, road, road_idx, st, &ra, &smem))) return rc;
  // persistent grid: at most 8 CTAs of 128 lanes per SM; scratch is [rows, ns] lane-interleaved
  const int block = 128;
  int64_t max_blocks = (int64_t)h->num_sms * 8;
  int64_t blocks = (B + block - 1) / block;
  if (blocks > max_blocks) blocks = max_blocks;
  int64_t ns = blocks * block;
  const KDesc& k = h->kd;
  size_t rows = (size_t)k.N * (k.model_id == MPCB_MODEL_PACEJKA6 ? fm_stage_doubles<MPCB_MODEL_PACEJKA6>()
                                                                   : fm_stage_doubles<MPCB_MODEL_KINEMATIC4>());
  if ((rc = grow(h, &h->ws, &h->ws_cap, rows * ns * sizeof(double), st))) return rc;
  if (k.model_id == MPCB_MODEL_PACEJKA6) {
    if ((rc = set_smem(h, k_value_grad<MPCB_MODEL_PACEJKA6>, smem))) return rc;
    k_value_grad<MPCB_MODEL_PACEJKA6><<<(unsigned)blocks, block, smem, st>>>(k, B, x0, U, p, ra, obs, lambda, Sigma, cost_out, grad_out, h->ws, ns);
  } else {
    if ((rc = set_smem(h, k_value_grad<MPCB_MODEL_KINEMATIC4>, smem))) return rc;
    k_value_grad<MPCB_MODEL_KINEMATIC4><<<(unsigned)blocks, block, smem, st>>>(k, B, x0, U, p, ra, obs, lambda, Sigma, cost_out, grad_out, h->ws, ns);
  }
  h->launches++;
  return leave(h, st);
}

int mpcb_solve(MpcHandle* h, const MpcSolverParams* sp, int64_t B, const double* x0, const double* p,
               const double* road, const int32_t* road_idx, const double* obs, const double* lb, const double* ub,
               double* U, double* lambda, int32_t* status, int32_t* iters, int32_t* evals, double* psi, double* res,
               void* stream) {
  if (!h || !sp || B < 0) return fail(h, MPCB_E_ARG, "mpcb_solve: bad argument");
  if (B == 0) return MPCB_OK;
  if (!x0 || !p || !road || !lb || !ub || !U || !status || !iters || !evals || !psi || !res)
    return fail(h, MPCB_E_ARG, "mpcb_solve: null argument");
  if (sp->lbfgs_mem < 1 || sp->lbfgs_mem > 2 * MPCB_MAX_N || sp->max_iter < 0 || sp->max_no_progress < 1 || sp->alm_max_iter < 1)
    return fail(h, MPCB_E_ARG, "mpcb_solve: bad solver params (lbfgs_mem in [1, 2 MPCB_MAX_N], max_iter >= 0, max_no_progress >= 1, alm_max_iter >= 1)");
  if ((h->kd.flags & (MPCB_FLAG_OBST | MPCB_FLAG_SAFE_DIST)) && h->kd.n_obs > 0 && !obs) return fail(h, MPCB_E_ARG, "obs required with MPCB_FLAG_OBST / MPCB_FLAG_SAFE_DIST");
  cudaStream_t st = (cudaStream_t)stream;
  const KDesc& k = h->kd;
  int rc;
  if ((rc = enter(h, st))) return rc;
  if (k.bounded && !lambda) {  // λ = 0 warm start (controller.py:21) when the caller keeps no multipliers
    if ((rc = grow(h, &h->zero_lam, &h->zero_lam_cap, (size_t)B * k.m * sizeof(double), st))) return rc;
    CK(cudaMemsetAsync(h->zero_lam, 0, (size_t)B * k.m * sizeof(double), st));
    lambda = h->zero_lam;
  }
  RoadArg ra; size_t smem;
  if ((rc = prepare_road(h, B, road, road_idx, st, &ra, &smem))) return rc;
  const size_t road_smem = smem;
  // Two execution modes (same state machine, same iterates):
  //  * persistent (k_solve<.., false>): one launch, tiles pull instances from a queue and evaluate in place;
  //  * rounds (k_solve<.., true> + k_eval_requests): every round is a transition launch that publishes each
  //    lane's evaluation request and an evaluation launch over the compacted request lists.  The host
  //    enqueues rounds ahead of the GPU and reads the request counters every few dozen rounds to stop.
  // Tile width: spare lanes are spent on speculative line-search candidates (mpcb_solve.cuh).
  int block = 64;
  // Round mode at every batch size (measured on a B200 with the bucketed transition launches, gpurun r02s11: B = 1
  // 55.8 vs 73.7 ms, B = 64 164 vs 207 ms, B = 512 186 vs 205 ms, B = 1023 323 vs 326 ms, B = 65536 1.69 vs 6.9 s);
  // MPCB_MODE=persistent selects the one-launch persistent kernel.
  const int mode = h->mode_override >= 0 ? h->mode_override : 1;
  const int64_t lane_cap = (int64_t)h->num_sms * h->warps_per_sm * 32;
  int G = 1;
  if (mode == 1) {
    // rounds: idle lanes cost nothing in the evaluation launch; speculation pays until the evaluation launch
    // is throughput-bound (about 4 requests per evaluating lane in flight)
    // measured (profiles/r01_round_mode.md): G = 8 wins at every batch size up to 65536 (fewest rounds; idle lanes are
    // free); narrower tiles only to keep the per-lane workspace of very large batches within ~8 GB
    // (per-lane rows + the per-tile L-BFGS history; tile widths are 8 or 1)
    G = 8;
    const double per_tile = 8.0 * ((double)G * (8.0 * k.n + (k.bounded ? k.N * k.nx : 0)) + 2.0 * sp->lbfgs_mem * k.n);
    if ((double)B * per_tile > 24e9) G = 1;
  } else {
    if (B * 8 <= (int64_t)h->num_sms * 8 * 32) G = 8;
  }
  if (h->tile_override > 0) G = h->tile_override;
  if (mode == 1) block = MPCB_ROUND_BLOCK;
  // + box [2n] + the two-loop recursion scalars of the CTA's tiles [block / G][2 mem]
  const size_t box_smem = ((size_t)2 * k.n + (size_t)(block / G) * 2 * sp->lbfgs_mem) * sizeof(double);
  smem += box_smem;
  if (smem > 200 * 1024) return fail(h, MPCB_E_UNSUPPORTED, "mpcb_solve: road table + L-BFGS scalars exceed shared memory");
  int64_t tiles_wanted = B;
  int64_t blocks = (tiles_wanted * G + block - 1) / block;
  const int64_t max_blocks = lane_cap / block;
  if (mode == 0 && blocks > max_blocks) blocks = max_blocks;
  const int64_t ns = blocks * block, nt = ns / G;
  // round mode: the transition grid covers every tile plus the padding of the buckets to whole warps
  if (mode == 1) blocks = ((nt + (int64_t)MPCB_NKEY * (32 / G)) * G + block - 1) / block;
  const bool bounded = k.bounded != 0;
  const int stage_doubles = k.model_id == MPCB_MODEL_PACEJKA6 ? fm_stage_doubles<MPCB_MODEL_PACEJKA6>()
                                                              : fm_stage_doubles<MPCB_MODEL_KINEMATIC4>();
  LaneMap LM(k.n, k.N, k.nx, mode == 1 ? 0 : stage_doubles, bounded);
  TileMap TM(k.n, G, sp->lbfgs_mem, k.m, bounded);
  if ((rc = grow(h, &h->ws, &h->ws_cap, (size_t)LM.total * ns * sizeof(double), st))) return rc;
  if ((rc = grow(h, &h->wst, &h->wst_cap, (size_t)TM.total * nt * sizeof(double), st))) return rc;
  CK(cudaMemsetAsync(h->queue, 0, sizeof(unsigned long long), st));
  if (h->poison) {
    CK(cudaMemsetAsync(h->ws, 0xFF, (size_t)LM.total * ns * sizeof(double), st));
    CK(cudaMemsetAsync(h->wst, 0xFF, (size_t)TM.total * nt * sizeof(double), st));
  }
  if ((rc = grow(h, &h->gam, &h->gam_cap, (size_t)B * sizeof(double), st))) return rc;
  h->gam_B = B;
  if ((rc = grow(h, &h->ifail, &h->ifail_cap, (size_t)B * sizeof(int32_t), st))) return rc;
  if ((rc = grow(h, &h->fin, &h->fin_cap, (size_t)(B + 1) * sizeof(unsigned long long), st))) return rc;
  CK(cudaMemsetAsync(h->fin, 0xFF, (size_t)(B + 1) * sizeof(unsigned long long), st));
  h->trace_B = 0;
  if (h->trace_cap > 0) {
    const size_t tb = (size_t)B * h->trace_cap * 4 * sizeof(double);
    if ((rc = grow(h, &h->trace, &h->trace_buf_cap, tb, st))) return rc;
    CK(cudaMemsetAsync(h->trace, 0xFF, tb, st));  // NaN = iteration not reached
    h->trace_B = B;
  }
  SolveArgs a;
  std::memset(&a, 0, sizeof(a));
  a.B = B; a.x0 = x0; a.p = p; a.obs = obs; a.lb = lb; a.ub = ub; a.ra = ra; a.U = U; a.lam = lambda;
  a.status = status; a.iters = iters; a.evals = evals; a.psi = psi; a.res = res; a.gam_out = h->gam; a.ifail_out = h->ifail; a.fin_ns = h->fin;
  a.trace = h->trace_cap > 0 ? h->trace : nullptr; a.trace_cap = h->trace_cap;
  a.ws = h->ws; a.wst = h->wst; a.ns = ns; a.nt = nt; a.queue = h->queue;
  a.dbg = nullptr;
  a.spec_w = h->spec_w; a.spec_live_wide = h->spec_live_wide; a.spec_live_mid = h->spec_live_mid; a.ctl = nullptr;
  h->kd.debug = h->debug_timers;
  h->last_tile = G;
  h->last_mode = mode;
  if (mode == 1) {
    // evaluation launch: one CTA of 256 lanes (255 registers each) per SM, never more lanes than requests can exist
    int64_t wblocks = (ns + 31) / 32;   // CTAs with evaluation scratch: one warp per SM sub-partition before any doubling up
    if (wblocks > (int64_t)h->num_sms) wblocks = (int64_t)h->num_sms;
    const int64_t ens = wblocks * 256;
    // the grid always covers every SM: CTAs beyond `wblocks` only take parallel-in-time requests (shared memory only)
    const int64_t eblocks = h->num_sms;
    const size_t pit_smem = sizeof(double) * (size_t)fm_pit_doubles<MPCB_MODEL_PACEJKA6>(k.N);  // (Fm<KINEMATIC4>::PIT is false)
    const size_t eval_smem = road_smem + (h->pit_ok && k.model_id == MPCB_MODEL_PACEJKA6 ? pit_smem : 0);
    if (eval_smem > 200 * 1024) return fail(h, MPCB_E_UNSUPPORTED, "mpcb_solve: road table + evaluation buffers exceed shared memory");
    if ((rc = grow(h, &h->rs_d, &h->rs_d_cap, (size_t)MPCB_SD_ROWS * ns * sizeof(double), st))) return rc;
    if ((rc = grow(h, &h->rs_i, &h->rs_i_cap, (size_t)MPCB_SI_ROWS * ns * sizeof(int32_t), st))) return rc;
    // requests: want[ns] i32, list_grad[ns] i32, list_val[ns] i32, (pad) b[ns] i64, p[ns], g[ns] ptr, val[ns] f64
    const size_t ns8 = ((size_t)ns + 1) & ~size_t(1);
    if ((rc = grow(h, (char**)&h->rs_req, &h->rs_req_cap, ns8 * (3 * 4 + 4 * 8) + 64, st))) return rc;
    if ((rc = grow(h, &h->rs_escr, &h->rs_escr_cap, (size_t)k.N * stage_doubles * ens * sizeof(double), st))) return rc;
    if (!h->rs_counters) {
      CK(cudaMalloc((void**)&h->rs_counters, 16 * sizeof(unsigned)));  // [2][4] round counters + [4] control words
      CK(cudaMallocHost((void**)&h->rs_counters_host, 64 * 4 * sizeof(unsigned)));
    }
    CK(cudaMemsetAsync(h->rs_counters, 0, 16 * sizeof(unsigned), st));
    if (h->poison) {
      CK(cudaMemsetAsync(h->rs_d, 0xFF, (size_t)MPCB_SD_ROWS * ns * sizeof(double), st));
      CK(cudaMemsetAsync(h->rs_i, 0xFF, (size_t)MPCB_SI_ROWS * ns * sizeof(int32_t), st));
      CK(cudaMemsetAsync(h->rs_escr, 0xFF, (size_t)k.N * stage_doubles * ens * sizeof(double), st));
    }
    const size_t key_ints = (size_t)2 * MPCB_NKEY * (1 + (size_t)nt);
    if ((rc = grow(h, &h->rs_keys, &h->rs_keys_cap, key_ints * sizeof(int32_t), st))) return rc;
    CK(cudaMemsetAsync(h->rs_keys, 0, (size_t)2 * MPCB_NKEY * sizeof(int32_t), st));
    a.key_cnt = h->rs_keys; a.key_list = h->rs_keys + 2 * MPCB_NKEY;
    char* rq = (char*)h->rs_req;
    a.req_b = (int64_t*)rq;
    a.req_p = (double**)(rq + ns8 * 8);
    a.req_g = (double**)(rq + ns8 * 16);
    a.req_val = (double*)(rq + ns8 * 24);
    a.req_want = (int32_t*)(rq + ns8 * 32);
    a.list_grad = (int32_t*)(rq + ns8 * 36);
    a.list_val = (int32_t*)(rq + ns8 * 40);
    a.sd = h->rs_d; a.si = h->rs_i; a.counters = h->rs_counters; a.ctl = h->rs_counters + 8;
    EvalArgs e;
    std::memset(&e, 0, sizeof(e));
    e.ns = ns; e.nt = nt; e.ens = ens; e.G = G;
    e.x0 = x0; e.p = p; e.obs = obs; e.lam = lambda; e.ra = ra;
    e.tp_base = h->wst + (size_t)TM.vec_total * nt; e.tp_rows = TM.total - TM.vec_total; e.sig_row = TM.SIG - TM.vec_total;
    e.erows = k.N * stage_doubles;
    e.req_want = a.req_want; e.req_b = a.req_b; e.req_p = a.req_p; e.req_g = a.req_g; e.req_val = a.req_val;
    e.list_grad = a.list_grad; e.list_val = a.list_val; e.escr = h->rs_escr; e.ctl = h->rs_counters + 8;
    e.split_ok = h->split_ok; e.pit_ok = h->pit_ok; e.work_blocks = (int)wblocks;
#define MPCB_LAUNCH_T(MODEL, GG)                                                               \
  do {                                                                                         \
    if (round == 0 && (rc = set_smem(h, k_solve<MODEL, GG, true>, box_smem))) return rc;       \
    k_solve<MODEL, GG, true><<<(unsigned)t_blocks, block, box_smem, st>>>(k, *sp, a);          \
  } while (0)
#define MPCB_LAUNCH_T_G(MODEL)                                                                 \
  switch (G) {                                                                                 \
    case 8: MPCB_LAUNCH_T(MODEL, 8); break;                                                    \
    default: MPCB_LAUNCH_T(MODEL, 1); break;                                                   \
  }
    if (k.model_id == MPCB_MODEL_PACEJKA6) {
      if ((rc = set_smem(h, k_eval_requests<MPCB_MODEL_PACEJKA6>, eval_smem))) return rc;
      if ((rc = set_carveout(h, k_eval_requests<MPCB_MODEL_PACEJKA6>, eval_smem, 1))) return rc;
    } else {
      if ((rc = set_smem(h, k_eval_requests<MPCB_MODEL_KINEMATIC4>, eval_smem))) return rc;
      if ((rc = set_carveout(h, k_eval_requests<MPCB_MODEL_KINEMATIC4>, eval_smem, 1))) return rc;
    }
    int64_t round = 0;
    // transition grid: every tile at first; after each chunk of rounds the live-tile count the host has just read bounds
    // it from above (tiles only finish), so the last rounds of a batch launch a few CTAs instead of thousands that exit
    int64_t t_blocks = blocks;
    // no instance can need more: <= ~40 rounds per PANOC iteration (9 trials x descent-lemma steps) over at most
    // max_iter iterations per PANOC run and ~8 runs of the tolerance continuation; a guard against a hang, not a limit
    const int64_t round_cap = std::min<int64_t>(4000000, 64 + (int64_t)40 * (sp->max_iter + 8) * 10);
    if (h->loop_graph && !h->round_log && h->eval_trace_round < 0) {
      // device-driven rounds: ONE graph launch, asynchronous on the caller's stream (include/mpcb200.h)
      if (!h->rs_stats) CK(cudaMalloc((void**)&h->rs_stats, kStatRounds * 2 * sizeof(unsigned)));
      void *t_func = nullptr, *e_func = nullptr;
      if (k.model_id == MPCB_MODEL_PACEJKA6) {
        e_func = (void*)k_eval_requests<MPCB_MODEL_PACEJKA6>;
        t_func = G == 8 ? (void*)k_solve<MPCB_MODEL_PACEJKA6, 8, true> : (void*)k_solve<MPCB_MODEL_PACEJKA6, 1, true>;
        if ((rc = G == 8 ? set_smem(h, k_solve<MPCB_MODEL_PACEJKA6, 8, true>, box_smem) : set_smem(h, k_solve<MPCB_MODEL_PACEJKA6, 1, true>, box_smem))) return rc;
      } else {
        e_func = (void*)k_eval_requests<MPCB_MODEL_KINEMATIC4>;
        t_func = G == 8 ? (void*)k_solve<MPCB_MODEL_KINEMATIC4, 8, true> : (void*)k_solve<MPCB_MODEL_KINEMATIC4, 1, true>;
        if ((rc = G == 8 ? set_smem(h, k_solve<MPCB_MODEL_KINEMATIC4, 8, true>, box_smem) : set_smem(h, k_solve<MPCB_MODEL_KINEMATIC4, 1, true>, box_smem))) return rc;
      }
      KDesc kk = k;
      MpcSolverParams spp = *sp;
      if ((rc = launch_graph_loop(h, st, t_func, e_func, kk, spp, a, e, blocks, block, box_smem, eblocks, eval_smem, G,
                                  (unsigned)std::min<int64_t>(round_cap, 0x7fffffff))))
        return rc;
      return leave(h, st);
    }
    int chunk = 48;
    for (cudaEvent_t& ev : h->rs_ev) if (!ev) CK(cudaEventCreate(&ev));
    if (h->debug_timers >= 2) {
      unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
      CK(cudaMemcpyToSymbolAsync(g_dbg_calls, z, sizeof(z), 0, cudaMemcpyHostToDevice, st));
    }
    struct LogFile {  // closed on every return path
      std::FILE* f;
      ~LogFile() { if (f) std::fclose(f); }
    } log_guard{h->round_log ? std::fopen(h->round_log, "w") : nullptr};
    std::FILE* log = log_guard.f;
    // MPCB_EVAL_TRACE=r: per-chunk cycle counts of evaluation launch r, printed to stderr (diagnostic)
    const long long trace_round = h->eval_trace_round;
    if (trace_round >= 0) {
      if ((rc = grow(h, &h->dbg, &h->dbg_cap, (size_t)(ns / 32 + 2) * 2 * sizeof(long long), st))) return rc;
      CK(cudaMemsetAsync(h->dbg, 0, (size_t)(ns / 32 + 2) * 2 * sizeof(long long), st));
    }
    h->last_eval_ms = 0.0; h->last_trans_ms = 0.0;
    if (!h->rs_stats) CK(cudaMalloc((void**)&h->rs_stats, kStatRounds * 2 * sizeof(unsigned)));
    h->rs_t_ms.clear(); h->rs_e_ms.clear();
    for (;;) {
      for (int i = 0; i < chunk; ++i, ++round) {
        a.round = (int)round;
        e.counters = h->rs_counters + 4 * (round & 1);
        e.key_cnt_zero = a.key_cnt + MPCB_NKEY * (round & 1);
        e.round_stats = round < kStatRounds ? h->rs_stats + 2 * round : nullptr;
        e.trace = (trace_round >= 0 && round == trace_round) ? (long long*)h->dbg : nullptr;
        CK(cudaEventRecord(h->rs_ev[3 * i], st));
        if (k.model_id == MPCB_MODEL_PACEJKA6) {
          MPCB_LAUNCH_T_G(MPCB_MODEL_PACEJKA6)
          CK(cudaEventRecord(h->rs_ev[3 * i + 1], st));
          k_eval_requests<MPCB_MODEL_PACEJKA6><<<(unsigned)eblocks, 256, eval_smem, st>>>(k, e);
        } else {
          MPCB_LAUNCH_T_G(MPCB_MODEL_KINEMATIC4)
          CK(cudaEventRecord(h->rs_ev[3 * i + 1], st));
          k_eval_requests<MPCB_MODEL_KINEMATIC4><<<(unsigned)eblocks, 256, eval_smem, st>>>(k, e);
        }
        CK(cudaEventRecord(h->rs_ev[3 * i + 2], st));
        if (log || i == chunk - 1)
          CK(cudaMemcpyAsync(h->rs_counters_host + 4 * i, h->rs_counters + 4 * (round & 1), 4 * sizeof(unsigned),
                             cudaMemcpyDeviceToHost, st));
      }
      h->launches += 2 * chunk;
      CK(cudaGetLastError());
      CK(cudaStreamSynchronize(st));
      for (int i = 0; i < chunk; ++i) {
        float t1 = 0.f, t2 = 0.f;
        CK(cudaEventElapsedTime(&t1, h->rs_ev[3 * i], h->rs_ev[3 * i + 1]));
        CK(cudaEventElapsedTime(&t2, h->rs_ev[3 * i + 1], h->rs_ev[3 * i + 2]));
        h->last_trans_ms += t1; h->last_eval_ms += t2;
        h->rs_t_ms.push_back(t1); h->rs_e_ms.push_back(t2);
        if (log) std::fprintf(log, "%lld %u %u %.4f %.4f\n", (long long)(round - chunk + i), h->rs_counters_host[4 * i],
                              h->rs_counters_host[4 * i + 1], t1, t2);
      }
      if (trace_round >= 0 && trace_round < round && trace_round >= round - chunk) {
        const int i = (int)(trace_round - (round - chunk));
        (void)i;
        std::vector<long long> tr((size_t)(ns / 32 + 2) * 2);
        CK(cudaMemcpy(tr.data(), h->dbg, tr.size() * sizeof(long long), cudaMemcpyDeviceToHost));
        std::vector<std::pair<long long, long long>> v;
        for (size_t c = 0; c < tr.size() / 2; ++c) if (tr[2 * c] > 0) v.push_back({tr[2 * c], (long long)c});
        std::sort(v.begin(), v.end());
        if (!v.empty()) {
          std::fprintf(stderr, "[mpcb trace] round %lld: %zu chunks, cycles min %lld median %lld p90 %lld max %lld\n", trace_round,
                       v.size(), v.front().first, v[v.size() / 2].first, v[v.size() * 9 / 10].first, v.back().first);
          long long nf_chunks = 0, nf_cyc = 0, ok_cyc = 0;
          for (auto& w : v) { if (tr[2 * w.second + 1] % 1000) { ++nf_chunks; nf_cyc += w.first; } else ok_cyc += w.first; }
          std::fprintf(stderr, "[mpcb trace]   chunks with a non-finite psi: %lld (mean cycles %lld), without: %zu (mean cycles %lld)\n",
                       nf_chunks, nf_chunks ? nf_cyc / nf_chunks : 0, v.size() - (size_t)nf_chunks,
                       v.size() > (size_t)nf_chunks ? ok_cyc / (long long)(v.size() - nf_chunks) : 0);
          for (size_t k = 0; k < 8 && k < v.size(); ++k) {
            const auto& w = v[v.size() - 1 - k];
            std::fprintf(stderr, "[mpcb trace]   chunk %lld cycles %lld kind+nonfinite lanes %lld\n", w.second, w.first, tr[2 * w.second + 1]);
          }
        }
      }
      const unsigned* last = h->rs_counters_host + 4 * (chunk - 1);
      if (last[0] == 0 && last[1] == 0) break;
      {
        const int64_t live = last[3];
        const int64_t need = ((live + (int64_t)MPCB_NKEY * (32 / G)) * G + block - 1) / block;
        t_blocks = need < blocks ? (need > 0 ? need : 1) : blocks;
      }
      if (round > round_cap) return fail(h, MPCB_E_CUDA, "mpcb_solve: round cap exceeded");
      chunk = 24;
    }
#undef MPCB_LAUNCH_T_G
#undef MPCB_LAUNCH_T
    if (h->debug_timers >= 2) {
      unsigned long long c[8];
      if (cudaMemcpyFromSymbol(c, g_dbg_calls, sizeof(c)) == cudaSuccess)
        std::fprintf(stderr, "[mpcb debug] model calls by lane: fast %llu wide %llu nan %llu slow %llu; warp calls %llu, "
                     "with a wide lane %llu, with a slow lane %llu\n", c[0], c[2], c[3], c[4], c[5], c[6], c[7]);
    }
    h->last_rounds = round;
    {  // bulk / tail split of the profile: which rounds carried more requests than the evaluation launch has lanes
      const int64_t nr = std::min<int64_t>(round, kStatRounds);
      h->rs_stats_host.resize((size_t)nr * 2);
      CK(cudaMemcpy(h->rs_stats_host.data(), h->rs_stats, (size_t)nr * 2 * sizeof(unsigned), cudaMemcpyDeviceToHost));
      h->bulk_eval_ms = h->bulk_trans_ms = h->bulk_grad = h->bulk_val = 0.0; h->bulk_rounds = 0;
      for (int64_t r = 0; r < nr && r < (int64_t)h->rs_e_ms.size(); ++r) {
        const double g = h->rs_stats_host[2 * r], v = h->rs_stats_host[2 * r + 1];
        if (g + v > (double)ens) {
          h->bulk_eval_ms += h->rs_e_ms[r]; h->bulk_trans_ms += h->rs_t_ms[r]; h->bulk_grad += g; h->bulk_val += v; h->bulk_rounds++;
        }
      }
    }
    return leave(h, st);
  }
  if (h->debug_timers) {
    unsigned long long z[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    CK(cudaMemcpyToSymbolAsync(g_dbg_calls, z, sizeof(z), 0, cudaMemcpyHostToDevice, st));
    if ((rc = grow(h, &h->dbg, &h->dbg_cap, (size_t)(ns / 32) * 4 * sizeof(long long), st))) return rc;
    CK(cudaMemsetAsync(h->dbg, 0, (size_t)(ns / 32) * 4 * sizeof(long long), st));
    a.dbg = h->dbg;
    h->dbg_warps = ns / 32;
  }
#define MPCB_LAUNCH_SOLVE(MODEL, GG)                                                           \
  do {                                                                                         \
    if ((rc = set_smem(h, k_solve<MODEL, GG, false>, smem))) return rc;                        \
    k_solve<MODEL, GG, false><<<(unsigned)blocks, block, smem, st>>>(k, *sp, a);               \
  } while (0)
#define MPCB_LAUNCH_SOLVE_G(MODEL)                                                             \
  switch (G) {                                                                                 \
    case 8: MPCB_LAUNCH_SOLVE(MODEL, 8); break;                                                \
    default: MPCB_LAUNCH_SOLVE(MODEL, 1); break;                                               \
  }
  if (k.model_id == MPCB_MODEL_PACEJKA6) { MPCB_LAUNCH_SOLVE_G(MPCB_MODEL_PACEJKA6) }
  else { MPCB_LAUNCH_SOLVE_G(MPCB_MODEL_KINEMATIC4) }
#undef MPCB_LAUNCH_SOLVE_G
#undef MPCB_LAUNCH_SOLVE
  h->last_tile = G;
  h->launches++;
  return leave(h, st);
}

}  // extern "C"

// ---- host-buffer entry point + FP64 peak probe ------------------------------------------------------
namespace {
struct Packer {
  size_t off = 0;
  size_t add(size_t bytes) { size_t o = off; off += (bytes + 255) & ~size_t(255); return o; }
};
}  // namespace

extern "C" {

int mpcb_solve_host(MpcHandle* h, const MpcSolverParams* sp, int64_t B, const double* x0, const double* p,
                    const double* road, const int32_t* road_idx, const double* obs, const double* lb,
                    const double* ub, double* U, double* lambda, int32_t* status, int32_t* iters, int32_t* evals,
                    double* psi, double* res) {
  if (!h || !sp || !x0 || !p || !road || !lb || !ub || !U || !status || !iters || !evals || !psi || !res || B < 0)
    if (!(h && sp && B == 0)) return fail(h, MPCB_E_ARG, "mpcb_solve_host: bad argument");
  if (B == 0) return MPCB_OK;
  const KDesc& k = h->kd;
  cudaStream_t st = h->own_stream;
  size_t road_elems = 0;
  switch (k.road_mode) {
    case MPCB_ROAD_SHARED: road_elems = 2 * (size_t)k.S; break;
    case MPCB_ROAD_TABLE: road_elems = (size_t)k.n_roads * 2 * k.S; break;
    case MPCB_ROAD_PER_INSTANCE: road_elems = (size_t)B * 2 * k.S; break;
    case MPCB_ROAD_BEZIER: road_elems = (size_t)B * 12; break;
  }
  const bool has_obs = (k.flags & MPCB_FLAG_OBST) && k.n_obs > 0 && obs;
  const bool has_idx = k.road_mode == MPCB_ROAD_TABLE && road_idx;
  // inputs first (one H2D), in/out + outputs last (one D2H)
  Packer pk;
  size_t o_x0 = pk.add(sizeof(double) * B * k.nx), o_p = pk.add(sizeof(double) * MPCB_NP);
  size_t o_road = pk.add(sizeof(double) * road_elems), o_idx = pk.add(has_idx ? sizeof(int32_t) * B : 0);
  size_t o_obs = pk.add(has_obs ? sizeof(double) * B * k.n_obs * MPCB_OBS_DIM : 0);
  size_t o_lb = pk.add(sizeof(double) * k.n), o_ub = pk.add(sizeof(double) * k.n);
  size_t o_U = pk.add(sizeof(double) * B * k.n);
  // multipliers travel only when D is bounded: with D = R^m (the reference's case, main.py:57) the ALM never reads or
  // writes them, and [B, nx N] doubles each way is 2.6x the rest of the call's traffic
  if (!k.bounded) lambda = nullptr;
  size_t o_lam = pk.add(lambda ? sizeof(double) * B * k.m : 0);
  size_t in_end = pk.off;
  size_t o_status = pk.add(sizeof(int32_t) * B), o_iters = pk.add(sizeof(int32_t) * 2 * B);
  size_t o_evals = pk.add(sizeof(int32_t) * 2 * B), o_psi = pk.add(sizeof(double) * B), o_res = pk.add(sizeof(double) * B);
  size_t total = pk.off;
  if (total > h->pin_cap) {
    if (h->pin) { CK(cudaStreamSynchronize(st)); CK(cudaFreeHost(h->pin)); h->pin = nullptr; h->pin_cap = 0; }
    CK(cudaMallocHost(&h->pin, total + total / 8));
    h->pin_cap = total + total / 8;
  }
  int rc;
  if ((rc = grow(h, (char**)&h->dev, &h->dev_cap, total, st))) return rc;
  char* hp = (char*)h->pin;
  char* dp = (char*)h->dev;
  std::memcpy(hp + o_x0, x0, sizeof(double) * B * k.nx);
  std::memcpy(hp + o_p, p, sizeof(double) * MPCB_NP);
  std::memcpy(hp + o_road, road, sizeof(double) * road_elems);
  if (has_idx) std::memcpy(hp + o_idx, road_idx, sizeof(int32_t) * B);
  if (has_obs) std::memcpy(hp + o_obs, obs, sizeof(double) * B * k.n_obs * MPCB_OBS_DIM);
  std::memcpy(hp + o_lb, lb, sizeof(double) * k.n);
  std::memcpy(hp + o_ub, ub, sizeof(double) * k.n);
  std::memcpy(hp + o_U, U, sizeof(double) * B * k.n);
  if (lambda) std::memcpy(hp + o_lam, lambda, sizeof(double) * B * k.m);
  CK(cudaMemcpyAsync(dp, hp, in_end, cudaMemcpyHostToDevice, st));
  rc = mpcb_solve(h, sp, B, (double*)(dp + o_x0), (double*)(dp + o_p), (double*)(dp + o_road),
                  has_idx ? (int32_t*)(dp + o_idx) : nullptr, has_obs ? (double*)(dp + o_obs) : nullptr,
                  (double*)(dp + o_lb), (double*)(dp + o_ub), (double*)(dp + o_U),
                  lambda ? (double*)(dp + o_lam) : nullptr, (int32_t*)(dp + o_status), (int32_t*)(dp + o_iters),
                  (int32_t*)(dp + o_evals), (double*)(dp + o_psi), (double*)(dp + o_res), st);
  if (rc) return rc;
  CK(cudaMemcpyAsync(hp + o_U, dp + o_U, total - o_U, cudaMemcpyDeviceToHost, st));
  CK(cudaStreamSynchronize(st));
  std::memcpy(U, hp + o_U, sizeof(double) * B * k.n);
  if (lambda) std::memcpy(lambda, hp + o_lam, sizeof(double) * B * k.m);
  std::memcpy(status, hp + o_status, sizeof(int32_t) * B);
  std::memcpy(iters, hp + o_iters, sizeof(int32_t) * 2 * B);
  std::memcpy(evals, hp + o_evals, sizeof(int32_t) * 2 * B);
  std::memcpy(psi, hp + o_psi, sizeof(double) * B);
  std::memcpy(res, hp + o_res, sizeof(double) * B);
  h->h2d_bytes = (int64_t)in_end;
  h->d2h_bytes = (int64_t)(total - o_U);
  return MPCB_OK;
}

/* Diagnostic (MPCB_DEBUG_TIMERS=1): per-warp [eval cycles, transition cycles, rounds, gradient rounds] of the
 * last mpcb_solve; returns the number of warps copied (<= max_warps), synchronises the device. */
int64_t mpcb_debug_timers(MpcHandle* h, long long* out, int64_t max_warps) {
  if (!h || !h->dbg || !out) return 0;
  int64_t nw = h->dbg_warps < max_warps ? h->dbg_warps : max_warps;
  if (cudaDeviceSynchronize() != cudaSuccess) return 0;
  if (cudaMemcpy(out, h->dbg, (size_t)nw * 4 * sizeof(long long), cudaMemcpyDeviceToHost) != cudaSuccess) return 0;
  unsigned long long c[2] = {0, 0};
  if (h->debug_timers >= 2 && cudaMemcpyFromSymbol(c, g_dbg_calls, sizeof(c)) == cudaSuccess)
    std::fprintf(stderr, "[mpcb debug] model calls: fast path %llu, general path %llu (%.2f%%)\n", c[0], c[1],
                 100.0 * (double)c[1] / (double)(c[0] + c[1] + 1));
  return nw;
}

void mpcb_default_car_params(double* c11) {  // game_theory.py:23-56
  const double v[11] = {4.2, 1.8, 3.0, 3.2 / 180.0 * 3.14159265358979323846, 5.17, 1.2, 0.15, 0.9, 7.0, 3.75, 1.0};
  for (int i = 0; i < 11; ++i) c11[i] = v[i];
}

int mpcb_game_payoff(MpcHandle* h, int64_t B, int32_t M, const double* ego, const double* cars, const double* car_params,
                     double* payoff_out, int32_t* decision_out, double* ctrl_out, double* tca_out, void* stream) {
  if (!h || B < 0 || M < 0 || M > 64) return fail(h, MPCB_E_ARG, "mpcb_game_payoff: bad argument (0 <= M <= 64)");
  if (B == 0) return MPCB_OK;
  if (!ego || (!cars && M > 0) || !payoff_out) return fail(h, MPCB_E_ARG, "mpcb_game_payoff: null argument");
  double v[11];
  mpcb_default_car_params(v);
  if (car_params) for (int i = 0; i < 11; ++i) v[i] = car_params[i];  // HOST pointer, like MpcSolverParams
  GamePar c{v[0], v[1], v[2], v[3], v[4], v[5], v[6], v[7], v[8], v[9], v[10]};
  cudaStream_t st = (cudaStream_t)stream;
  k_game_payoff<<<(unsigned)((B + 127) / 128), 128, 0, st>>>(B, M, c, ego, cars, payoff_out, decision_out, ctrl_out, tca_out);
  h->launches++;
  CK(cudaGetLastError());
  return MPCB_OK;
}

int64_t mpcb_check_guards(MpcHandle* h) {
  if (!h) return -1;
  if (!h->guards) return -2;  // created without MPCB_GUARDS=1
  if (cudaDeviceSynchronize() != cudaSuccess) return -3;
  std::vector<unsigned char> buf(kGuardBytes);
  int64_t bad = 0;
  for (const auto& g : h->guarded) {
    for (int side = 0; side < 2; ++side) {
      const char* src = (const char*)g.base + (side ? kGuardBytes + g.bytes : 0);
      if (cudaMemcpy(buf.data(), src, kGuardBytes, cudaMemcpyDeviceToHost) != cudaSuccess) return -3;
      for (unsigned char c : buf) bad += c != 0xA5;
    }
  }
  return bad;
}

int mpcb_last_finish_ns(MpcHandle* h, int64_t B, uint64_t* out, void* stream) {
  if (!h || !out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_last_finish_ns: bad argument");
  if (B != h->gam_B || !h->fin) return fail(h, MPCB_E_ARG, "mpcb_last_finish_ns: B must be the batch of the last solve on this handle");
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if ((rc = enter(h, st))) return rc;
  CK(cudaMemcpyAsync(out, h->fin, (size_t)(B + 1) * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
  return leave(h, st);
}

int mpcb_set_trace(MpcHandle* h, int32_t cap) {
  if (!h || cap < 0 || cap > (1 << 20)) return fail(h, MPCB_E_ARG, "mpcb_set_trace: bad argument");
  h->trace_cap = cap;
  return MPCB_OK;
}

int mpcb_last_trace(MpcHandle* h, int64_t B, double* trace_out, void* stream) {
  if (!h || !trace_out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_last_trace: bad argument");
  if (B > h->trace_B || !h->trace || h->trace_cap <= 0) return fail(h, MPCB_E_ARG, "mpcb_last_trace: the last solve on this handle recorded no trace for that many instances (mpcb_set_trace)");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if ((rc = enter(h, st))) return rc;
  CK(cudaMemcpyAsync(trace_out, h->trace, (size_t)B * h->trace_cap * 4 * sizeof(double), cudaMemcpyDeviceToDevice, st));
  return leave(h, st);
}

int mpcb_last_inner_failures(MpcHandle* h, int64_t B, int32_t* out, void* stream) {
  if (!h || !out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_last_inner_failures: bad argument");
  if (B > h->gam_B || !h->ifail) return fail(h, MPCB_E_ARG, "mpcb_last_inner_failures: the last solve on this handle had fewer instances");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if ((rc = enter(h, st))) return rc;
  CK(cudaMemcpyAsync(out, h->ifail, (size_t)B * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
  return leave(h, st);
}

int mpcb_last_final_gamma(MpcHandle* h, int64_t B, double* gamma_out, void* stream) {
  if (!h || !gamma_out || B < 0) return fail(h, MPCB_E_ARG, "mpcb_last_final_gamma: bad argument");
  if (B > h->gam_B || !h->gam) return fail(h, MPCB_E_ARG, "mpcb_last_final_gamma: the last solve on this handle had fewer instances");
  if (B == 0) return MPCB_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  if ((rc = enter(h, st))) return rc;
  CK(cudaMemcpyAsync(gamma_out, h->gam, (size_t)B * sizeof(double), cudaMemcpyDeviceToDevice, st));
  return leave(h, st);
}

void mpcb_last_transfer_bytes(MpcHandle* h, int64_t* h2d, int64_t* d2h) {
  if (h2d) *h2d = h ? h->h2d_bytes : 0;
  if (d2h) *d2h = h ? h->d2h_bytes : 0;
}

/* Dependent-chain latencies (SM cycles per op, one warp): [0] DFMA [1] div [2] sqrt [3] sincos [4] atan
 * [5] atan2 [6] sin [7] one Pacejka model call.  Diagnostic for DESIGN.md / profiles/. */
int mpcb_latency_probe(double* cycles8) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MPCB_E_NODEVICE;
  double* out = nullptr;
  if (cudaMalloc((void**)&out, sizeof(double) * 64) != cudaSuccess) return MPCB_E_CUDA;
  k_latency_probe<<<1, 32>>>(out, 0.37, 2000);
  k_latency_probe<<<1, 32>>>(out, 0.37, 2000);
  cudaError_t e = cudaMemcpy(cycles8, out, sizeof(double) * 8, cudaMemcpyDeviceToHost);
  cudaFree(out);
  return e == cudaSuccess ? MPCB_OK : MPCB_E_CUDA;
}

int mpcb_fp64_peak(double* tflops_out, double* ms_out) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) return MPCB_E_NODEVICE;
  int dev = 0, sms = 148;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int block = 256, blocks = sms * 8, iters = 4000;
  double* out = nullptr;
  if (cudaMalloc((void**)&out, sizeof(double) * block * blocks) != cudaSuccess) return MPCB_E_CUDA;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  float best = 1e30f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(e0);
    k_fp64_peak<<<blocks, block>>>(out, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(out); return MPCB_E_CUDA; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(out);
  double flops = 2.0 * 64.0 * iters * (double)block * blocks;
  if (tflops_out) *tflops_out = flops / (best * 1e-3) / 1e12;
  if (ms_out) *ms_out = best;
  return MPCB_OK;
}

}  // extern "C"

// ---- allocation-free psi + grad psi (no counterpart in the reference) ----
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
// Same kernel as mpcb_value_grad (k_value_grad), but its stage-Jacobian scratch [N * stage_doubles, lanes] lives in
// a file-scope static device array instead of the handle's grow-only cudaMalloc'ed workspace.  Every GPU that loads
// the library holds MPCB_STATIC_SCRATCH_BYTES until exit.  The array is shared by EVERY call on a device (all
// handles, all streams) and keeps the last call's contents: nothing in it has to start at zero (a lane writes its
// stage rows in the forward pass before its adjoint reads them back), but two calls that overlap on the GPU race.
#include <mutex>

namespace {
constexpr size_t kStaticScratchDoubles = MPCB_STATIC_SCRATCH_BYTES / sizeof(double);
static __device__ __align__(256) double g_static_scratch[kStaticScratchDoubles];
// In host code `g_static_scratch` is a host-side stand-in, not a device address: the device address is looked up
// once per device (with that device current) and cached here.
constexpr int kMaxDevices = 64;
double* g_static_scratch_ptr[kMaxDevices] = {};
std::mutex g_static_scratch_mu;

int static_scratch_ptr(MpcHandle* h, double** out) {
  int dev = -1;
  CK(cudaGetDevice(&dev));
  if (dev != h->device) return fail(h, MPCB_E_ARG, "mpcb_value_grad_static: the handle's device is not the current device");
  if (dev < 0 || dev >= kMaxDevices) return fail(h, MPCB_E_UNSUPPORTED, "mpcb_value_grad_static: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_static_scratch_mu);
  if (!g_static_scratch_ptr[dev]) {
    double* ptr = nullptr;
    CK(cudaGetSymbolAddress((void**)&ptr, g_static_scratch));
    if (!ptr) return fail(h, MPCB_E_CUDA, "mpcb_value_grad_static: cudaGetSymbolAddress returned a null address");
    g_static_scratch_ptr[dev] = ptr;
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}
}  // namespace

extern "C" {

int mpcb_value_grad_static(MpcHandle* h, int64_t B, const double* x0, const double* U, const double* p,
                           const double* road, const int32_t* road_idx, const double* obs, const double* lambda,
                           const double* Sigma, double* cost_out, double* grad_out, void* stream) {
  if (!h || B < 0) return fail(h, MPCB_E_ARG, "mpcb_value_grad_static: bad argument");
  if (B == 0) return MPCB_OK;
  if (!x0 || !U || !p || !road || !cost_out || !grad_out) return fail(h, MPCB_E_ARG, "mpcb_value_grad_static: null argument");
  if (B > (int64_t)MPCB_STATIC_MAX_B)
    return fail(h, MPCB_E_UNSUPPORTED, "mpcb_value_grad_static: B = " + std::to_string(B) + " exceeds MPCB_STATIC_MAX_B = " +
                                           std::to_string((long long)MPCB_STATIC_MAX_B) + " (use mpcb_value_grad)");
  const KDesc& k = h->kd;
  if ((k.flags & MPCB_FLAG_OBST) && k.n_obs > 0 && !obs) return fail(h, MPCB_E_ARG, "obs required with MPCB_FLAG_OBST");
  // per-instance and Bezier roads need a [2S, B] transposed copy in grow-only (cudaMalloc'ed) scratch
  if (k.road_mode != MPCB_ROAD_SHARED && k.road_mode != MPCB_ROAD_TABLE)
    return fail(h, MPCB_E_UNSUPPORTED, "mpcb_value_grad_static: only MPCB_ROAD_SHARED / MPCB_ROAD_TABLE roads (the other "
                                       "layouts need allocated scratch; use mpcb_value_grad)");
  cudaStream_t st = (cudaStream_t)stream;
  int rc;
  double* scratch = nullptr;
  if ((rc = static_scratch_ptr(h, &scratch))) return rc;
  const int block = 128;
  const size_t rows = (size_t)k.N * (k.model_id == MPCB_MODEL_PACEJKA6 ? fm_stage_doubles<MPCB_MODEL_PACEJKA6>()
                                                                       : fm_stage_doubles<MPCB_MODEL_KINEMATIC4>());
  int64_t blocks = (B + block - 1) / block;
  const int64_t max_blocks = (int64_t)h->num_sms * 8;
  const int64_t cap_blocks = (int64_t)(kStaticScratchDoubles / (rows * block));  // lanes the static scratch holds
  if (cap_blocks < 1)
    return fail(h, MPCB_E_UNSUPPORTED, "mpcb_value_grad_static: one CTA's stage rows (" + std::to_string(rows * block * 8) +
                                           " bytes) exceed MPCB_STATIC_SCRATCH_BYTES");
  if (blocks > max_blocks) blocks = max_blocks;
  if (blocks > cap_blocks) blocks = cap_blocks;
  const int64_t ns = blocks * block;
  if ((rc = enter(h, st))) return rc;
  RoadArg ra; size_t smem;
  if ((rc = prepare_road(h, B, road, road_idx, st, &ra, &smem))) return rc;  // table modes: no launch, no allocation
  if (k.model_id == MPCB_MODEL_PACEJKA6) {
    if ((rc = set_smem(h, k_value_grad<MPCB_MODEL_PACEJKA6>, smem))) return rc;
    k_value_grad<MPCB_MODEL_PACEJKA6><<<(unsigned)blocks, block, smem, st>>>(k, B, x0, U, p, ra, obs, lambda, Sigma, cost_out, grad_out, scratch, ns);
  } else {
    if ((rc = set_smem(h, k_value_grad<MPCB_MODEL_KINEMATIC4>, smem))) return rc;
    k_value_grad<MPCB_MODEL_KINEMATIC4><<<(unsigned)blocks, block, smem, st>>>(k, B, x0, U, p, ra, obs, lambda, Sigma, cost_out, grad_out, scratch, ns);
  }
  h->launches++;
  return leave(h, st);
}

}  // extern "C"
