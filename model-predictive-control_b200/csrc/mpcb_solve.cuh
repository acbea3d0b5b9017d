// mpcb_solve.cuh — the fused, persistent ALM + structured-PANOC + masked-L-BFGS kernel.
//
// Replaces `self.solver(problem, U, λ)` (reference controller.py:57): alpaqa's ALMSolver over
// StructuredPANOCLBFGSSolver calling CasADi-generated f/grad_f.  Canonical algorithm: SURVEY.md
// Appendix A (restated on the CPU in oracle/orc_solver.hpp).
//
// Mapping ("lane tile per instance").  One solve is a chain of ~300 dependent rollouts, each a chain
// of 16 N dependent model calls: the kernel is bound by that latency, not by throughput, until the
// batch exceeds ~20 k instances per GPU.  A tile of G lanes (G = 8, 4, 2, 1, picked by the host from
// the batch size) therefore owns one instance and spends its spare lanes on SPECULATION that removes
// evaluations from the dependent chain:
//   * line search: lane j evaluates the candidate tau = 2^-j (all 8 back-tracking trials of
//     Appendix A 3d at once for G = 8), then its own psi(x̂+); the first accepted candidate in
//     sequential order wins — the same decisions as sequential back-tracking because every trial
//     restarts from the iterate's (L, gamma);
//   * the finite-difference Hessian-vector gradient of the NEXT iteration (Appendix A 3c) is evaluated by the
//     upper half of the tile in the same round in which the lower half evaluates psi(x̂+) of the first
//     G/2 candidates: a typical iteration is two rounds (candidates, then psi(x̂+) || FD) instead of three;
//   * descent lemma at k = 0: lane j evaluates psi(x̂(gamma / 2^j));
//   * Lipschitz estimate: lane 1 evaluates grad psi(x + h) while lane 0 evaluates grad psi(x).
// Every lane of a tile runs the same state machine on the same data (redundantly — free in SIMT), so
// control flow is tile-uniform; per-lane candidate results only enter through tile ballots/shuffles.
// All lanes of a warp meet at ONE evaluation site per loop trip, so the dominant fp64 work never
// diverges even though tiles are at different PANOC iterations.  Finished tiles pull the next
// instance from a global queue.
//
// Workspace.  Per-lane rows (iterate banks, candidate, stage Jacobians) are lane-interleaved in BLOCKS of 32 lanes
// ([block][row][lane in block], MPCB_LB): a warp's access to element i of its 32 lanes is one 256-byte transaction, and the
// rows of a block are contiguous, so everything a tile touches in a round lies within a few dozen KB.  (Round 1 interleaved
// over ALL lanes, [row][lane]: consecutive elements of a vector were ns * 8 bytes = 4 MB apart at 64K instances, every
// access of a tile's serial chains landed on another 2 MB page, and a transition launch for a handful of tiles cost
// 95 us against 32 us for the same work at B = 1.)  Per-tile vectors (q, L-BFGS
// history) use a tile-vector layout in which the G lanes of a tile own G consecutive doubles, so the
// two-loop recursion is split over the lanes with coalesced accesses and tile-shuffle dot products.
// Locations shared by a tile are either written by their owner lane, or written by every lane with
// the same value that does not depend on the old content; read-modify-writes are done by lane 0
// between tile barriers.
#pragma once
#include <float.h>
#include <type_traits>
#include "mpcb_fwd.cuh"

namespace mpcb {

struct SolveArgs {
  int64_t B;
  const double *x0, *p, *obs, *lb, *ub;
  RoadArg ra;
  double *U, *lam;
  int32_t *status, *iters, *evals;
  double *psi, *res;
  unsigned long long* fin_ns;  // optional [B + 1]: %globaltimer (ns) when instance b finished; [B] = start of the solve
  int32_t* ifail_out;  // [B] PANOC runs of the ALM loop that did not converge (alpaqa stats: inner_convergence_failures)
  double* trace;       // optional [B, trace_cap, 4]: (psi, phi_gamma, gamma, accepted candidate) at the top of every inner iteration (the progress callback's info)
  int trace_cap;
  double* gam_out;  // [B] final step size gamma of every instance's last PANOC run (alpaqa's inner stats: final γ)
  double* ws;    // per-lane rows   [ns / MPCB_LB][LaneMap.total][MPCB_LB]
  double* wst;   // per-tile data: tile vectors [ns / MPCB_LB][vec rows][MPCB_LB], then plain rows [ns / MPCB_LB][rows][MPCB_LB / G]
  int64_t ns, nt;
  unsigned long long* queue;
  long long* dbg;  // optional [warps, 4]: cycles in evaluation / transition phases, rounds, gradient rounds
  // ---- round mode (k_solve<.., ROUND = true> + k_eval_requests, see below) ----
  int round;                 // index of this transition launch (host-driven loop)
  const unsigned* round_ptr; // graph-driven loop: the round index lives in device memory (ctl[1]); null otherwise
  double* sd;                // saved lane state, doubles  [ns / MPCB_LB][MPCB_SD_ROWS][MPCB_LB]
  int32_t* si;               // saved lane state, ints     [ns / MPCB_LB][MPCB_SI_ROWS][MPCB_LB]
  int32_t* req_want;         // [ns] evaluation request of every lane: 0 none, 1 psi, 2 psi + grad psi, 3 psi + states
  int64_t* req_b;            // [ns] instance
  double** req_p;            // [ns] point to evaluate (element i at [i * MPCB_LB])
  double** req_g;            // [ns] gradient row (want 2) / state row (want 3)
  double* req_val;           // [ns] psi written by the evaluation kernel
  int32_t* list_grad;        // compacted lanes whose request wants a gradient
  int32_t* list_val;         // compacted lanes whose request wants psi only
  unsigned* counters;        // [2][4] by round parity: gradient requests, value requests, evaluation work counter, live tiles
  // Speculation window (tau candidates evaluated per round, 1 .. G-1): spec_w > 0 fixes it (MPCB_SPECW); 0 = picked on
  // the device from the number of tiles still solving (ctl[0], written by the previous round's evaluation launch).
  // Tile order of the transition launches: tiles are bucketed by what they will do next (state + line-search phase)
  // so that the tiles of a warp, and the warps that follow each other on an SM, run the same code.  key_cnt[parity]
  // [MPCB_NKEY] and key_list[parity][MPCB_NKEY][nt] (tile indices): launch r reads parity r & 1 and fills the other.
  int32_t* key_cnt;
  int32_t* key_list;
  int spec_w;
  int spec_live_wide, spec_live_mid;  // live tiles up to which the window is G-1 wide / 3 wide (1 beyond)
  const unsigned* ctl;       // [4] device control words: [0] live tiles of the previous round
};
constexpr int MPCB_SD_ROWS = 25, MPCB_SI_ROWS = 28;
constexpr int MPCB_LB = 32;         // lanes per block of the lane-interleaved arrays
constexpr int MPCB_NKEY = 16;       // transition buckets: one per SolveState, ST_LS_STEP split by phase
#ifndef MPCB_ROUND_BLOCK_N
#define MPCB_ROUND_BLOCK_N 128
#endif
constexpr int MPCB_ROUND_BLOCK = MPCB_ROUND_BLOCK_N;  // lanes per CTA of a transition launch (64 measured 2 % slower: 1 510 vs 1 478 ms per 64K batch)

enum SolveState {
  ST_IDLE = 0, ST_START_PANOC, ST_INIT_G1, ST_LIP_DONE, ST_XHAT_REQ, ST_XHAT_DONE, ST_ITER_TOP, ST_FD_DONE,
  ST_LS_START, ST_LS_STEP, ST_PANOC_EXIT, ST_ALM_V, ST_ALM_UPDATE
};

// Per-lane workspace rows (units: doubles per lane).
struct LaneMap {
  int n, SCR, ST, total;
  __host__ __device__ LaneMap(int n_, int N, int nx, int stage_doubles, bool bounded) : n(n_) {
    SCR = 8 * n; ST = SCR + N * stage_doubles; total = ST + (bounded ? N * nx : 0);
  }
  __host__ __device__ int X(int bank) const { return bank * 4 * n; }
  __host__ __device__ int G(int bank) const { return bank * 4 * n + n; }
  __host__ __device__ int XH(int bank) const { return bank * 4 * n + 2 * n; }
  __host__ __device__ int P(int bank) const { return bank * 4 * n + 3 * n; }
};
// Per-tile workspace (units: doubles per tile).  The first `vec_total` doubles are tile-vectors of
// padded length np (Q, XB, S_0.., Y_0..); the rest are plain per-tile rows.
struct TileMap {
  int np, Q, XB, LS, LY, vec_total, AL, RH, SIG, E1, E2, total;
  __host__ __device__ TileMap(int n, int G, int mem, int m, bool bounded) {
    np = (n + G - 1) / G * G;
    Q = 0; XB = np; LS = 2 * np; LY = LS + mem * np; vec_total = LY + mem * np;
    AL = vec_total; RH = AL + mem; SIG = RH + mem;
    E1 = SIG + (bounded ? m : 0); E2 = E1 + (bounded ? m : 0); total = E2 + (bounded ? m : 0);
  }
};

#ifndef MPCB_T_MINB
#define MPCB_T_MINB 1  // min resident CTAs per SM asked of the transition launch (register cap); A/B with -DMPCB_T_MINB=n
#endif
template <int MODEL, int G, bool ROUND>
__global__ void __launch_bounds__(ROUND ? MPCB_ROUND_BLOCK : 64, ROUND ? MPCB_T_MINB : 1)
k_solve(KDesc d, MpcSolverParams sp, SolveArgs a) {
  constexpr int NX = Fm<MODEL>::NX;
  constexpr int K = ROUND ? 0 : fm_stage_doubles<MODEL>();  // round mode: the stage Jacobians live in the evaluation kernel's scratch
  extern __shared__ double smem[];
  const int n = d.n, m = d.m, mem = sp.lbfgs_mem;
  double* sroad = smem;
  double* sbox = smem + (!ROUND && a.ra.in_smem ? a.ra.R * road_smem_doubles(d.S) : 0);  // lb[n], ub[n]
  const int rnd = (ROUND && a.round_ptr) ? (int)*a.round_ptr : a.round;  // this launch's round
  if (a.fin_ns && threadIdx.x == 0 && (!ROUND || rnd == 0)) {  // start of the solve: earliest CTA of the first launch
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    atomicMin(&a.fin_ns[a.B], t);
  }
  // ---- which tile does this group of G lanes run? ----
  // persistent kernel: tile = position in the grid.  Round mode: launch r takes the tiles in the order launch r-1
  // bucketed them (key_list), every bucket padded to whole warps; groups beyond the last live tile have none.
  int64_t tile_sel = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / G;
  bool notile = false;
  if constexpr (!ROUND) {
    fm_load_smem();
    if (a.ra.in_smem) stage_roads(sroad, a.ra.road, a.ra.R, d.S);
  } else {
    constexpr int TPW = 32 / G;  // tiles per warp
    if (rnd > 0) {
      const int32_t* kc = a.key_cnt + MPCB_NKEY * (rnd & 1);
      int64_t g = tile_sel, start = 0;
      int key = -1;
      int64_t off = 0;
      int kcv[MPCB_NKEY];  // all counts requested at once (16 dependent L2 round trips otherwise, in every launch)
#pragma unroll
      for (int kq = 0; kq < MPCB_NKEY; ++kq) kcv[kq] = kc[kq];
#pragma unroll
      for (int kq = 0; kq < MPCB_NKEY; ++kq) {
        const int64_t c = kcv[kq];
        const int64_t cp = (c + TPW - 1) / TPW * TPW;
        if (key < 0 && g < start + cp) { key = kq; off = g - start; if (off >= c) key = MPCB_NKEY; }
        start += cp;
      }
      // `start` = padded number of groups: a CTA past it has nothing to do (late in a batch: most of the grid)
      if ((int64_t)blockIdx.x * (blockDim.x / G) >= start) {
        if (blockIdx.x == 0 && threadIdx.x == 0) {  // (no live tile at all: the batch is done)
          unsigned* nx = a.counters + 4 * ((rnd + 1) & 1);
          nx[0] = 0u; nx[1] = 0u; nx[2] = 0u; nx[3] = 0u;
        }
        return;
      }
      if (key < 0 || key >= MPCB_NKEY) notile = true;
      else tile_sel = a.key_list[((int64_t)(rnd & 1) * MPCB_NKEY + key) * a.nt + off];
    } else if (tile_sel >= a.nt) {
      notile = true;
    }
    if (notile) tile_sel = 0;
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) { sbox[i] = a.lb[i]; sbox[n + i] = a.ub[i]; }
  if constexpr (ROUND) __syncthreads();
  else stage_problem(d, a.p);
  const double* lbv = sbox;
  const double* ubv = sbox + n;
  // two-loop recursion scalars (rho_i, alpha_i) of this CTA's tiles: [tile in CTA][2 * mem]
  double* const s_ra = sbox + 2 * n + (threadIdx.x / G) * 2 * mem;

  const int64_t nt = a.nt;
  const int lane = threadIdx.x & 31;
  const int l = lane % G;  // lane within the tile
  const int tbase = lane - l;
  const unsigned tmask = ((G >= 32) ? 0xffffffffu : ((1u << G) - 1u)) << tbase;
  const int64_t tile = tile_sel;
  const int64_t slot = tile * G + l;  // this lane's column in every lane-interleaved array
  const LaneMap M(n, d.N, NX, K, d.bounded != 0);
  const TileMap T(n, G, mem, m, d.bounded != 0);
  const int ne = T.np / G;  // tile-vector elements owned by each lane: e = l + G j
  // row `row` of tile lane `ll` (lane-interleaved)
  const int64_t blk = slot / MPCB_LB;              // block of 32 lanes (tiles never straddle blocks: G divides 32)
  const int lane32 = (int)(slot % MPCB_LB);        // = this thread's lane: the tile groups of a warp map to one block
  constexpr int TB = MPCB_LB / G;                  // tiles per block
  double* const wsb = a.ws + blk * ((int64_t)M.total * MPCB_LB) + (lane32 - l);
  auto WL = [&](int row, int ll) { return Vec{wsb + (int64_t)row * MPCB_LB + ll, MPCB_LB}; };
  // element i of the tile-vector at offset `off` (off is a multiple of G): [block][(off + i) / G][tile in block][(off + i) % G]
  const int VR = T.vec_total / G;                  // rows of G elements in the tile-vector region
  double* const wtv = a.wst + blk * ((int64_t)VR * MPCB_LB) + (lane32 - l);
  auto tv = [&](int off, int i) -> double& { return wtv[(int64_t)((off + i) / G) * MPCB_LB + ((off + i) % G)]; };
  // plain per-tile rows behind the tile-vector region: [block][row][tile in block]
  const int PR = T.total - T.vec_total;
  double* const wtp = a.wst + (int64_t)T.vec_total * nt + blk * ((int64_t)PR * TB) + (lane32 - l) / G;
  auto TR = [&](int off, int i) -> double& { return wtp[(int64_t)(off - T.vec_total + i) * TB]; };
  // saved lane state (round mode): [block][row][lane in block]
  int32_t* const sib = ROUND ? a.si + blk * ((int64_t)MPCB_SI_ROWS * MPCB_LB) + lane32 : nullptr;
  double* const sdb = ROUND ? a.sd + blk * ((int64_t)MPCB_SD_ROWS * MPCB_LB) + lane32 : nullptr;
  auto tsync = [&]() { __syncwarp(tmask); };
  auto tsum = [&](double x) {
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) x += __shfl_xor_sync(tmask, x, o);
    return x;
  };
  auto tbcast = [&](double x, int src) { return __shfl_sync(tmask, x, tbase + src); };
  // Dot products over a tile-vector in ONE summation order for every tile width, so that the iterates do not
  // depend on G: element e belongs to class e % 8, a class is summed in increasing e, and the 8 class sums are
  // combined as ((c0 + c4) + (c2 + c6)) + ((c1 + c5) + (c3 + c7)) — the order of tsum's xor-4, xor-2, xor-1 steps
  // for G = 8, where lane l owns exactly class l.  A lane of a narrower tile owns CL = 8 / G classes (element
  // e = l + G j is class l + G (j % CL)) and folds them locally before the shuffles.
  constexpr int CL = G >= 8 ? 1 : 8 / G;
  auto tsum_c = [&](const double* pc) {
    double x;
    if (CL == 1) x = pc[0];
    else if (CL == 2) x = pc[0] + pc[1];
    else if (CL == 4) x = (pc[0] + pc[2]) + (pc[1] + pc[3]);
    else x = ((pc[0] + pc[4 % CL]) + (pc[2 % CL] + pc[6 % CL])) + ((pc[1 % CL] + pc[5 % CL]) + (pc[3 % CL] + pc[7 % CL]));
#pragma unroll
    for (int o = (G >= 8 ? 4 : G / 2); o > 0; o >>= 1) x += __shfl_xor_sync(tmask, x, o);
    return x;
  };

  // ---- tile-uniform registers ----
  int state = ST_IDLE;
  bool exhausted = false;
  int64_t b = -1;
  int cl = 0, cb = 0;  // the iterate (x, g, x̂, p) lives in lane cl's bank cb
  constexpr int H = G >= 2 ? G / 2 : G;  // candidates whose psi(x̂+) is evaluated first; lanes >= H speculate the FD gradient
  bool fd_ready = false;                 // the next iteration's FD gradient is already in lane fd_lane's P row of bank cb
  int fd_lane = 0, ge_row = 0, ge_lane = 0;
  int k = 0, no_prog = 0, lb_idx = 0, lb_full = 0, hb = 0, lsw = 0;
  int lsW = G >= 2 ? G - 1 : 1;          // tau candidates per window of the running line search (1 .. G-1)
  unsigned specmask = 0;                 // candidates (by lane) whose FD speculation has been handed out
  double psi = 0, psi_hat = 0, gTp = 0, pTp = 0, L = 0, gam = 0, phi = 0, tau0 = 0, sigma = 0, hfd = 0, hh = 0;
  double eps_tol = 0, peps = 0, ppsi = 0, norm_e2 = 0;
  int alm_i = 0, inner_total = 0, pstatus = 0, n_ifail = 0;
  int last_c = -1;  // line-search candidate that produced the current iterate: 0..7 -> tau = 2^-c, 8 = the prox step, -1 = start of a run
  unsigned long long mask0 = 0, mask1 = 0;  // inactive-set bitmask (n <= 128)
  auto mbit = [&](int j) { return j < 64 ? (mask0 >> j) & 1ull : (mask1 >> (j - 64)) & 1ull; };
  // ---- per-lane registers (this lane's candidate / evaluation request) ----
  int want = 0, n_val = 0, n_grad = 0, t_rounds = 0, dbg_fd = 0, dbg_lemma = 0, dbg_slide = 0;
  Vec evp{nullptr, 0}, evg{nullptr, 0};
  double val = 0;
  bool c_valid = false, c_pend = false, c_rq = false, c_fb = false, c_acc = false, c_started = false, fd_done = false;
  bool c_needg = false;
  int c_idx = 0, spec_for = -1;
  double psi_n = 0, psi_hat_n = 0, gTp_n = 0, pTp_n = 0, L_n = 0, gam_n = 0, phi_n = 0;

  // ---- round mode: the registers above live in global memory between transition launches ----
#define MPCB_STATE_D(X) X(psi) X(psi_hat) X(gTp) X(pTp) X(L) X(gam) X(phi) X(tau0) X(sigma) X(hfd) X(hh) X(eps_tol) X(peps) \
  X(ppsi) X(norm_e2) X(psi_n) X(psi_hat_n) X(gTp_n) X(pTp_n) X(L_n) X(gam_n) X(phi_n)
#define MPCB_STATE_I(X) X(state) X(cl) X(cb) X(fd_lane) X(ge_row) X(ge_lane) X(k) X(no_prog) X(lb_idx) X(lb_full) X(hb) X(lsw) \
  X(alm_i) X(inner_total) X(pstatus) X(n_val) X(n_grad) X(t_rounds) X(dbg_fd) X(dbg_lemma) X(dbg_slide) X(c_idx) X(spec_for) X(lsW) X(n_ifail) X(last_c)
  bool dead = notile;  // round mode: this group has no tile / its tile has no instance left
  if constexpr (ROUND) {
    if (rnd > 0 && !notile) {
      const int fl = sib[0];
      exhausted = fl & 1; fd_ready = fl & 2; c_valid = fl & 4; c_pend = fl & 8; c_rq = fl & 16; c_fb = fl & 32;
      c_acc = fl & 64; c_started = fl & 128; fd_done = fl & 256; c_needg = fl & 512;
      dead = fl & 1024;
      if (!dead) {
        int r = 1;
#define X(v) v = sib[(r++) * MPCB_LB];
        MPCB_STATE_I(X)
#undef X
        specmask = (unsigned)sib[(r++) * MPCB_LB];
        r = 0;
#define X(v) v = sdb[(r++) * MPCB_LB];
        MPCB_STATE_D(X)
#undef X
        mask0 = (unsigned long long)__double_as_longlong(sdb[(r++) * MPCB_LB]);
        mask1 = (unsigned long long)__double_as_longlong(sdb[(r++) * MPCB_LB]);
        b = __double_as_longlong(sdb[(r++) * MPCB_LB]);
        val = a.req_val[slot];
      }
    }
  }
  const bool was_dead = dead;

  // x̂ = Π_C(x − γ g), p = x̂ − x (stored into XH, P); returns gᵀp and pᵀp
  auto prox = [&](const Vec& X, const Vec& Gv, double gamma, const Vec& XH, const Vec& P, double& gtp, double& ptp) {
    // Loads of 8 elements are issued together, then the arithmetic and the stores: the compiler cannot move a
    // load above an earlier store through another Vec, which serialised n DRAM round trips per call (ncu:
    // 12 % of the transition launch's samples on this loop's first use of a loaded value).
    double s1 = 0.0, s2 = 0.0;
    for (int i0 = 0; i0 < n; i0 += 8) {
      double xv[8], gv[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const bool in = i0 + k < n;
        xv[k] = in ? X(i0 + k) : 0.0;
        gv[k] = in ? Gv(i0 + k) : 0.0;
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int i = i0 + k;
        if (i < n) {
          double t = fma(-gamma, gv[k], xv[k]);
          t = t < lbv[i] ? lbv[i] : (t > ubv[i] ? ubv[i] : t);
          double pi = t - xv[k];
          XH.set(i, t);
          P.set(i, pi);
          s1 = fma(gv[k], pi, s1);
          s2 = fma(pi, pi, s2);
        }
      }
    }
    gtp = s1;
    ptp = s2;
  };

  // q <- H q on the inactive set (masked two-loop recursion, Appendix A 3c), split over the tile: lane l owns the
  // elements e = l + G j.  With at most NQ of them per lane (G = 8: n <= 80) q stays in registers for the whole
  // recursion — the global read-modify-write of q per history pair (and the store/load ordering it imposed)
  // was a third of the transition launch's samples — and the dot products run in the same order as before.
  constexpr int NQ = 10;
  auto lbfgs_apply = [&]() {
    const int cnt = lb_full ? mem : lb_idx;
    double H0 = -1.0;
    // The history of a large batch lives in DRAM (13 KB per tile): ask for all of it at once, so the
    // dependent chain below (every step needs the q of the previous one) waits on L2 hits instead.
    for (int kk = 0; kk < cnt; ++kk) {
      int i = lb_idx - 1 - kk;
      if (i < 0) i += mem;
      for (int j = 0; j < ne; j += (G >= 16 ? 1 : 16 / G)) {  // one request per 128-byte line of the tile's row
        int e = l + G * j;
        if (e < n) {
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&tv(T.LS + i * T.np, e)));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(&tv(T.LY + i * T.np, e)));
        }
      }
    }
    if (ROUND && ne <= NQ) {  // (the persistent kernel has no registers to spare: it spilled with this path)
      // The rows of a pair are loaded UNCONDITIONALLY, all at once (padding lanes read the row's padding, elements outside
      // the inactive set are read and not used): written as `own ? load : 0` every load sat behind its own branch next
      // to its use and the ten of them took ten L2 round trips, 2 000 cycles per pair and loop on a lone tile (B200,
      // clock64 around the statements) — 46 us of an 80 us transition launch at 20 pairs.
      auto run = [&](auto nqc) {
        constexpr int NQc = decltype(nqc)::value;
        double q[NQc];
        bool own[NQc];
#pragma unroll
        for (int j = 0; j < NQc; ++j) {
          const int e = l + G * j;
          own[j] = j < ne && e < n && mbit(e);
          q[j] = tv(T.Q, l + G * (j < ne ? j : ne - 1));
        }
#pragma unroll
        for (int j = 0; j < NQc; ++j) q[j] = own[j] ? q[j] : 0.0;
        auto load_pair = [&](int i, double* sv, double* yv) {
#pragma unroll
          for (int j = 0; j < NQc; ++j) {
            const int e = l + G * (j < ne ? j : ne - 1);
            sv[j] = tv(T.LS + i * T.np, e);
            yv[j] = tv(T.LY + i * T.np, e);
          }
        };
        for (int kk = 0; kk < cnt; ++kk) {
          int i = lb_idx - 1 - kk;
          if (i < 0) i += mem;
          double sv[NQc], yv[NQc];
          load_pair(i, sv, yv);
          double sy = 0.0, yy = 0.0, sq = 0.0;
#pragma unroll
          for (int j = 0; j < NQc; ++j)
            if (own[j]) { sy = fma(sv[j], yv[j], sy); yy = fma(yv[j], yv[j], yy); sq = fma(sv[j], q[j], sq); }
          sy = tsum(sy); yy = tsum(yy); sq = tsum(sq);
          if (!(sy > 0.0)) { s_ra[i] = nan(""); continue; }
          double rho = 1.0 / sy;
          s_ra[i] = rho;
          if (H0 < 0.0) H0 = sy / yy;
          double al = rho * sq;
          s_ra[mem + i] = al;
#pragma unroll
          for (int j = 0; j < NQc; ++j)
            if (own[j]) q[j] = fma(-al, yv[j], q[j]);
        }
        if (!(H0 >= 0.0)) H0 = gam;
#pragma unroll
        for (int j = 0; j < NQc; ++j)
          if (own[j]) q[j] *= H0;
        for (int kk = cnt - 1; kk >= 0; --kk) {
          int i = lb_idx - 1 - kk;
          if (i < 0) i += mem;
          double rho = s_ra[i];
          if (isnan(rho)) continue;
          double sv[NQc], yv[NQc];
          load_pair(i, sv, yv);
          double yq = 0.0;
#pragma unroll
          for (int j = 0; j < NQc; ++j)
            if (own[j]) yq = fma(yv[j], q[j], yq);
          yq = tsum(yq);
          double coef = s_ra[mem + i] - rho * yq;
#pragma unroll
          for (int j = 0; j < NQc; ++j)
            if (own[j]) q[j] = fma(coef, sv[j], q[j]);
        }
#pragma unroll
        for (int j = 0; j < NQc; ++j)
          if (own[j]) tv(T.Q, l + G * j) = q[j];
      };
      if (ne <= 5) run(std::integral_constant<int, 5>{});
      else run(std::integral_constant<int, NQ>{});
      return;
    }
    for (int kk = 0; kk < cnt; ++kk) {
      int i = lb_idx - 1 - kk;
      if (i < 0) i += mem;
      double syc[CL], yyc[CL], sqc[CL];
#pragma unroll
      for (int t = 0; t < CL; ++t) { syc[t] = 0.0; yyc[t] = 0.0; sqc[t] = 0.0; }
      for (int j0 = 0; j0 < ne; j0 += CL) {
#pragma unroll
        for (int t = 0; t < CL; ++t) {
          const int e = l + G * (j0 + t);
          if (j0 + t < ne && e < n && mbit(e)) {
            double s = tv(T.LS + i * T.np, e), y = tv(T.LY + i * T.np, e);
            syc[t] = fma(s, y, syc[t]); yyc[t] = fma(y, y, yyc[t]); sqc[t] = fma(s, tv(T.Q, e), sqc[t]);
          }
        }
      }
      const double sy = tsum_c(syc), yy = tsum_c(yyc), sq = tsum_c(sqc);
      if (!(sy > 0.0)) { s_ra[i] = nan(""); continue; }
      double rho = 1.0 / sy;
      s_ra[i] = rho;
      if (H0 < 0.0) H0 = sy / yy;
      double al = rho * sq;
      s_ra[mem + i] = al;
      for (int j = 0; j < ne; ++j) {
        int e = l + G * j;
        if (e < n && mbit(e)) tv(T.Q, e) = fma(-al, tv(T.LY + i * T.np, e), tv(T.Q, e));
      }
    }
    if (!(H0 >= 0.0)) H0 = gam;
    for (int j = 0; j < ne; ++j) {
      int e = l + G * j;
      if (e < n && mbit(e)) tv(T.Q, e) *= H0;
    }
    for (int kk = cnt - 1; kk >= 0; --kk) {
      int i = lb_idx - 1 - kk;
      if (i < 0) i += mem;
      double rho = s_ra[i];
      if (isnan(rho)) continue;
      double yqc[CL];
#pragma unroll
      for (int t = 0; t < CL; ++t) yqc[t] = 0.0;
      for (int j0 = 0; j0 < ne; j0 += CL) {
#pragma unroll
        for (int t = 0; t < CL; ++t) {
          const int e = l + G * (j0 + t);
          if (j0 + t < ne && e < n && mbit(e)) yqc[t] = fma(tv(T.LY + i * T.np, e), tv(T.Q, e), yqc[t]);
        }
      }
      const double yq = tsum_c(yqc);
      double coef = s_ra[mem + i] - rho * yq;
      for (int j = 0; j < ne; ++j) {
        int e = l + G * j;
        if (e < n && mbit(e)) tv(T.Q, e) = fma(coef, tv(T.LS + i * T.np, e), tv(T.Q, e));
      }
    }
  };

  // Line-search candidates (Appendix A 3d): c = 0..7 -> tau = 2^-c, c = 8 -> the prox step x̂ (forced
  // accept).  Lane order is trial order: the last lane of the tile holds c = 8, the lanes below it hold a
  // window of lsW consecutive tau candidates (1 <= lsW <= G-1), right-aligned, that slides when all of its
  // candidates are rejected.  The window width is the amount of SPECULATION, never a decision: candidates are
  // decided in trial order whatever the width, so every width gives the same iterates (tests: W = 1 vs W = 7
  // bit-identical).  Wide (lsW = G-1, picked when the GPU has lanes to spare): all trials of Appendix A 3d in two
  // windows and the prox step evaluated from the start (long runs of iterations in which every tau fails are what
  // the slowest instances of a batch are made of).  Narrow (lsW = 1, the bulk of a large batch, where every
  // evaluation costs throughput): one trial at a time like sequential back-tracking, the prox step only once every
  // tau has failed (`fb_now`).  G = 1: one candidate at a time, c = lsw.  tau0 = 0 (k = 0 or a non-finite q):
  // only the prox step.  Newly assigned candidates request psi and grad psi.
  auto ls_assign = [&](bool first, bool fb_now) {
    const int nb = cb ^ 1;
    bool mine = false, valid = false;
    if (G == 1) {
      c_idx = tau0 == 0.0 ? 8 : lsw;
      mine = true; valid = true;
    } else if (l == G - 1) {
      if (fb_now || (first && (lsW == G - 1 || tau0 == 0.0))) { c_idx = 8; mine = true; valid = true; }
      else if (first) { c_idx = 8; mine = true; valid = false; }  // narrow window: handed out later (fb_now)
    } else if (!fb_now) {
      const int c_first = lsw * lsW;
      const int cnt = 8 - c_first < lsW ? 8 - c_first : lsW;
      const int j = l - (G - 1 - cnt);
      c_idx = c_first + j;
      mine = true; valid = tau0 != 0.0 && j >= 0;
    }
    if (!mine) return;
    c_fb = c_idx == 8;
    c_valid = valid; c_pend = valid; c_needg = valid;
    c_rq = false; c_acc = false; c_started = false;
    if (valid) {
      Vec XN = WL(M.X(nb), l);
      if (c_fb) {
        Vec XH = WL(M.XH(cb), cl);
        for (int i0 = 0; i0 < n; i0 += 8) {
          double v[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) v[k] = i0 + k < n ? XH(i0 + k) : 0.0;
#pragma unroll
          for (int k = 0; k < 8; ++k) if (i0 + k < n) XN.set(i0 + k, v[k]);
        }
      } else {
        const double tau = ldexp(1.0, -c_idx);
        Vec X = WL(M.X(cb), cl), P = WL(M.P(cb), cl);
        for (int i0 = 0; i0 < n; i0 += 8) {  // loads first, stores after (see prox)
          double xv[8], pv[8], qv[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const bool in = i0 + k < n;
            xv[k] = in ? X(i0 + k) : 0.0;
            pv[k] = in ? P(i0 + k) : 0.0;
            qv[k] = in ? tv(T.Q, i0 + k) : 0.0;
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) if (i0 + k < n) XN.set(i0 + k, fma(tau, qv[k], fma(1.0 - tau, pv[k], xv[k])));
        }
      }
      evp = XN; evg = WL(M.G(nb), l); want = 2;
    }
  };

  // write the instance's results (from the backup x when the solve failed numerically)
  auto finish_instance = [&](int fin_status, bool from_backup) {
    int nv = n_val, ng = n_grad;
#pragma unroll
    for (int o = G / 2; o > 0; o >>= 1) { nv += __shfl_xor_sync(tmask, nv, o); ng += __shfl_xor_sync(tmask, ng, o); }
    if (l == 0) {
      Vec X0 = WL(M.X(0), 0);
      for (int i = 0; i < n; ++i) a.U[b * n + i] = from_backup ? tv(T.XB, i) : X0(i);
      a.status[b] = fin_status;
      a.iters[2 * b] = alm_i + 1; a.iters[2 * b + 1] = inner_total;
      a.evals[2 * b] = nv; a.evals[2 * b + 1] = ng;
      // MPCB_DEBUG_TIMERS=3: psi carries the instance's evaluation rounds + 1e4 non-speculated FD rounds + 1e8 descent-lemma
      // re-evaluations at x+ + 1e12 window slides
      a.psi[b] = d.debug >= 3 ? (double)t_rounds + 1e4 * dbg_fd + 1e8 * dbg_lemma + 1e12 * dbg_slide : ppsi;
      a.res[b] = peps;
      if (a.gam_out) a.gam_out[b] = gam;
      if (a.fin_ns) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        a.fin_ns[b] = t;
      }
      if (a.ifail_out) a.ifail_out[b] = n_ifail;
    }
    n_val = 0; n_grad = 0; t_rounds = 0; dbg_fd = 0; dbg_lemma = 0; dbg_slide = 0;
  };

  long long t_eval = 0, t_trans = 0, n_rounds = 0, n_grounds = 0, t_mark = clock64();
  for (;;) {
    // ================= transition phase: run the tile's state machine up to its next evaluation ====
    want = 0;
    for (; !dead;) {
      if (state == ST_IDLE) {
        if (exhausted) break;
        long long bb = 0;
        if (l == 0) bb = (long long)atomicAdd(a.queue, 1ull);
        bb = __shfl_sync(tmask, bb, tbase);
        if (bb >= a.B) { exhausted = true; break; }
        b = bb;
        // controller.py:20-21,57: warm start (U, λ) comes in through U_inout / lambda_inout
        cl = 0; cb = 0;
        if (l == 0) {
          Vec X = WL(M.X(0), 0);
          for (int i = 0; i < n; ++i) X.set(i, a.U[b * n + i]);
        }
        tsync();
        eps_tol = sp.eps0;
        alm_i = 0; inner_total = 0; n_val = 0; n_grad = 0; n_ifail = 0;
        if (d.bounded) {
          for (int r = 0; r < m; ++r) { TR(T.SIG, r) = sp.Sigma0; TR(T.E1, r) = 0.0; }
        }
        state = ST_START_PANOC;
      } else if (state == ST_START_PANOC) {
        if (d.bounded) {  // clamp multipliers to ±M, 0 on infinite sides
          if (l == 0) {
            for (int r = 0; r < m; ++r) {
              const int ri = r % d.rows;  // state rows: their box; ellipse rows (MPCB_FLAG_SAFE_DIST): D = (-inf, 0]
              double lo = ri < NX ? (isinf(d.D_lo[ri]) ? 0.0 : -sp.M) : 0.0, hi = ri < NX ? (isinf(d.D_hi[ri]) ? 0.0 : sp.M) : sp.M;
              double lv = a.lam[b * m + r];
              a.lam[b * m + r] = lv < lo ? lo : (lv > hi ? hi : lv);
            }
          }
          tsync();
        }
        Vec X = WL(M.X(cb), cl);
        // the Lipschitz probe point x + h goes to the probing lane's free bank
        const int pl = G >= 2 ? 1 : 0;
        Vec XE = WL(M.X(cb ^ 1), pl);
        hh = 0.0;
        for (int i = 0; i < n; ++i) {
          double xi = X(i);
          tv(T.XB, i) = xi;
          double h = fmax(fabs(xi) * sp.lip_eps, sp.lip_delta);
          if (l == pl) XE.set(i, xi + h);
          hh = fma(h, h, hh);
        }
        k = 0; no_prog = 0; lb_idx = 0; lb_full = 0; fd_ready = false;
        if (G >= 2) {
          if (l == 0) { evp = X; evg = WL(M.G(cb), cl); want = 2; }
          else if (l == 1) { evp = XE; evg = WL(M.G(cb ^ 1), 1); want = 2; }
          state = ST_LIP_DONE;
        } else {
          evp = X; evg = WL(M.G(cb), cl); want = 2;
          state = ST_INIT_G1;
        }
        break;
      } else if (state == ST_INIT_G1) {  // G == 1 only: second gradient of the Lipschitz estimate
        psi = val;
        evp = WL(M.X(cb ^ 1), 0); evg = WL(M.G(cb ^ 1), 0); want = 2;
        state = ST_LIP_DONE;
        break;
      } else if (state == ST_LIP_DONE) {
        if (G >= 2) psi = tbcast(val, 0);
        const int pl = G >= 2 ? 1 : 0;
        Vec Gc = WL(M.G(cb), cl), GE = WL(M.G(cb ^ 1), pl);
        double dd = 0.0;
        for (int i = 0; i < n; ++i) { double t = GE(i) - Gc(i); dd = fma(t, t, dd); }
        L = sqrt(dd) / sqrt(hh);
        if (!isfinite(L)) { pstatus = MPCB_STATUS_NOT_FINITE; peps = nan(""); ppsi = psi; state = ST_PANOC_EXIT; continue; }
        L = L < sp.L_min ? sp.L_min : (L > sp.L_max ? sp.L_max : L);
        gam = sp.Lgamma_factor / L;
        hb = 0;
        state = ST_XHAT_REQ;
      } else if (state == ST_XHAT_REQ) {
        // descent lemma at k = 0 (Appendix A 3a): candidate c has halved gamma c times
        const int c = hb * G + l;
        L_n = ldexp(L, c); gam_n = ldexp(gam, -c);
        c_valid = c == 0 || L_n <= sp.L_max;
        if (c_valid) {
          prox(WL(M.X(cb), cl), WL(M.G(cb), cl), gam_n, WL(M.XH(cb ^ 1), l), WL(M.P(cb ^ 1), l), gTp_n, pTp_n);
          evp = WL(M.XH(cb ^ 1), l); want = 1;
        }
        state = ST_XHAT_DONE;
        break;
      } else if (state == ST_XHAT_DONE) {
        bool done = false;
        if (c_valid) {
          psi_hat_n = val;
          double margin = (1.0 + fabs(psi)) * sp.quad_tol;
          bool violated = psi_hat_n - psi > gTp_n + 0.5 * L_n * pTp_n + margin && L_n * 2.0 <= sp.L_max;
          done = !violated;
        }
        const unsigned dm = (__ballot_sync(tmask, done) & tmask) >> tbase;
        if (!dm) { ++hb; state = ST_XHAT_REQ; continue; }
        const int jw = __ffs(dm) - 1;
        L = tbcast(L_n, jw); gam = tbcast(gam_n, jw);
        gTp = tbcast(gTp_n, jw); pTp = tbcast(pTp_n, jw);
        psi_hat = tbcast(psi_hat_n, jw);
        if (l == jw) {  // the winner's x̂, p become the iterate's
          Vec XHs = WL(M.XH(cb ^ 1), l), Ps = WL(M.P(cb ^ 1), l), XHd = WL(M.XH(cb), cl), Pd = WL(M.P(cb), cl);
          for (int i = 0; i < n; ++i) { XHd.set(i, XHs(i)); Pd.set(i, Ps(i)); }
        }
        tsync();
        phi = psi + pTp / (2.0 * gam) + gTp;
        state = ST_ITER_TOP;
      } else if (state == ST_ITER_TOP) {
        if (a.trace && l == 0 && inner_total + k < a.trace_cap) {  // (psi, FBE phi_gamma, gamma, accepted candidate) of iterate k
          double* tr = a.trace + ((int64_t)b * a.trace_cap + inner_total + k) * 4;
          tr[0] = psi; tr[1] = phi; tr[2] = gam; tr[3] = (double)(k == 0 ? -1 : last_c);
        }
        double eps_k = sqrt(pTp);
        int stop = 0;
        if (!isfinite(eps_k)) stop = MPCB_STATUS_NOT_FINITE;
        else if (eps_k <= eps_tol) stop = MPCB_STATUS_CONVERGED;
        else if (k == sp.max_iter) stop = MPCB_STATUS_MAX_ITER;
        else if (no_prog > sp.max_no_progress) stop = MPCB_STATUS_NO_PROGRESS;
        if (stop) { pstatus = stop; peps = eps_k; ppsi = psi_hat; state = ST_PANOC_EXIT; continue; }
        if (k == 0) { state = ST_LS_START; continue; }
        // direction (Appendix A 3c)
        Vec X = WL(M.X(cb), cl), Gc = WL(M.G(cb), cl), P = WL(M.P(cb), cl);
        int nJ = 0;
        mask0 = 0; mask1 = 0;
        double xx = 0.0;
        for (int i = 0; i < n; ++i) {
          double xi = X(i);
          double gd = fma(-gam, Gc(i), xi);
          bool active = gd < lbv[i] || ubv[i] < gd;
          if (!active) { ++nJ; if (i < 64) mask0 |= 1ull << i; else mask1 |= 1ull << (i - 64); }
          xx = fma(xi, xi, xx);
        }
        const bool use_fd = sp.hess_vec_fd && nJ > 0 && nJ < n;
        for (int j = 0; j < ne; ++j) {
          int e = l + G * j;
          if (e < n) tv(T.Q, e) = mbit(e) ? (use_fd ? 0.0 : -Gc(e)) : P(e);
        }
        if (use_fd) {
          hfd = 6.0554544523933395e-06 /* cbrt(DBL_EPSILON) */ * (1.0 + sqrt(xx));
          if (fd_ready) {  // evaluated speculatively during the previous line search (same point, bit for bit)
            fd_ready = false;
            ge_row = M.P(cb); ge_lane = fd_lane;
            state = ST_FD_DONE;
            continue;
          }
          if (l == 0) {  // q is p on the active set and 0 on the inactive set at this point
            Vec XE = WL(M.X(cb ^ 1), 0);
            for (int i = 0; i < n; ++i) XE.set(i, fma(hfd, mbit(i) ? 0.0 : P(i), X(i)));
            evp = XE; evg = WL(M.G(cb ^ 1), 0); want = 2;
          }
          ge_row = M.G(cb ^ 1); ge_lane = 0;
          ++dbg_fd;
          state = ST_FD_DONE;
          break;
        }
        fd_ready = false;
        if (nJ > 0) lbfgs_apply();
        tsync();  // every lane reads the whole q next
        state = ST_LS_START;
      } else if (state == ST_FD_DONE) {
        Vec Gc = WL(M.G(cb), cl), GE = WL(ge_row, ge_lane);
        for (int j = 0; j < ne; ++j) {
          int e = l + G * j;
          if (e < n && mbit(e)) { double gi = Gc(e); tv(T.Q, e) = -gi - (GE(e) - gi) / hfd; }
        }
        lbfgs_apply();
        tsync();
        state = ST_LS_START;
      } else if (state == ST_LS_START) {
        tau0 = 0.0;
        if (k > 0) {
          tau0 = 1.0;
          bool fin = true;
          for (int i = 0; i < n; ++i) fin = fin && isfinite(tv(T.Q, i));
          if (!fin) { tau0 = 0.0; lb_idx = 0; lb_full = 0; }
        }
        sigma = (1.0 - gam * L) * pTp / (2.0 * gam);
        lsw = 0; specmask = 0;
        fd_done = false; spec_for = -1;
        // speculation width of this line search (constant until its end)
        if (G >= 2) {
          int w = a.spec_w;
          if (w <= 0) {
            // device policy: a tile in a wide window costs ~8 evaluations per round; while more tiles are solving
            // than the evaluation launch has lanes for, every one of them costs throughput, so trials go one at a time
            const unsigned live = a.ctl ? a.ctl[0] : 0u;
            w = live <= (unsigned)a.spec_live_wide ? G - 1 : (live <= (unsigned)a.spec_live_mid ? 3 : 1);
          }
          lsW = w > G - 1 ? G - 1 : w;
        }
        ls_assign(true, false);
        state = ST_LS_STEP;
        break;
      } else if (state == ST_LS_STEP) {
        // One decision step of the line search (Appendix A 3d), run after every evaluation round.
        const int nb = cb ^ 1;
        const double thr = (1.0 + fabs(phi)) * sp.ls_tol;
        // phi_gamma(x+) = psi + min_z {g'(z - x+) + |z - x+|^2 / (2 gamma)} can only grow when the descent
        // lemma shrinks gamma, so a candidate that fails the line-search test with its current gamma fails
        // it for every smaller gamma: it is rejected at once, without evaluating psi(x̂+) or finishing its
        // descent-lemma loop — the decision sequential back-tracking reaches at the end of that loop.  The
        // guard covers the rounding of the two gamma-dependent terms (40-term dot products) and of the
        // two additions; inside it the candidate is evaluated in full.  NaN fails (as in the sequential test).
        auto ls_hopeless = [&]() {
          if (d.exact) return false;
          const double s2 = pTp_n / (2.0 * gam_n);
          const double lc = psi_n + s2 + gTp_n - (phi - sigma);
          const double guard = 8.9e-16 * (1.0 + fabs(phi) + fabs(psi_n)) + 1e-12 * (fabs(gTp_n) + s2);
          return !(lc <= thr + guard);
        };
        // After a prox step at (L_n, gam_n): consume every descent-lemma step whose outcome needs no
        // evaluation.  With psi >= 0 (KDesc.psi_nonneg) the lemma test psi(x̂+) - psi+ > g'p + L/2 |p|^2 + margin
        // is certainly violated when it is violated for psi(x̂+) = 0 — a back-tracking trial that landed
        // where psi is huge halves gamma a dozen times before anything else can happen, and none of those
        // halvings depends on psi(x̂+).  (Sequential back-tracking evaluates psi(x̂+) each time and reaches
        // the same (L_n, gam_n); the one difference: a NaN psi(x̂+) would have stopped its loop.)
        auto ls_settle = [&]() {
          for (;;) {
            if (!c_fb && ls_hopeless()) { c_pend = false; c_acc = false; return; }
            const double margin = (1.0 + fabs(psi_n)) * sp.quad_tol;
            if (d.exact || !(d.psi_nonneg && 0.0 - psi_n > gTp_n + 0.5 * L_n * pTp_n + margin && L_n * 2.0 <= sp.L_max)) return;
            L_n *= 2.0; gam_n *= 0.5;
            prox(WL(M.X(nb), l), WL(M.G(nb), l), gam_n, WL(M.XH(nb), l), WL(M.P(nb), l), gTp_n, pTp_n);
          }
        };
        if (c_valid && c_pend) {
          if (c_needg) {  // psi, grad psi of this lane's candidate arrived: x̂+, p+ and the tentative phi+
            c_needg = false;
            psi_n = val; L_n = L; gam_n = gam;
            prox(WL(M.X(nb), l), WL(M.G(nb), l), gam_n, WL(M.XH(nb), l), WL(M.P(nb), l), gTp_n, pTp_n);
            c_started = true; fd_done = false;
            ls_settle();
          } else if (c_rq) {  // psi(x̂+) arrived: descent lemma at x+, then the line-search test
            psi_hat_n = val;
            const double margin = (1.0 + fabs(psi_n)) * sp.quad_tol;
            if (psi_hat_n - psi_n > gTp_n + 0.5 * L_n * pTp_n + margin && L_n * 2.0 <= sp.L_max) {
              L_n *= 2.0; gam_n *= 0.5;  // shrink this candidate's step and re-evaluate
              ++dbg_lemma;
              prox(WL(M.X(nb), l), WL(M.G(nb), l), gam_n, WL(M.XH(nb), l), WL(M.P(nb), l), gTp_n, pTp_n);
              fd_done = false;
              ls_settle();
            } else {
              c_pend = false;
              phi_n = psi_n + pTp_n / (2.0 * gam_n) + gTp_n;
              // NaN-safe: a non-finite phi+ is a failed trial (keeps back-tracking to the prox step)
              c_acc = c_fb || (phi_n - (phi - sigma) <= thr);
            }
          }
        }
        c_rq = false;
        unsigned pm = (__ballot_sync(tmask, c_valid && c_pend) & tmask) >> tbase;
        const unsigned am = (__ballot_sync(tmask, c_valid && !c_pend && c_acc) & tmask) >> tbase;
        const int first_acc = am ? __ffs(am) - 1 : G;
        const unsigned below = (1u << first_acc) - 1u;  // first_acc <= G <= 8
        // the window of tau candidates slides once all of its lanes are rejected (the prox step on the last
        // lane cannot be accepted before every tau has been tried)
        const unsigned wmask = (1u << (G - 1)) - 1u;
        const bool wdead = G >= 2 && !((pm | am) & wmask);  // every tau candidate of the window has been rejected
        const bool slide = wdead && tau0 != 0.0 && (lsw + 1) * lsW <= 7;
        // narrow windows hand out the prox-step candidate only now, when no tau is left to try
        const bool fb_live = G >= 2 && ((__ballot_sync(tmask, c_valid && c_fb) & tmask) != 0u);
        const bool need_fb = wdead && !slide && !fb_live;
        if (pm & below || first_acc == G || slide) {
          // earlier candidates are undecided (or, G = 1, the only candidate was rejected): another round
          if (G == 1) {
            if (!c_pend) { ++lsw; ls_assign(false, false); }
          } else {
            if (need_fb) {
              ls_assign(false, true);
              pm = (__ballot_sync(tmask, c_valid && c_pend) & tmask) >> tbase;
            }
            if (slide) {
              ++lsw; ++dbg_slide;
              ls_assign(false, false);
              // speculations made for the window's previous candidates are void
              if (spec_for >= 0 && spec_for < G - 1) { fd_done = false; spec_for = -1; }
              specmask &= ~wmask;
              pm = (__ballot_sync(tmask, c_valid && c_pend) & tmask) >> tbase;
            }
          }
          const unsigned needg = (__ballot_sync(tmask, c_valid && c_pend && c_needg) & tmask) >> tbase;
          // primary candidates: the first H undecided ones evaluate psi(x̂+) in this round
          unsigned primary = 0;
          {
            unsigned cand = pm & below & ~needg;
            for (int r = 0; r < H && cand; ++r) { unsigned low = cand & (0u - cand); primary |= low; cand ^= low; }
          }
          if ((primary >> l) & 1u) {
            if (!c_started) {  // deferred candidate (its rows were lent to an FD speculation)
              prox(WL(M.X(nb), l), WL(M.G(nb), l), gam_n, WL(M.XH(nb), l), WL(M.P(nb), l), gTp_n, pTp_n);
              c_started = true; fd_done = false;
            }
            evp = WL(M.XH(nb), l); want = 1; c_rq = true;
          }
          if (G >= 2 && sp.hess_vec_fd) {
            // Speculate the next iteration's FD gradient for primary candidates on lanes with nothing to
            // evaluate in this round, exactly as ST_ITER_TOP would form it if the candidate is accepted
            // with its step size unchanged: the r-th free lane serves the r-th primary candidate.
            const unsigned holds = (__ballot_sync(tmask, fd_done) & tmask) >> tbase;
            const unsigned tbits = (G >= 32) ? 0xffffffffu : ((1u << G) - 1u);
            unsigned fr = tbits & ~(primary | needg | holds | am);  // decided-and-waiting lanes keep their x̂+, p+
            unsigned ts = primary & ~specmask;
            int target = -1;
            while (fr && ts) {
              const int fl = __ffs(fr) - 1, tl = __ffs(ts) - 1;
              if (fl == l) target = tl;
              fr &= fr - 1; ts &= ts - 1;
              specmask |= 1u << tl;
            }
            if (target >= 0) {
              Vec Xj = WL(M.X(nb), target), Gj = WL(M.G(nb), target);
              int nJ = 0;
              double xx = 0.0;
              for (int i = 0; i < n; ++i) {
                double xi = Xj(i);
                double gd = fma(-gam, Gj(i), xi);
                if (!(gd < lbv[i] || ubv[i] < gd)) ++nJ;
                xx = fma(xi, xi, xx);
              }
              if (nJ > 0 && nJ < n) {
                const double hf = 6.0554544523933395e-06 * (1.0 + sqrt(xx));
                Vec XE = WL(M.XH(nb), l);
                for (int i0 = 0; i0 < n; i0 += 8) {  // loads first, stores after (see prox)
                  double xv[8], gv[8];
#pragma unroll
                  for (int k = 0; k < 8; ++k) {
                    const bool in = i0 + k < n;
                    xv[k] = in ? Xj(i0 + k) : 0.0;
                    gv[k] = in ? Gj(i0 + k) : 0.0;
                  }
#pragma unroll
                  for (int k = 0; k < 8; ++k) {
                    const int i = i0 + k;
                    if (i < n) {
                      double xi = xv[k];
                      double gd = fma(-gam, gv[k], xi);
                      bool active = gd < lbv[i] || ubv[i] < gd;
                      double t = gd < lbv[i] ? lbv[i] : (gd > ubv[i] ? ubv[i] : gd);
                      XE.set(i, fma(hf, active ? t - xi : 0.0, xi));
                    }
                  }
                }
                evp = XE; evg = WL(M.P(nb), l); want = 2; fd_done = true; spec_for = target;
                c_started = false;  // this lane's own x̂+, p+ rows now hold the speculation
              }
            }
          }
          break;
        }
        // accept candidate jw: L-BFGS update with s = x⁺ − x, y = ∇ψ⁺ − ∇ψ (Appendix A 3e)
        const int jw = first_acc;
        {
          Vec X = WL(M.X(cb), cl), XN = WL(M.X(nb), jw), Gc = WL(M.G(cb), cl), GN = WL(M.G(nb), jw);
          // The pair enters the history only if it is accepted (Appendix A 3e; alpaqa stores accepted pairs only):
          // with a full buffer slot lb_idx holds the OLDEST live pair, which a rejected one must not destroy.
          double syc[CL], ssc[CL];
#pragma unroll
          for (int t = 0; t < CL; ++t) { syc[t] = 0.0; ssc[t] = 0.0; }
          bool same = true;
          for (int j0 = 0; j0 < ne; j0 += CL) {
#pragma unroll
            for (int t = 0; t < CL; ++t) {
              const int e = l + G * (j0 + t);
              if (j0 + t < ne && e < n) {
                double xi = X(e), xni = XN(e);
                double s = xni - xi, y = GN(e) - Gc(e);
                syc[t] = fma(s, y, syc[t]);
                ssc[t] = fma(s, s, ssc[t]);
                same = same && (xi == xni);
              }
            }
          }
          const double sy = tsum_c(syc), ss = tsum_c(ssc);
          same = (__ballot_sync(tmask, !same) & tmask) == 0;
          const double tiny = 1.4916681462400413e-154;  // sqrt(DBL_MIN)
          if (isfinite(sy) && ss >= tiny && sy >= tiny) {
            for (int j = 0; j < ne; ++j) {
              const int e = l + G * j;
              if (e < n) {
                tv(T.LS + lb_idx * T.np, e) = XN(e) - X(e);
                tv(T.LY + lb_idx * T.np, e) = GN(e) - Gc(e);
              }
            }
            if (++lb_idx == mem) { lb_idx = 0; lb_full = 1; }
          }
          if (no_prog > 0 || k % sp.max_no_progress == 0) no_prog = same ? no_prog + 1 : 0;
        }
        {
          const unsigned fdm = (__ballot_sync(tmask, fd_done && spec_for == jw) & tmask) >> tbase;
          const double gam_w = tbcast(gam_n, jw);
          fd_ready = G >= 2 && fdm != 0 && gam_w == gam;
          fd_lane = fdm ? __ffs(fdm) - 1 : 0;
        }
        if (d.debug >= 16 && b == d.debug - 16) {  // MPCB_DEBUG_TIMERS=16+i: trace of instance i
          const int cw = __shfl_sync(tmask, c_idx, tbase + jw);
          const double gw = tbcast(gam_n, jw), pw = tbcast(pTp_n, jw);
          if (l == 0)
            printf("[trace] outer %d k %d accepted c=%d gamma %.17g pTp %.17g psi %.17g phi %.17g lbfgs %d/%d\n", alm_i, k, cw, gw, pw,
                   tbcast(psi_n, jw), tbcast(phi_n, jw), lb_idx, lb_full);
        }
        last_c = __shfl_sync(tmask, c_idx, tbase + jw);
        psi = tbcast(psi_n, jw); psi_hat = tbcast(psi_hat_n, jw);
        gTp = tbcast(gTp_n, jw); pTp = tbcast(pTp_n, jw);
        L = tbcast(L_n, jw); gam = tbcast(gam_n, jw);
        phi = tbcast(phi_n, jw);
        cl = jw; cb = nb;
        tsync();
        ++k;
        state = ST_ITER_TOP;
      } else if (state == ST_PANOC_EXIT) {
        inner_total += k;
        if (pstatus == MPCB_STATUS_NOT_FINITE) {  // failed instances keep their warm start
          finish_instance(MPCB_STATUS_NOT_FINITE, true);
          state = ST_IDLE;
          continue;
        }
        // return x̂_k: it becomes lane 0 / bank 0's x (the next PANOC run of the ALM loop starts there)
        if (l == 0) {
          Vec XH = WL(M.XH(cb), cl), X0 = WL(M.X(0), 0);
          for (int i = 0; i < n; ++i) X0.set(i, XH(i));
        }
        cl = 0; cb = 0;
        tsync();
        norm_e2 = 0.0;
        if (d.bounded) {
          if (l == 0) { evp = WL(M.X(0), 0); want = 3; }
          state = ST_ALM_V;
          break;
        }
        state = ST_ALM_UPDATE;
      } else if (state == ST_ALM_V) {
        // err_z = g(x̂) − Π_D(ζ), λ ← ŷ(x̂); row r = (stage, state component | obstacle)
        Vec STt = WL(M.ST, 0);
        auto row_at = [&](int r, double* yh, double* ez) {
          const int R = d.rows, nn = r / R, ri = r - nn * R;
          double g, lo, hi;
          if (ri < NX) {
            const double y = STt(nn * NX + ri);
            g = y * y - d.g_off[ri]; lo = d.D_lo[ri]; hi = d.D_hi[ri];
          } else {
            double yy[NX];
#pragma unroll
            for (int j = 0; j < NX; ++j) yy[j] = STt(nn * NX + j);
            const double* ob = a.obs + (b * d.n_obs + (ri - NX)) * MPCB_OBS_DIM;
            const double t = (double)(nn + 1) * d.Ts;
            double so, co;
            sincos(ob[2], &so, &co);
            const double sdist = ob[3] * t + 0.5 * ob[4] * t * t;  // event_triggered.py:26
            g = safe_dist_row<MODEL>(d, yy, ob[0] + sdist * co, ob[1] + sdist * so, ob[2], ob[3] + ob[4] * t, 0.0, nullptr);
            lo = __longlong_as_double(0xfff0000000000000LL); hi = 0.0;
          }
          alm_row(g, a.lam[b * m + r], TR(T.SIG, r), lo, hi, yh, ez);
        };
        norm_e2 = 0.0;
        for (int r = 0; r < m; ++r) {
          double yh, ez;
          row_at(r, &yh, &ez);
          TR(T.E2, r) = ez;
          norm_e2 = fmax(norm_e2, fabs(ez));
        }
        tsync();  // every lane has read the old multipliers before lane 0 overwrites them
        if (l == 0) {
          for (int r = 0; r < m; ++r) {
            double yh, ez;
            row_at(r, &yh, &ez);
            a.lam[b * m + r] = yh;
          }
        }
        tsync();
        state = ST_ALM_UPDATE;
      } else /* ST_ALM_UPDATE */ {
        bool inner_conv = pstatus == MPCB_STATUS_CONVERGED;
        if (!inner_conv) ++n_ifail;
        int fin = 0;
        if (inner_conv && peps <= sp.eps_target && norm_e2 <= sp.delta_target) fin = MPCB_STATUS_CONVERGED;
        else if (alm_i + 1 == sp.alm_max_iter) fin = MPCB_STATUS_MAX_ITER;
        else if (!inner_conv && eps_tol <= sp.eps_target) fin = pstatus;
        if (fin) {
          finish_instance(fin, false);
          state = ST_IDLE;
          continue;
        }
        if (d.bounded) {
          tsync();
          if (l == 0) {
            for (int r = 0; r < m; ++r) {
              double e2 = TR(T.E2, r);
              if (norm_e2 > sp.delta_target && (alm_i == 0 || fabs(e2) > sp.theta * fabs(TR(T.E1, r))))
                TR(T.SIG, r) = fmin(sp.Sigma_max, fmax(sp.Delta * fabs(e2) / norm_e2, 1.0) * TR(T.SIG, r));
              TR(T.E1, r) = e2;
            }
          }
          tsync();
        }
        eps_tol = fmax(sp.rho * eps_tol, sp.eps_target);
        ++alm_i;
        state = ST_START_PANOC;
      }
    }
    if constexpr (ROUND) {
      // ================= round mode: publish the request, save the registers, leave ===============
      // (the evaluation runs in k_eval_requests over the compacted request lists)
      unsigned* cnt = a.counters + 4 * (rnd & 1);
      if (blockIdx.x == 0 && threadIdx.x == 0) {  // the other parity's counters: last read by the previous evaluation launch
        unsigned* nx = a.counters + 4 * ((rnd + 1) & 1);
        nx[0] = 0u; nx[1] = 0u; nx[2] = 0u; nx[3] = 0u;
      }
      if (!dead) {
        if (__ballot_sync(tmask, want != 0) & tmask) ++t_rounds;
        if (exhausted && state == ST_IDLE) dead = true;
      }
      const unsigned mg = __ballot_sync(0xffffffffu, want == 2);
      const unsigned mv = __ballot_sync(0xffffffffu, want == 1 || want == 3);
      const unsigned ml = __ballot_sync(0xffffffffu, l == 0 && !dead);  // tiles still solving (speculation policy)
      unsigned bg = 0, bv = 0;
      if (lane == 0) {
        if (mg) bg = atomicAdd(&cnt[0], (unsigned)__popc(mg));
        if (mv) bv = atomicAdd(&cnt[1], (unsigned)__popc(mv));
        if (ml) atomicAdd(&cnt[3], (unsigned)__popc(ml));
      }
      bg = __shfl_sync(0xffffffffu, bg, 0);
      bv = __shfl_sync(0xffffffffu, bv, 0);
      const unsigned below_me = (1u << lane) - 1u;
      if (want) {
        a.req_want[slot] = want; a.req_b[slot] = b; a.req_p[slot] = evp.p;
        a.req_g[slot] = want == 3 ? WL(M.ST, l).p : evg.p;
        if (want == 2) { a.list_grad[bg + __popc(mg & below_me)] = (int32_t)slot; ++n_grad; }
        else { a.list_val[bv + __popc(mv & below_me)] = (int32_t)slot; if (want == 1) ++n_val; }
      }
      // ---- bucket the tile for the next transition launch: what it does there is decided by its state, and in
      // ST_LS_STEP by whether it waits for psi(x̂+) (the round that usually ends the iteration: L-BFGS update, next
      // direction, next candidates) or for the gradients of new candidates (prox steps only)
      {
        __shared__ int s_kcnt[MPCB_NKEY], s_kbase[MPCB_NKEY];
        if (threadIdx.x < MPCB_NKEY) s_kcnt[threadIdx.x] = 0;
        __syncthreads();
        const bool rq_any = (__ballot_sync(tmask, c_rq) & tmask) != 0u;
        const int key = state == ST_LS_STEP ? (rq_any ? (int)ST_LS_STEP : (int)ST_ALM_UPDATE + 1) : state;
        int rank = -1;
        if (l == 0 && !dead) rank = atomicAdd(&s_kcnt[key], 1);
        __syncthreads();
        int32_t* kc_out = a.key_cnt + MPCB_NKEY * ((rnd + 1) & 1);
        if (threadIdx.x < MPCB_NKEY && s_kcnt[threadIdx.x] > 0)
          s_kbase[threadIdx.x] = atomicAdd(&kc_out[threadIdx.x], s_kcnt[threadIdx.x]);
        __syncthreads();
        if (rank >= 0)
          a.key_list[((int64_t)((rnd + 1) & 1) * MPCB_NKEY + key) * nt + s_kbase[key] + rank] = (int32_t)tile;
      }
      if (!was_dead) {  // tiles that were already dead keep their saved flags
        sib[0] = (int)exhausted | (int)fd_ready << 1 | (int)c_valid << 2 | (int)c_pend << 3 | (int)c_rq << 4 | (int)c_fb << 5 |
                     (int)c_acc << 6 | (int)c_started << 7 | (int)fd_done << 8 | (int)c_needg << 9 | (int)dead << 10;
        if (!dead) {
          int r = 1;
#define X(v) sib[(r++) * MPCB_LB] = v;
          MPCB_STATE_I(X)
#undef X
          sib[(r++) * MPCB_LB] = (int)specmask;
          r = 0;
#define X(v) sdb[(r++) * MPCB_LB] = v;
          MPCB_STATE_D(X)
#undef X
          sdb[(r++) * MPCB_LB] = __longlong_as_double((long long)mask0);
          sdb[(r++) * MPCB_LB] = __longlong_as_double((long long)mask1);
          sdb[(r++) * MPCB_LB] = __longlong_as_double((long long)b);
        }
      }
      return;
    } else {
    // ================= evaluation phase: the one convergent site ================================
    if (!__any_sync(0xffffffffu, want != 0)) break;
    const bool anyg = __any_sync(0xffffffffu, want == 2);
    { long long t = clock64(); t_trans += t - t_mark; t_mark = t; }
    if (__ballot_sync(0xffffffffu, want != 0) & tmask) ++t_rounds;
    if (want) {
      RoadView r = road_view(a.ra, sroad, d.S, b);
      Extras ex{a.obs ? a.obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, d.bounded ? a.lam + b * m : nullptr,
                d.bounded ? &TR(T.SIG, 0) : nullptr, 1, TB};
      double y0[NX];
#pragma unroll
      for (int i = 0; i < NX; ++i) y0[i] = a.x0[b * NX + i];
      const Vec scr = WL(M.SCR, l);
      const Vec stv = want == 3 ? WL(M.ST, l) : Vec{nullptr, 0};
      // ONE variant for every request of the persistent kernel (psi-only requests skip the stores): the bits of a
      // psi must not depend on which other requests happen to share its warp in this trip.  (Until round 2 a psi-only
      // request took the cheaper value-only variant when its whole warp was psi-only; the sin kernels of the two
      // variants differ in the last bits, so results depended on co-scheduling: tests/test_gpu_exactness.py.)
      val = eval_fm<MODEL, true>(r, ex, y0, evp, scr, evg, want == 2, stv, IVec{nullptr, 0}, !d.exact);
      if (want == 2) ++n_grad;
      else if (want == 1) ++n_val;
    }
    __syncwarp();
    { long long t = clock64(); t_eval += t - t_mark; t_mark = t; ++n_rounds; n_grounds += anyg; }
    }
  }
  if (a.dbg && lane == 0) {
    long long* o = a.dbg + 4 * (slot >> 5);
    o[0] = t_eval; o[1] = t_trans; o[2] = n_rounds; o[3] = n_grounds;
  }
}

// ---- round mode: evaluation of the published requests -------------------------------------------------
// One launch per round, after the transition launch.  Warps pull chunks of 32 requests from the two
// compacted lists (gradient requests first, then psi-only requests), so every warp runs ONE variant of
// the evaluation with all its lanes busy, whatever PANOC iteration the owning tiles are in, and the
// requests of the few instances still running at the end of a batch spread over the whole GPU.  Only the
// evaluation code is resident (the transition kernel's 400 KB never compete for the instruction cache).
struct EvalArgs {
  int64_t ns, nt, ens;
  int G;
  const double *x0, *p, *obs, *lam;
  const double* tp_base;    // plain per-tile rows [block][row][tile in block]; ALM penalties = rows sig_row ..
  int tp_rows, sig_row;     // rows per tile of that region, first penalty row
  int erows;                // rows of the evaluation scratch per lane (N * stage doubles), blocked like the workspace
  RoadArg ra;
  const int32_t* req_want;
  const int64_t* req_b;
  double* const* req_p;
  double* const* req_g;
  double* req_val;
  const int32_t *list_grad, *list_val;
  unsigned* counters;       // this round's [4]
  unsigned* ctl;            // [4] control words read by the next transition launch: [0] live tiles
  int32_t* key_cnt_zero;    // [MPCB_NKEY] bucket counts the next transition launch appends to (zeroed here)
  unsigned* round_stats;    // [2] this round's (gradient, psi-only) request counts, kept for the solve profile; or null
  // graph-driven loop (round_ptr != null): the per-round pointers above are derived in the kernel from the round index
  const unsigned* round_ptr;
  unsigned* counters_base;        // [2][4]
  int32_t* key_cnt_base;          // [2][MPCB_NKEY]
  unsigned* round_stats_base;     // [stat_rounds][2]
  unsigned long long* stamps;     // [stat_rounds][2] %globaltimer at the start of the evaluation launch / end of the round
  int stat_rounds;
  double* escr;             // stage-Jacobian scratch of the evaluating lanes [N * K, ens]
  long long* trace;         // diagnostic (MPCB_EVAL_TRACE=round): per chunk [cycles, first slot] or null
  int split_ok;             // 0: never use the split evaluation (MPCB_SPLIT=0)
  int pit_ok;               // 0: never use the parallel-in-time evaluation (MPCB_PIT=0)
  int work_blocks;          // CTAs that own evaluation scratch and take chunks (the grid may be larger: CTAs beyond only take
                            // parallel-in-time requests, which live in shared memory)
};

// Launch shape: ONE CTA of 256 lanes per SM, so a CTA's warps w and w + 4 share an SM sub-partition and
// the first chunk of every warp is assigned statically, sub-partition by sub-partition across the whole GPU
// (chunk c -> SM c % gridDim, warp (c / gridDim) % 8): up to 4 x SMs chunks run one per sub-partition, the
// next 4 x SMs double up, and only beyond that do warps pull further chunks from the work counter.  (With
// two 128-lane CTAs per SM and a purely dynamic queue ncu showed sub-partitions holding 4 busy warps next
// to idle ones: 2.2 M cycles elapsed for 0.95 M cycles of average SM activity, profiles/r01_round_mode.md.)
template <int MODEL>
__global__ void __launch_bounds__(256, 1)
k_eval_requests(KDesc d, EvalArgs e) {
  constexpr int NX = Fm<MODEL>::NX;
  extern __shared__ double smem[];
  double* sroad = smem;
  fm_load_smem();
  if (e.ra.in_smem) stage_roads(sroad, e.ra.road, e.ra.R, d.S);
  stage_problem(d, e.p);
  const int lane = threadIdx.x & 31;
  const int64_t eslot = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const Vec scr{e.escr + (eslot / MPCB_LB) * ((int64_t)e.erows * MPCB_LB) + (eslot % MPCB_LB), MPCB_LB};
  // ALM penalties of the tile that owns lane `slot`: first row and stride between rows
  auto sig_of = [&](int64_t slot) -> const double* {
    const int TBe = MPCB_LB / e.G;
    return e.tp_base + (slot / MPCB_LB) * ((int64_t)e.tp_rows * TBe) + (int64_t)e.sig_row * TBe + (slot % MPCB_LB) / e.G;
  };
  const int64_t sig_stride = MPCB_LB / e.G;
  if (e.round_ptr) {  // graph-driven loop: this round's buffers
    const unsigned r = *e.round_ptr;
    e.counters = e.counters_base + 4 * (r & 1u);
    e.key_cnt_zero = e.key_cnt_base + MPCB_NKEY * (r & 1u);
    e.round_stats = r < (unsigned)e.stat_rounds ? e.round_stats_base + 2 * r : nullptr;
    if (blockIdx.x == 0 && threadIdx.x == 0 && r < (unsigned)e.stat_rounds) {
      unsigned long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      e.stamps[2 * r] = t;
    }
  }
  const unsigned ng = e.counters[0], nv = e.counters[1];
  if (blockIdx.x == 0 && threadIdx.x == 0 && e.ctl) e.ctl[0] = e.counters[3];
  if (blockIdx.x == 0 && threadIdx.x == 0 && e.round_stats) { e.round_stats[0] = ng; e.round_stats[1] = nv; }
  if (blockIdx.x == 0 && threadIdx.x < MPCB_NKEY && e.key_cnt_zero) e.key_cnt_zero[threadIdx.x] = 0;
  const unsigned cv = (nv + 31u) >> 5;
  // Few gradient requests (each can have a CTA of its own next to the psi-only chunks): parallel-in-time evaluation,
  // one request per CTA from the END of the grid (the psi-only chunks are handed out from its start).
  // Fewer requests than CTAs: the psi-only requests are evaluated that way too, every request on an SM of its own.
  const bool pit_all = Fm<MODEL>::PIT && e.pit_ok && e.pit_ok != 2 && ng + nv <= gridDim.x;
  const bool pit = pit_all || (Fm<MODEL>::PIT && e.pit_ok && ng > 0 && ng + (cv < 8u ? cv : 8u) <= gridDim.x);
  if (pit) {
    double* pit_sm = smem + (e.ra.in_smem ? e.ra.R * road_smem_doubles(d.S) : 0);
    const unsigned ridx = gridDim.x - 1u - blockIdx.x;
    if (ridx < ng || (pit_all && ridx < ng + nv)) {
      const bool grad = ridx < ng;
      const int64_t slot = grad ? e.list_grad[ridx] : e.list_val[ridx - ng];
      const int64_t b = e.req_b[slot];
      RoadView r = road_view(e.ra, sroad, d.S, b);
      Extras ex{e.obs ? e.obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, d.bounded ? e.lam + b * d.m : nullptr,
                d.bounded ? sig_of(slot) : nullptr, 1, sig_stride};
      double v = 0.0;
      if constexpr (Fm<MODEL>::PIT) {
        if (grad) v = eval_fm_pit<MODEL, true>(r, ex, e.x0 + b * NX, Vec{e.req_p[slot], MPCB_LB}, Vec{e.req_g[slot], MPCB_LB}, pit_sm, !d.exact);
        else v = eval_fm_pit<MODEL, false>(r, ex, e.x0 + b * NX, Vec{e.req_p[slot], MPCB_LB},
                                           e.req_want[slot] == 3 ? Vec{e.req_g[slot], MPCB_LB} : Vec{nullptr, 0}, pit_sm, !d.exact);
      }
      if (threadIdx.x == 0) e.req_val[slot] = v;
      return;  // (a CTA with a parallel-in-time request takes no chunks: nb below keeps them apart)
    }
  }
  if (pit_all) return;
  // CTAs that take chunks: those with evaluation scratch, minus the ones busy with parallel-in-time requests
  unsigned nb = (unsigned)e.work_blocks < gridDim.x ? (unsigned)e.work_blocks : gridDim.x;
  if (pit && gridDim.x - ng < nb) nb = gridDim.x - ng;
  if (blockIdx.x >= nb) return;
  const unsigned n_warps = nb * (blockDim.x >> 5);
  // few gradient requests (every chunk still gets a warp slot of its own on half the sub-partitions): 8 lanes per
  // request, one tangent direction each (eval_fm_split) — the latency floor of a round drops with the
  // per-warp instruction count
  const bool split = !pit && e.split_ok && ng > 0 && ((ng + 3u) >> 2) + cv <= (n_warps >> 1);
  const unsigned cg = pit ? 0u : (split ? (ng + 3u) >> 2 : (ng + 31u) >> 5);
  bool first = true;
  for (;;) {
    unsigned c = 0;
    if (first) {
      c = (threadIdx.x >> 5) * nb + blockIdx.x;
      first = false;
    } else {
      if (lane == 0) c = n_warps + atomicAdd(&e.counters[2], 1u);
      c = __shfl_sync(0xffffffffu, c, 0);
    }
    if (c >= cg + cv) break;
    const bool grad = c < cg;
    const unsigned idx = ((grad ? c : c - cg) << 5) + (unsigned)lane;
    const long long t_chunk = e.trace ? clock64() : 0;
    bool nonfinite = false;
    if (grad && split) {
      const int J = lane & 7;
      const unsigned ridx = (c << 2) + (unsigned)(lane >> 3);
      if (ridx < ng) {
        const int64_t slot = e.list_grad[ridx];
        const int64_t b = e.req_b[slot];
        RoadView r = road_view(e.ra, sroad, d.S, b);
        Extras ex{e.obs ? e.obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, d.bounded ? e.lam + b * d.m : nullptr,
                  d.bounded ? sig_of(slot) : nullptr, 1, sig_stride};
        double y0[NX];
#pragma unroll
        for (int i = 0; i < NX; ++i) y0[i] = e.x0[b * NX + i];
        const Vec scr_lead{scr.p - J, MPCB_LB};
        const double v = eval_fm_split<MODEL>(r, ex, y0, Vec{e.req_p[slot], MPCB_LB}, scr_lead, Vec{e.req_g[slot], MPCB_LB}, J,
                                              0xffu << (lane & 24));
        if (J == 0) e.req_val[slot] = v;
        nonfinite = J == 0 && !(fabs(v) < 1e300);
      }
    } else if (idx < (grad ? ng : nv)) {
      const int64_t slot = grad ? e.list_grad[idx] : e.list_val[idx];
      const int want = e.req_want[slot];
      const int64_t b = e.req_b[slot];
      RoadView r = road_view(e.ra, sroad, d.S, b);
      Extras ex{e.obs ? e.obs + b * d.n_obs * MPCB_OBS_DIM : nullptr, d.bounded ? e.lam + b * d.m : nullptr,
                d.bounded ? sig_of(slot) : nullptr, 1, sig_stride};
      double y0[NX];
#pragma unroll
      for (int i = 0; i < NX; ++i) y0[i] = e.x0[b * NX + i];
      const Vec pt{e.req_p[slot], MPCB_LB};
      const Vec out{e.req_g[slot], MPCB_LB};
      double v;
      if (grad) v = eval_fm<MODEL, true>(r, ex, y0, pt, scr, out, true, Vec{nullptr, 0}, IVec{nullptr, 0}, !d.exact);
      else v = eval_fm<MODEL, false>(r, ex, y0, pt, scr, Vec{nullptr, 0}, false, want == 3 ? out : Vec{nullptr, 0},
                                     IVec{nullptr, 0}, !d.exact);
      e.req_val[slot] = v;
      nonfinite = !(fabs(v) < 1e300);
    }
    __syncwarp();
    if (e.trace) {
      const unsigned nf = __ballot_sync(0xffffffffu, nonfinite);
      if (lane == 0) {
        e.trace[2 * c] = clock64() - t_chunk;
        e.trace[2 * c + 1] = (grad ? 1000 : 0) + __popc(nf);  // lanes whose psi came out NaN / Inf / > 1e300
      }
    }
  }
}

// ---- graph-driven loop: the node that closes a round ----------------------------------------------------
// Runs after the evaluation launch of round r inside the WHILE body of the solve's CUDA graph: decides whether another
// round is needed (any request published), advances the round index the kernels read, picks the transition launch's
// grid for the next round (the IF node: a small grid once few tiles are alive) and stamps the time.
__global__ void k_round_advance(unsigned* ctl, const unsigned* counters_base, cudaGraphConditionalHandle more_h,
                                cudaGraphConditionalHandle small_h, unsigned small_tiles, unsigned round_cap,
                                unsigned long long* stamps, int stat_rounds) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const unsigned r = ctl[1];
  const unsigned* c = counters_base + 4 * (r & 1u);
  const bool more = (c[0] + c[1]) > 0u && r + 1u < round_cap;
  if (r < (unsigned)stat_rounds) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    stamps[2 * r + 1] = t;
  }
  ctl[1] = r + 1u;
  if (!more && (c[0] + c[1]) > 0u) ctl[2] = 1u;  // stopped by the round cap
  cudaGraphSetConditional(more_h, more ? 1u : 0u);
  cudaGraphSetConditional(small_h, c[3] <= small_tiles ? 1u : 0u);  // c[3]: tiles still solving after this round
}

}  // namespace mpcb
