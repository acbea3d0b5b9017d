/* mpcb200.h — C ABI of libmpcb200.so: batched nonlinear MPC (PANOC + L-BFGS, ALM) on B200.
 *
 * This is the drop-in boundary for the one hot path of panagiotou23/model-predictive-control:
 * the NLP solve inside MPCController.__call__ (reference controller.py:51-69), which the
 * reference hands to alpaqa's ALMSolver(StructuredPANOCLBFGSSolver) over CasADi-generated
 * C functions f/grad_f/g/grad_g (reference main.py:25-59).  The reference's real FFI for
 * this path is the CasADi generated-C ABI `int f(const double** arg, double** res,
 * long long* iw, double* w, int mem)` dlopen'ed by alpaqa (main.py:54); it is per-instance
 * and CPU-only, so it is replaced — not mirrored — by the batched entry points below.
 *
 * Conventions
 *   - every data pointer is a DEVICE pointer owned by the caller (torch tensors), fp64,
 *     row-major [B, ...]; the *_host entry points take HOST pointers and do the copies.
 *   - every call is asynchronous on the cudaStream_t passed as `void* stream` (0 = legacy
 *     default stream); *_host entry points synchronise before returning.
 *   - return value: 0 = ok, negative = MPCB_E_* ; never throws.
 *   - U is the flat decision vector [d0, δ0, d1, δ1, ...] (column-major 2×N, reference
 *     car_dynamics.py:149-156); the road is the flat column-major centerline
 *     [x0..xS-1, y0..yS-1] (reference main.py:113, car_dynamics.py:204).
 */
#ifndef MPCB200_H
#define MPCB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MPCB_VERSION 100 /* 0.1.0 */

/* model_id */
#define MPCB_MODEL_PACEJKA6 0   /* reference car_dynamics.py:93-132 / dynamics.py:67-119 */
#define MPCB_MODEL_KINEMATIC4 1 /* reference dynamics.py:144-173 */

/* road_mode: how the `road` argument is laid out */
#define MPCB_ROAD_SHARED 0       /* road[2S]: one polyline for every instance (main.py:112-113) */
#define MPCB_ROAD_TABLE 1        /* road[R,2S] + road_idx[B] (int32): a few road shapes shared by the batch */
#define MPCB_ROAD_PER_INSTANCE 2 /* road[B,2S] */
#define MPCB_ROAD_BEZIER 3       /* road[B,12]: quintic control points Px0..5,Py0..5 (bezier_curves.py:19-48),
                                    sampled on device to S vertices at j=i/(S-1) */

/* flags */
#define MPCB_FLAG_LANE 1      /* add u_trgl + u_rbd (dpf_test.py:45-54) to every stage */
#define MPCB_FLAG_OBST 2      /* add dnf (dpf_test.py:81-102) for n_obs obstacles to every stage */
#define MPCB_FLAG_BOUNDED_D 4 /* state constraints g (main.py:43-52) are bounded by D_lower/D_upper */
#define MPCB_FLAG_SAFE_DIST 16 /* n_obs more rows of g per stage: the safe-distance ellipse around every obstacle,
                                 g = 1 - (dx / x_s)^2 - (dy / y_s)^2 <= 0 with (x_s, y_s) = get_safe_distances
                                 (dpf_test.py:57-64), D = (-inf, 0]; handled by the ALM like the state rows */
#define MPCB_FLAG_EXACT 8     /* diagnostic: solver shortcuts off (early line-search rejection, descent-lemma steps
                                 decided without psi(x̂+), evaluations cut at a blown-up state); same iterates, slower;
                                 MPCB_EXACT=1 in the environment sets it for every handle */

/* status: alpaqa 0.0.1 SolverStatus order (reference controller.py:64 compares against Converged) */
#define MPCB_STATUS_UNKNOWN 0
#define MPCB_STATUS_CONVERGED 1
#define MPCB_STATUS_MAX_TIME 2
#define MPCB_STATUS_MAX_ITER 3
#define MPCB_STATUS_NOT_FINITE 4
#define MPCB_STATUS_NO_PROGRESS 5
#define MPCB_STATUS_INTERRUPTED 6

/* error codes */
#define MPCB_OK 0
#define MPCB_E_ARG (-1)      /* bad argument (null pointer, size out of range) */
#define MPCB_E_CUDA (-2)     /* CUDA runtime error; see mpcb_last_error */
#define MPCB_E_UNSUPPORTED (-3)
#define MPCB_E_NODEVICE (-4) /* no CUDA device: there is no CPU fallback */

#define MPCB_MAX_N 64   /* horizon cap */
#define MPCB_MAX_OBS 8  /* obstacles per instance cap */
#define MPCB_NP 22      /* vehicle parameter vector length (car_dynamics.py:65-88) */
#define MPCB_OBS_DIM 5  /* per obstacle: x, y, phi, v, a */

/* Problem description: everything that CasADi would bake into the generated C
 * (reference main.py:25-59, car_dynamics.py:93-147,230-258). */
typedef struct MpcProblemDesc {
  int32_t model_id; /* MPCB_MODEL_* */
  int32_t N;        /* horizon length N_horiz (main.py:68; 12 in the demo) */
  int32_t S;        /* centerline vertices (main.py:70; 100) */
  int32_t n_sub;    /* RK4 finite elements per stage (car_dynamics.py:136; 4) */
  int32_t road_mode;
  int32_t n_roads;  /* R for MPCB_ROAD_TABLE */
  int32_t flags;
  int32_t n_obs;
  double Ts;        /* stage length (car_dynamics.py:93; 0.05) */
  double v_ref;     /* target speed (main.py:65; 1.0) */
  double c[6];      /* stage-cost weights (car_dynamics.py:230): speed, cte, pos, heading, steer, drive */
  double kin_friction; /* dynamics.py:34 (1)  — Kinematic-4 only */
  double kin_accel;    /* dynamics.py:35 (2)  — Kinematic-4 only */
  /* lane potentials, dpf_test.py:5,45-54 */
  double lane_a, lane_y_t;           /* u_trgl: a (0.5), y_trgl (1.75) */
  double lane_b, lane_y_br, lane_y_bl; /* u_rbd: b (100), y_br (1), y_bl (6) */
  /* obstacle potential, dpf_test.py:81 */
  double obs_Af, obs_b, obs_sigma_x, obs_sigma_y, obs_alpha_div; /* 1000, 1, 2, 0.5, 5 */
  /* state constraints g_i = y_i^2 - g_offset[i] (main.py:46-51) and their box D */
  double g_offset[6];
  double D_lower[6], D_upper[6]; /* +-inf = unbounded (the reference leaves D at +-inf, main.py:57) */
  /* safe distances, dpf_test.py:6,57-64: x_s = x_0/2 + (v_x - v_x,obs)^2 / (2 a_x_max), same in y */
  double safe_x0, safe_y0, safe_ax_max, safe_ay_max; /* 5, 3, 3, 1 */
} MpcProblemDesc;

/* Solver parameters: alpaqa 0.0.1 defaults with the overrides of reference controller.py:27-48.
 * Wall-clock caps (controller.py:30,44) are deliberately absent: they make results time-dependent. */
typedef struct MpcSolverParams {
  /* inner: StructuredPANOCLBFGS */
  int32_t max_iter;        /* 1000 (controller.py:31) */
  int32_t lbfgs_mem;       /* N_horiz (controller.py:36) */
  int32_t max_no_progress; /* 10 */
  int32_t hess_vec_fd;     /* 1: finite-difference Hessian-vector correction on the active set */
  double Lgamma_factor;    /* 0.95 */
  double L_min, L_max;     /* 1e-5, 1e20 */
  double lip_eps, lip_delta; /* 1e-6, 1e-12 */
  double tau_min;          /* 1/256 */
  double quad_tol;         /* 10*DBL_EPSILON: descent-lemma rounding tolerance factor */
  double ls_tol;           /* 10*DBL_EPSILON: line-search rounding tolerance factor */
  /* outer: ALM (controller.py:39-48) */
  int32_t alm_max_iter;    /* 1000 */
  int32_t reserved0;
  double eps_target;       /* ε  = 1e-6 */
  double delta_target;     /* δ  = 1e-4 */
  double Sigma0;           /* Σ_0 = 1e5 */
  double eps0;             /* ε_0 = 1 */
  double rho;              /* ρ = 0.1 */
  double theta;            /* θ = 0.1 */
  double Delta;            /* Δ = 10 */
  double M;                /* 1e9 */
  double Sigma_max;        /* 1e9 */
} MpcSolverParams;

typedef struct MpcHandle MpcHandle;

int mpcb_version(void);
/* Fill with the reference's constants (main.py:65-86, car_dynamics.py:230, controller.py:27-48). */
void mpcb_default_desc(MpcProblemDesc* desc, int32_t model_id, int32_t N, int32_t S);
void mpcb_default_solver_params(MpcSolverParams* sp, int32_t N);
/* The 22 vehicle constants of main.py:82-111 in car_dynamics.py:65-88 order. */
void mpcb_default_vehicle_params(double* p22);

int mpcb_create(const MpcProblemDesc* desc, MpcHandle** out);
void mpcb_destroy(MpcHandle* h);
const char* mpcb_last_error(MpcHandle* h);

/* Continuous-time model ẏ = f(y,u;p): replaces dynamics.py:67 / :144 __call__ (clip=1 mirrors
 * clip_inputs, dynamics.py:99) and the CasADi Function `f` (car_dynamics.py:132; clip=0). */
int mpcb_dynamics(MpcHandle* h, int64_t B, const double* x /*[B,nx]*/, const double* u /*[B,2]*/,
                  const double* p /*[22]*/, int32_t clip, double* xdot_out /*[B,nx]*/, void* stream);

/* Nearest vertex + tracking errors: replaces car_dynamics.py:174-228 (normalised=0: raw cross
 * products, scan i=1..S-2) and road.py:36-81 (normalised=1: divided by segment length, plain argmin). */
int mpcb_road_errors(MpcHandle* h, int64_t B, const double* pos /*[B,2]*/, const double* heading /*[B]*/,
                     const double* road, const int32_t* road_idx, int32_t normalised,
                     int32_t* idx_out /*[B]*/, double* err_out /*[B,3] cte, heading, pos*/, void* stream);

/* Sample the Bezier road of every instance to S vertices: replaces bezier_curves.py:19-25. */
int mpcb_bezier_sample(MpcHandle* h, int64_t B, const double* ctrl /*[B,12]*/, double* road_out /*[B,2S]*/,
                       void* stream);

/* Horizon rollout + cost: replaces model.simulate (car_dynamics.py:158-166) and the cost sum
 * (main.py:34-40).  states_out[B,N,nx] = state AFTER each input; idx_out[B,N] nearest vertex. */
int mpcb_rollout(MpcHandle* h, int64_t B, const double* x0, const double* U, const double* p,
                 const double* road, const int32_t* road_idx, const double* obs /*[B,n_obs,5] or NULL*/,
                 double* states_out, int32_t* idx_out, double* cost_out, void* stream);

/* ψ and ∇_U ψ (forward rollout + reverse sweep): replaces the generated C `f` + `grad_f`
 * (and g/grad_g through lambda/Sigma when MPCB_FLAG_BOUNDED_D; pass NULL for plain f). */
int mpcb_value_grad(MpcHandle* h, int64_t B, const double* x0, const double* U, const double* p,
                    const double* road, const int32_t* road_idx, const double* obs,
                    const double* lambda /*[B,nx*N] or NULL*/, const double* Sigma /*[B,nx*N] or NULL*/,
                    double* cost_out /*[B]*/, double* grad_out /*[B,2N]*/, void* stream);

/* Allocation-free ψ and ∇_U ψ.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same arguments, kernel and results as mpcb_value_grad; the stage-Jacobian scratch is a file-scope static
 * device array of MPCB_STATIC_SCRATCH_BYTES (held by every GPU that loads the library until exit) instead of the
 * handle's grow-only workspace, so the call makes no cudaMalloc / cudaFree and launches one kernel.
 * Limits (rejected with MPCB_E_UNSUPPORTED and a message in mpcb_last_error): B <= MPCB_STATIC_MAX_B; road_mode
 * MPCB_ROAD_SHARED or MPCB_ROAD_TABLE only (the per-instance and Bezier layouts need allocated scratch).  The
 * handle's device must be the current device.  The array is shared by every call on the device and keeps the last
 * call's contents (nothing in it needs to start at zero): calls that overlap on the GPU — two streams, two
 * handles — race; serialise them. */
#define MPCB_STATIC_SCRATCH_BYTES (32u << 20)
#define MPCB_STATIC_MAX_B (1 << 20)
int mpcb_value_grad_static(MpcHandle* h, int64_t B, const double* x0, const double* U, const double* p,
                           const double* road, const int32_t* road_idx, const double* obs,
                           const double* lambda /*[B,nx*N] or NULL*/, const double* Sigma /*[B,nx*N] or NULL*/,
                           double* cost_out /*[B]*/, double* grad_out /*[B,2N]*/, void* stream);

/* The MPC step's NLP solve: replaces `self.solver(problem, U, λ)` (controller.py:57).
 * U_inout / lambda_inout carry the warm start in and the solution out (controller.py:20-21,57). */
int mpcb_solve(MpcHandle* h, const MpcSolverParams* sp, int64_t B, const double* x0, const double* p,
               const double* road, const int32_t* road_idx, const double* obs,
               const double* lb /*[2N]*/, const double* ub /*[2N]*/,
               double* U_inout /*[B,2N]*/, double* lambda_inout /*[B,nx*N]*/,
               int32_t* status_out /*[B]*/, int32_t* iters_out /*[B,2] outer, inner*/,
               int32_t* evals_out /*[B,2] psi-only, psi+grad*/, double* psi_out /*[B]*/,
               double* res_out /*[B] final ||p||_2*/, void* stream);

/* Same solve with HOST buffers: pinned staging, H2D, solve, D2H, synchronise.  This is the call the
 * reference-facing controller makes for numpy inputs; road_idx/obs/lambda may be NULL. */
int mpcb_solve_host(MpcHandle* h, const MpcSolverParams* sp, int64_t B, const double* x0, const double* p,
                    const double* road, const int32_t* road_idx, const double* obs,
                    const double* lb, const double* ub, double* U_inout, double* lambda_inout,
                    int32_t* status_out, int32_t* iters_out, int32_t* evals_out, double* psi_out,
                    double* res_out);

/* Final step size gamma of every instance's last PANOC run in the last mpcb_solve / mpcb_solve_host on this handle
 * (the `final γ` of alpaqa's inner-solver statistics, which controller.py:59-61 prints a subset of): the scale of the
 * stopping test ||x̂ - x||_2 <= eps, needed to re-check a returned U* as a fixed point of the forward-backward map.
 * gamma_out is a DEVICE pointer [B]; B at most the last solve's batch. */
int mpcb_last_final_gamma(MpcHandle* h, int64_t B, double* gamma_out, void* stream);

/* Game-theoretic lane-change decision for a batch of scenarios: replaces Car.get_total_payoff for target lanes 1 and 2
 * (game_theory.py:205-244, with get_safety_payoff :155-176, get_safety_distance :115-153, get_velocity_payoff :178-189
 * and the car-behind term that reads the module-global ego, :222-233), the decision of the reference's sweep
 * (:377-379: change when the lane-2 payoff is larger) and get_bezier_control_points towards the car in front in ego's
 * lane (:92-113, i = 5).  ego [B,3] and cars [B,M,3] are DEVICE arrays of (x, v, lane) records; lane <= 0 marks an empty
 * slot.  car_params: HOST pointer to the 11 Car constants L, W, l, theta_max, tlc, td, ti, tau, a_max, h, Lf
 * (game_theory.py:23-56) or NULL for the defaults.  Outputs (device): payoff_out [B,2]; decision_out [B] (1 = change to
 * lane 2; nullable); ctrl_out [B,12] = Px0..5, Py0..5 relative to ego (NaN when nobody is in front or the speeds are
 * equal; nullable) — the layout MPCB_ROAD_BEZIER takes; tca_out [B] (nullable). */
void mpcb_default_car_params(double* c11);
int mpcb_game_payoff(MpcHandle* h, int64_t B, int32_t M, const double* ego, const double* cars, const double* car_params,
                     double* payoff_out, int32_t* decision_out, double* ctrl_out, double* tca_out, void* stream);

/* Bounds evidence without compute-sanitizer (closed on the B200 pool): a handle created with MPCB_GUARDS=1 in the
 * environment puts every scratch buffer it allocates between two 4 KiB canary regions; this synchronises the device
 * and returns the number of canary bytes that no longer hold their pattern (0 = no kernel wrote outside its buffers),
 * -2 if the handle was created without guards, -3 on a CUDA error. */
int64_t mpcb_check_guards(MpcHandle* h);

/* Per-solve latency of the last solve on this handle (the metric's "p50 solve latency"; alpaqa reports elapsed_time per
 * solve, controller.py:60): out is a DEVICE buffer of B + 1 uint64 — the GPU's %globaltimer (ns) at which instance b
 * returned its answer, and in out[B] the time the solve started on the GPU.  B must equal the last solve's batch. */
int mpcb_last_finish_ns(MpcHandle* h, int64_t B, uint64_t* out, void* stream);

/* Per-iteration trace in place of the reference's host progress callback (controller.py:23-24,38 prints info.ψ every
 * PANOC iteration): mpcb_set_trace(h, cap) makes every later solve on the handle record (psi, phi_gamma, gamma, c) — cost,
 * forward-backward envelope, step size (fields of alpaqa's progress info) and the line-search candidate that produced the
 * iterate (c = 0..7: tau = 2^-c, 8: the forced prox step, -1: start of a PANOC run) — at the top of each inner iteration
 * (all PANOC runs of the ALM loop, in order) into a handle-owned buffer [B, cap, 4]; cap = 0 turns it off.
 * mpcb_last_trace copies it to a DEVICE buffer [B, cap, 4]; iterations that were not reached hold NaN. */
int mpcb_set_trace(MpcHandle* h, int32_t cap);
int mpcb_last_trace(MpcHandle* h, int64_t B, double* trace_out, void* stream);

/* alpaqa's stats['inner_convergence_failures'] (printed by controller.py:61) of the last solve on this handle: PANOC
 * runs of the ALM loop that ended without Converged, per instance.  out is a DEVICE pointer [B] (int32). */
int mpcb_last_inner_failures(MpcHandle* h, int64_t B, int32_t* out, void* stream);

/* Bytes moved by the last mpcb_solve_host call (bench.py's e2e.h2d/d2h_bytes_per_step). */
void mpcb_last_transfer_bytes(MpcHandle* h, int64_t* h2d, int64_t* d2h);

/* Plant step of the closed loop (main.py:145: model.simulate(1, y_n, u_n, param)). */
int mpcb_plant_step(MpcHandle* h, int64_t B, const double* x /*[B,nx]*/, const double* u /*[B,2]*/,
                    const double* p, double* x_out /*[B,nx]*/, void* stream);

/* Measured FP64 FMA throughput of this device (dependent DFMA chains, all SMs), TFLOP/s:
 * the roofline denominator bench.py reports against. */
int mpcb_fp64_peak(double* tflops_out, double* ms_out);

/* Dependent-chain latencies on this device, SM cycles per op measured with one warp:
 * [0] DFMA [1] div [2] sqrt [3] sincos [4] atan [5] atan2 [6] sin [7] one Pacejka model call. */
int mpcb_latency_probe(double* cycles8);

/* Diagnostic, enabled by MPCB_DEBUG_TIMERS=1 in the environment at mpcb_create: per-warp
 * [evaluation cycles, transition cycles, rounds, gradient rounds] of the last mpcb_solve.  Returns the number of
 * warps written (<= max_warps). */
int64_t mpcb_debug_timers(MpcHandle* h, long long* out4, int64_t max_warps);

/* Execution profile of the last mpcb_solve on this handle: out6 = [mode (0 persistent kernel, 1 rounds), tile width G,
 * rounds, evaluation-kernel ms (sum over rounds), transition-kernel ms (sum over rounds), evaluation launches].
 * Round mode (the default; MPCB_MODE=persistent|rounds overrides) runs a transition launch and an evaluation launch per
 * round.  The loop is driven ON THE DEVICE by a CUDA graph — WHILE(any request published) { transition; evaluation;
 * advance } — so mpcb_solve is one asynchronous graph launch on the caller's stream, like every other call; the per-kernel
 * times are then differences of %globaltimer stamps the kernels write (CUDA events cannot bracket the nodes of a WHILE
 * body), and this function synchronises that stream to read them.  MPCB_LOOP=host (or a per-round diagnostic:
 * MPCB_ROUND_LOG, MPCB_EVAL_TRACE) drives the same launches from the host instead — that call blocks the host until the
 * batch is done and times every launch with CUDA events.  Same kernels, same bits either way. */
void mpcb_last_solve_profile(MpcHandle* h, double* out6);

/* Bulk phase of the last round-mode solve — the rounds that carried more evaluation requests than the evaluation launch
 * has lanes (throughput-bound; the remaining rounds serve a few slow instances at the latency floor of one evaluation):
 * out5 = [rounds, evaluation-kernel ms, transition-kernel ms, gradient requests, psi-only requests] summed over them. */
void mpcb_last_solve_bulk_profile(MpcHandle* h, double* out5);

/* Number of kernels launched through this handle since creation (bench.py's gpu_launches). */
int64_t mpcb_launch_count(MpcHandle* h);

#ifdef __cplusplus
}
#endif
#endif /* MPCB200_H */
