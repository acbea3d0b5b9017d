// ubench_model.cu — cycles per Pacejka model call of the solver's RK4 element (mpcb_fwd.cuh), for one
// lone warp (latency regime) and for a full GPU (throughput regime), value-only and with tangents.
// Built once per MPCB_FM_MODE (0 literals, 1 __constant__, 2 shared) by tools/build_ubench.sh.
#include <cstdio>
#include <cuda_runtime.h>
#include "../model-predictive-control_b200/csrc/mpcb_fwd.cuh"

using namespace mpcb;

template <bool GRAD>
__global__ void __launch_bounds__(128) k_chain(double* out, int iters, long long* cyc) {
  __shared__ VehP v;
  fm_load_smem();
  if (threadIdx.x == 0) {
    v.lf = 4.7e-2; v.lr = 5e-2; v.m = 0.1735; v.iz = 18.3e-5; v.inv_m = 1 / 0.1735; v.inv_iz = 1 / 18.3e-5;
    v.cm1 = 0.266; v.cm2 = 0.1; v.cr0 = 0.1025; v.cr2 = 0.0011; v.bf = 0.268; v.cf = 2.165; v.df = 3.47;
    v.br = 0.242; v.cr = 2.38; v.dr = 2.84;
  }
  __syncthreads();
  int t = blockIdx.x * blockDim.x + threadIdx.x;
  StageU su;
  v.trig_ok = 1;
  make_stage_u<MPCB_MODEL_PACEJKA6>(su, 0.6, 0.05 + 1e-4 * (t & 31), v, GRAD);
  double p[2] = {0, 0}, q[4] = {0.1 + 1e-3 * (t & 7), 1.0, 0.01, 0.1};
  double tp[5][2], tq[5][4];
  for (int j = 0; j < 5; ++j) { tp[j][0] = tp[j][1] = 0; for (int i = 0; i < 4; ++i) tq[j][i] = (i == j + 1) ? 1.0 : 0.0; }
  long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < iters; ++it) rk4_fm<MPCB_MODEL_PACEJKA6, GRAD>(p, q, tp, tq, su, v, 0.0125);
  long long t1 = clock64();
  double s = p[0] + p[1] + q[0] + q[1] + q[2] + q[3];
  if (GRAD) for (int j = 0; j < 5; ++j) s += tp[j][0] + tp[j][1] + tq[j][0] + tq[j][1] + tq[j][2] + tq[j][3];
  out[t] = s;
  if (t == 0) *cyc = t1 - t0;
}

template <bool GRAD>
void run(const char* name, int blocks, int threads, int iters, double* out, long long* cyc, bool last) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  k_chain<GRAD><<<blocks, threads>>>(out, iters, cyc);
  cudaEventRecord(e0);
  k_chain<GRAD><<<blocks, threads>>>(out, iters, cyc);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
  double calls = 4.0 * iters;
  printf("  {\"case\": \"%s\", \"grad\": %d, \"blocks\": %d, \"threads\": %d, \"cycles_per_call_warp0\": %.1f, "
         "\"ms\": %.4f, \"lane_calls_per_s\": %.4g}%s\n", name, (int)GRAD, blocks, threads, (double)c / calls, ms,
         calls * blocks * threads / (ms * 1e-3), last ? "" : ",");
}

int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, sizeof(double) * 148 * 16 * 128);
  cudaMalloc(&cyc, 8);
  printf("{\"fm_mode\": %d, \"results\": [\n", MPCB_FM_MODE);
  run<false>("lone_warp", 1, 32, 2000, out, cyc, false);
  run<true>("lone_warp", 1, 32, 2000, out, cyc, false);
  run<false>("1_warp_per_smsp", 148, 128, 2000, out, cyc, false);
  run<true>("1_warp_per_smsp", 148, 128, 2000, out, cyc, false);
  run<false>("2_warps_per_smsp", 148 * 2, 128, 2000, out, cyc, false);
  run<true>("2_warps_per_smsp", 148 * 2, 128, 2000, out, cyc, false);
  run<false>("4_warps_per_smsp", 148 * 4, 128, 2000, out, cyc, false);
  run<true>("2_warps_per_smsp_x", 148 * 2, 128, 2000, out, cyc, true);
  printf("]}\n");
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { fprintf(stderr, "cuda: %s\n", cudaGetErrorString(e)); return 1; }
  return 0;
}
