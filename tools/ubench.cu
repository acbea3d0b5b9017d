// ubench.cu — micro-benchmarks that size the solver design (run on the B200 box).
//   1. FP64 issue cost of a partially populated warp: DFMA chains (ILP 8) with 32/16/8/4 active lanes.
//   2. Accuracy of rcp.approx.ftz.f64 and of its Newton refinements.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/ubench tools/ubench.cu
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>

__global__ void k_dfma(double* out, int iters, int active, long long* cyc) {
  int lane = threadIdx.x & 31;
  double a[8];
  for (int i = 0; i < 8; ++i) a[i] = 1.0 + 1e-9 * (threadIdx.x + i);
  double b = 1.0000001, c = 1e-9;
  long long t0 = 0, t1 = 0;
  if (lane < active) {
    t0 = clock64();
#pragma unroll 1
    for (int it = 0; it < iters; ++it) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = fma(a[i], b, c);
    }
    t1 = clock64();
    double s = 0;
    for (int i = 0; i < 8; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (lane == 0 && blockIdx.x == 0 && threadIdx.x == 0) *cyc = t1 - t0;
  }
}

__device__ __forceinline__ double rcp_approx(double x) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(x));
  return r;
}

__global__ void k_rcp(double* err4, int n) {
  double e0 = 0, e1 = 0, e2 = 0, e3 = 0;
  for (int i = threadIdx.x + blockIdx.x * blockDim.x; i < n; i += blockDim.x * gridDim.x) {
    double x = 0.3 + 3.7 * (double)i / n;
    if (i & 1) x = -x * 1e3;
    double ex = 1.0 / x;
    double r = rcp_approx(x);
    e0 = fmax(e0, fabs(r - ex) / fabs(ex));
    double e = fma(-x, r, 1.0);
    double r1 = fma(r, e, r);
    e1 = fmax(e1, fabs(r1 - ex) / fabs(ex));
    double r2 = fma(r, fma(e, e, e), r);  // third order
    e2 = fmax(e2, fabs(r2 - ex) / fabs(ex));
    double ee = fma(-x, r1, 1.0);
    double r3 = fma(r1, ee, r1);          // two Newton steps
    e3 = fmax(e3, fabs(r3 - ex) / fabs(ex));
  }
  // crude max-reduce
  for (int o = 16; o; o >>= 1) {
    e0 = fmax(e0, __shfl_xor_sync(~0u, e0, o)); e1 = fmax(e1, __shfl_xor_sync(~0u, e1, o));
    e2 = fmax(e2, __shfl_xor_sync(~0u, e2, o)); e3 = fmax(e3, __shfl_xor_sync(~0u, e3, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax((unsigned long long*)&err4[0], __double_as_longlong(e0));
    atomicMax((unsigned long long*)&err4[1], __double_as_longlong(e1));
    atomicMax((unsigned long long*)&err4[2], __double_as_longlong(e2));
    atomicMax((unsigned long long*)&err4[3], __double_as_longlong(e3));
  }
}

int main() {
  double* out; long long* cyc; double* err;
  cudaMalloc(&out, sizeof(double) * 1 << 20);
  cudaMalloc(&cyc, 8);
  cudaMalloc(&err, 32);
  const int iters = 20000;
  printf("{\"dfma_partial_warp\": [");
  bool first = true;
  for (int wps : {1, 2, 4}) {          // warps per SMSP (block = 4*wps warps on one SM)
    for (int active : {32, 16, 8, 4}) {
      k_dfma<<<1, 128 * wps>>>(out, iters, active, cyc);
      k_dfma<<<1, 128 * wps>>>(out, iters, active, cyc);
      long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
      printf("%s{\"warps_per_smsp\": %d, \"active_lanes\": %d, \"cycles_per_warp_dfma\": %.3f}", first ? "" : ", ", wps,
             active, (double)c / (iters * 8.0));
      first = false;
    }
  }
  printf("],\n\"dfma_warps_per_sm\": [");
  first = true;
  for (int w : {1, 2, 3, 4, 6, 8}) {   // warps on ONE SM (round-robin over its 4 sub-partitions), all lanes active
    k_dfma<<<1, 32 * w>>>(out, iters, 32, cyc);
    k_dfma<<<1, 32 * w>>>(out, iters, 32, cyc);
    long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost);
    printf("%s{\"warps_per_sm\": %d, \"cycles_per_warp_dfma\": %.3f}", first ? "" : ", ", w, (double)c / (iters * 8.0));
    first = false;
  }
  printf("],\n");
  cudaMemset(err, 0, 32);
  k_rcp<<<148, 256>>>(err, 1 << 24);
  double e[4]; cudaMemcpy(e, err, 32, cudaMemcpyDeviceToHost);
  printf("\"rcp_approx_relerr\": {\"raw\": %.3e, \"newton1\": %.3e, \"third_order\": %.3e, \"newton2\": %.3e}}\n", e[0], e[1], e[2], e[3]);
  cudaError_t ce = cudaDeviceSynchronize();
  if (ce != cudaSuccess) { fprintf(stderr, "cuda error %s\n", cudaGetErrorString(ce)); return 1; }
  return 0;
}
