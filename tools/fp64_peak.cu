// FP64 pipe microbenchmark for B200 (sm_100a): DMMA.8x8x4 and DFMA issue rates.
// Used once per pod to establish the FP64 roofline denominator;
// MEASURED_PEAKS.json holds no FP64 figure.
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

template <int NACC>
__global__ void __launch_bounds__(1024) dmma_kernel(double* out, int iters, long long* cyc) {
  double c[NACC][2];
#pragma unroll
  for (int i = 0; i < NACC; ++i) { c[i][0] = 0.0; c[i][1] = 0.0; }
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

template <int NACC>
__global__ void __launch_bounds__(1024) dfma_kernel(double* out, int iters, long long* cyc) {
  double c[NACC];
#pragma unroll
  for (int i = 0; i < NACC; ++i) c[i] = i;
  double a = 1.0 + threadIdx.x * 1e-12, b = 1e-9 * threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < NACC; ++i) c[i] = fma(c[i], a, b);
  }
  long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < NACC; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0 && blockIdx.x == 0) *cyc = t1 - t0;
}

int main() {
  int dev = 0; cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, dev));
  int nsm = p.multiProcessorCount;
  printf("device %s, %d SMs, clock %d kHz\n", p.name, nsm, p.clockRate);
  double* out; long long* cyc; CK(cudaMalloc(&out, sizeof(double) * nsm * 4 * 1024)); CK(cudaMalloc(&cyc, 8));
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  const int iters = 20000;
  int warps_list[] = {4, 8, 16, 32};
  for (int wi = 0; wi < 4; ++wi) {
    int warps = warps_list[wi];
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaEventRecord(e0));
      dmma_kernel<16><<<nsm, warps * 32>>>(out, iters, cyc);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost));
      double nmma = (double)iters * 16 * warps * nsm;
      double flops = nmma * 512.0;
      if (rep == 2)
        printf("DMMA.8x8x4 warps/SM=%2d: %.3f ms, %.2f TFLOP/s, %.2f cyc per DMMA per SMSP, eff clock %.0f MHz\n",
               warps, ms, flops / ms * 1e-9, (double)c / ((double)iters * 16 * warps / 4.0), c / (ms * 1e3));
    }
  }
  for (int wi = 0; wi < 4; ++wi) {
    int warps = warps_list[wi];
    for (int rep = 0; rep < 3; ++rep) {
      CK(cudaEventRecord(e0));
      dfma_kernel<16><<<nsm, warps * 32>>>(out, iters, cyc);
      CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
      float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
      long long c; CK(cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost));
      double nf = (double)iters * 16 * warps * 32 * nsm;
      if (rep == 2)
        printf("DFMA       warps/SM=%2d: %.3f ms, %.2f TFLOP/s, %.2f cyc per warp-DFMA per SMSP, eff clock %.0f MHz\n",
               warps, ms, nf * 2 / ms * 1e-9, (double)c / ((double)iters * 16 * warps / 4.0), c / (ms * 1e3));
    }
  }
  // sustained: ~2 s of DMMA back to back
  {
    CK(cudaEventRecord(e0));
    int n = 0;
    for (; n < 40; ++n) dmma_kernel<16><<<nsm, 16 * 32>>>(out, iters * 4, cyc);
    CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
    double flops = (double)n * iters * 4 * 16 * 16 * nsm * 512.0;
    printf("DMMA sustained (%d launches, %.0f ms): %.2f TFLOP/s\n", n, ms, flops / ms * 1e-9);
  }
  return 0;
}
