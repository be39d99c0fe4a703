// FP64 pipe microbenchmark v2 (B200, sm_100a): GEMM-like register-tile inner loops.
//  mode 0: DMMA warp tile (MT x NT fragments of m8n8k4), operands in registers
//  mode 1: DFMA thread tile 8x8 outer product
//  mode 2: mixed — even warps DMMA, odd warps DFMA (are the pipes shared?)
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { \
  fprintf(stderr, "CUDA error %s at %s:%d\n", cudaGetErrorString(e), __FILE__, __LINE__); exit(1); } } while (0)

template <int MT, int NT>
__device__ __forceinline__ double dmma_body(int iters, double seed) {
  double c[MT][NT][2];
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j) { c[i][j][0] = 0; c[i][j][1] = 0; }
  double a[MT], b[NT];
#pragma unroll
  for (int i = 0; i < MT; ++i) a[i] = seed + i;
#pragma unroll
  for (int j = 0; j < NT; ++j) b[j] = seed - j;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < MT; ++i)
#pragma unroll
      for (int j = 0; j < NT; ++j)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
                     : "+d"(c[i][j][0]), "+d"(c[i][j][1]) : "d"(a[i]), "d"(b[j]));
#pragma unroll
    for (int i = 0; i < MT; ++i) a[i] += 1e-30;   // keep operands live / changing like a k-loop
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j) s += c[i][j][0] + c[i][j][1];
  return s;
}

__device__ __forceinline__ double dfma_body(int iters, double seed) {
  double c[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) c[i][j] = 0;
  double a[8], b[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { a[i] = seed + i; b[i] = seed - i; }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) c[i][j] = fma(a[i], b[j], c[i][j]);
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] += 1e-30;
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) s += c[i][j];
  return s;
}

template <int MT, int NT>
__global__ void __launch_bounds__(256) k_dmma(double* out, int iters) {
  out[blockIdx.x * blockDim.x + threadIdx.x] = dmma_body<MT, NT>(iters, 1.0 + threadIdx.x * 1e-9);
}
__global__ void __launch_bounds__(256) k_dfma(double* out, int iters) {
  out[blockIdx.x * blockDim.x + threadIdx.x] = dfma_body(iters, 1.0 + threadIdx.x * 1e-9);
}
__global__ void __launch_bounds__(256) k_mixed(double* out, int it_dmma, int it_dfma) {
  int w = threadIdx.x >> 5;
  double r;
  if (w & 1) r = dfma_body(it_dfma, 1.0 + threadIdx.x * 1e-9);
  else r = dmma_body<4, 4>(it_dmma, 1.0 + threadIdx.x * 1e-9);
  out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <typename F>
float timeit(F f) {
  cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
  f(); CK(cudaDeviceSynchronize());
  float best = 1e30f;
  for (int r = 0; r < 3; ++r) {
    CK(cudaEventRecord(e0)); f(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
    float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
  }
  return best;
}

int main() {
  cudaDeviceProp p; CK(cudaGetDeviceProperties(&p, 0));
  int nsm = p.multiProcessorCount; double clk = p.clockRate * 1e3;
  double* out; CK(cudaMalloc(&out, sizeof(double) * nsm * 4 * 512));
  printf("%s %d SMs\n", p.name, nsm);
  auto report = [&](const char* name, int warps, float ms, double fma_total) {
    double tf = 2 * fma_total / ms * 1e-9;
    printf("%-28s warps/SM=%2d  %8.3f ms  %6.2f TFLOP/s  %.1f FMA/clk/SM @1965\n", name, warps, ms, tf,
           fma_total / nsm / (ms * 1e-3 * clk));
  };
  int wl[] = {4, 8};
  for (int w : wl) {
    int it = 40000;
    { float ms = timeit([&] { k_dmma<8, 4><<<nsm, w * 32>>>(out, it / 2); });
      report("DMMA 8x4 frags (64x32 tile)", w, ms, (double)(it / 2) * 32 * 256 * w * nsm); }
    { float ms = timeit([&] { k_dmma<4, 4><<<nsm, w * 32>>>(out, it); });
      report("DMMA 4x4 frags (32x32 tile)", w, ms, (double)it * 16 * 256 * w * nsm); }
    { float ms = timeit([&] { k_dmma<4, 2><<<nsm, w * 32>>>(out, it * 2); });
      report("DMMA 4x2 frags (32x16 tile)", w, ms, (double)it * 2 * 8 * 256 * w * nsm); }
    { float ms = timeit([&] { k_dfma<<<nsm, w * 32>>>(out, it * 2); });
      report("DFMA 8x8 thread tile", w, ms, (double)it * 2 * 64 * 32 * w * nsm); }
  }
  // two CTAs per SM of 8 warps
  { int it = 40000; float ms = timeit([&] { k_dmma<4, 4><<<nsm * 2, 8 * 32>>>(out, it); });
    report("DMMA 4x4, 2 CTA/SM x 8 warps", 16, ms, (double)it * 16 * 256 * 8 * nsm * 2); }
  { int it = 40000; float ms = timeit([&] { k_dmma<4, 4><<<nsm, 16 * 32>>>(out, it); });
    report("DMMA 4x4, 16 warps", 16, ms, (double)it * 16 * 256 * 16 * nsm); }
  { int it = 20000; float ms = timeit([&] { k_dmma<8, 8><<<nsm, 4 * 32>>>(out, it); });
    report("DMMA 8x8 frags (64x64), 4 warps", 4, ms, (double)it * 64 * 256 * 4 * nsm); }
  // mixed: 8 warps (4 DMMA + 4 DFMA) and 16 warps
  for (int w : {8}) {
    int itm = 40000, itf = 40000 * 2;
    // balance so both halves take similar time at their solo rates (DMMA 4x4: 16*256 FMA/iter/warp; DFMA: 64*32)
    float ms = timeit([&] { k_mixed<<<nsm, w * 32>>>(out, itm, itf * 2); });
    double fma = ((double)itm * 16 * 256 + (double)itf * 2 * 64 * 32) * (w / 2) * nsm;
    report("mixed DMMA+DFMA", w, ms, fma);
  }
  return 0;
}
