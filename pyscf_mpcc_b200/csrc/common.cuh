// Shared device/host helpers for the DF-MPCC sm_100a kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

namespace mpccb {

typedef long long i64;

// ---------------------------------------------------------------------------
// FP64 tensor-core atom. On sm_100a every mma.sync f64 shape lowers to SASS
// DMMA.8x8x4 (checked with cuobjdump), so the m8n8k4 form is used directly.
//   A (8x4, row):  a  = A[lane>>2][lane&3]
//   B (4x8, col):  b  = B[lane&3][lane>>2]
//   C (8x8):       c0 = C[lane>>2][2*(lane&3)], c1 = C[lane>>2][2*(lane&3)+1]
// ---------------------------------------------------------------------------
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1},{%2},{%3},{%0,%1};"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
  return v;
}

// Deterministic block sum (fixed tree). `scratch` holds >= 32 doubles of shared memory.
__device__ __forceinline__ double block_sum(double v, double* scratch) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) scratch[w] = v;
  __syncthreads();
  const int nw = (blockDim.x + 31) >> 5;
  double r = 0.0;
  if (w == 0) {
    r = (lane < nw) ? scratch[lane] : 0.0;
    r = warp_sum(r);
  }
  return r;  // valid in warp 0
}

// streaming (read-once) 8-byte load that does not allocate in L1
__device__ __forceinline__ double ld_stream(const double* p) {
  double v;
  asm volatile("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(v) : "l"(p));
  return v;
}

}  // namespace mpccb
