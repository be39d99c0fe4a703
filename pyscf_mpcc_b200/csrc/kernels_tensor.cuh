// Tensor-engine kernels for the CCSD-type residual terms (the einsum list of
// pyscf/cc/rmpccsd_slow.py:73-180 and pyscf/cc/rintermediates.py:29-130). Every einsum there is a
// permute -> GEMM (DMMA) -> permute/accumulate; these are the permute/elementwise halves. HBM-bound.
#pragma once
#include "common.cuh"

namespace mpccb {

constexpr int MAX_RANK = 6;

struct PermuteDesc {
  int rank;
  int out_dim[MAX_RANK];       // extents of the output, slowest first
  i64 in_stride[MAX_RANK];     // stride (elements) in the INPUT of the index that is output axis d
  i64 total;
};

// out[o0,o1,..] = alpha * in[perm(...)] + beta * out[...]   (beta == 0: out is not read)
// Generic gather with coalesced writes; the last TWO output axes are walked by a 32x32 shared-memory
// tile when the input's unit-stride axis is not the output's (so both sides stay coalesced).
__global__ void __launch_bounds__(256)
k_permute_axpby(const double* __restrict__ in, double* __restrict__ out, PermuteDesc d, double alpha, double beta) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < d.total; idx += stride) {
    i64 rem = idx, off = 0;
#pragma unroll
    for (int k = MAX_RANK - 1; k >= 0; --k) {
      if (k < d.rank) {
        const i64 q = rem / d.out_dim[k];
        off += (rem - q * d.out_dim[k]) * d.in_stride[k];
        rem = q;
      }
    }
    const double v = alpha * in[off];
    out[idx] = (beta == 0.0) ? v : v + beta * out[idx];
  }
}

// Tiled variant: output viewed as [outer][R][C] with C the output's fastest axis and R the output axis
// whose input stride is 1 (the input's fastest axis). One 32x32 tile per block iteration.
struct PermuteTiled {
  int rank;                    // rank of the OUTER index space (all output axes except R and C)
  int outer_dim[MAX_RANK];
  i64 outer_in_stride[MAX_RANK];
  i64 outer_out_stride[MAX_RANK];
  i64 n_outer;
  int R, C;                    // extents
  i64 r_out_stride;            // output stride of axis R (C has output stride 1)
  i64 c_in_stride;             // input stride of axis C (R has input stride 1)
};

__global__ void __launch_bounds__(256)
k_permute_tiled(const double* __restrict__ in, double* __restrict__ out, PermuteTiled d, double alpha, double beta) {
  __shared__ double tile[32][33];
  const int tiles_r = (d.R + 31) / 32, tiles_c = (d.C + 31) / 32;
  const i64 per_outer = (i64)tiles_r * tiles_c;
  const i64 total = d.n_outer * per_outer;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;      // 32 x 8
  for (i64 t = blockIdx.x; t < total; t += gridDim.x) {
    const i64 ou = t / per_outer;
    const int tt = (int)(t - ou * per_outer);
    const int r0 = (tt / tiles_c) * 32, c0 = (tt % tiles_c) * 32;
    i64 rem = ou, ioff = 0, ooff = 0;
    for (int k = d.rank - 1; k >= 0; --k) {
      const i64 q = rem / d.outer_dim[k];
      const i64 x = rem - q * d.outer_dim[k];
      ioff += x * d.outer_in_stride[k];
      ooff += x * d.outer_out_stride[k];
      rem = q;
    }
    // read: lanes along R (input unit stride)
    for (int dy = ty; dy < 32; dy += 8) {
      const int c = c0 + dy, r = r0 + tx;
      if (c < d.C && r < d.R) tile[dy][tx] = in[ioff + (i64)c * d.c_in_stride + r];
    }
    __syncthreads();
    // write: lanes along C (output unit stride)
    for (int dy = ty; dy < 32; dy += 8) {
      const int r = r0 + dy, c = c0 + tx;
      if (r < d.R && c < d.C) {
        const i64 o = ooff + (i64)r * d.r_out_stride + c;
        const double v = alpha * tile[tx][dy];
        out[o] = (beta == 0.0) ? v : v + beta * out[o];
      }
    }
    __syncthreads();
  }
}

// t1[i,a] /= eo[i] - ev[a]                      (rmpccsd_slow.py:166-169: t1new /= eia)
// t2[i,j,a,b] /= eo[i] + eo[j] - ev[a] - ev[b]  (t2new /= eijab)
__global__ void __launch_bounds__(256)
k_div_denominators(double* __restrict__ t, const double* __restrict__ eo, const double* __restrict__ ev,
                   int o, int v, int rank) {
  const i64 total = rank == 2 ? (i64)o * v : (i64)o * o * v * v;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    double den;
    if (rank == 2) {
      den = eo[idx / v] - ev[idx % v];
    } else {
      const int b = (int)(idx % v);
      i64 q = idx / v;
      const int a = (int)(q % v); q /= v;
      const int j = (int)(q % o);
      const int i = (int)(q / o);
      den = eo[i] + eo[j] - ev[a] - ev[b];
    }
    t[idx] /= den;
  }
}

// Active-space overwrite (rmpccsd_slow.py:171-178): tnew[idx] = told[idx] for the index classes selected by
// `class_mask`. A class is the tuple of active(1)/inactive(0) flags of the indices: singles bit = 2*act_h[i] + act_p[a];
// doubles bit = 8*act_h[i] + 4*act_h[j] + 2*act_p[a] + act_p[b].
__global__ void __launch_bounds__(256)
k_class_overwrite(double* __restrict__ tnew, const double* __restrict__ told, const int* __restrict__ act_h,
                  const int* __restrict__ act_p, unsigned class_mask, int o, int v, int rank) {
  const i64 total = rank == 2 ? (i64)o * v : (i64)o * o * v * v;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    int cls;
    if (rank == 2) {
      cls = 2 * act_h[idx / v] + act_p[idx % v];
    } else {
      const int b = (int)(idx % v);
      i64 q = idx / v;
      const int a = (int)(q % v); q /= v;
      const int j = (int)(q % o);
      const int i = (int)(q / o);
      cls = 8 * act_h[i] + 4 * act_h[j] + 2 * act_p[a] + act_p[b];
    }
    if ((class_mask >> cls) & 1u) tnew[idx] = told[idx];
  }
}

// partial[block] = sum_i x[i]*y[i] over a grid-stride range (fixed order; second pass k_final_sums)
__global__ void __launch_bounds__(256)
k_dot_partial(const double* __restrict__ x, const double* __restrict__ y, double* __restrict__ partial, i64 n) {
  __shared__ double scratch[32];
  double s = 0.0;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) s += x[i] * y[i];
  s = block_sum(s, scratch);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

}  // namespace mpccb

namespace mpccb {
// t[x0,..,x_{r-1}] /= sum_d sign[d] * e[d][x_d]   — orbital-energy denominators in any index order and spin
// combination (reference pyscf/cc/umpcc_lowlevel.py:468-476 D_aa/D_ab/D_bb in [a,i,b,j] order, :350-354 fock[0/1])
struct DenomDesc {
  int rank;
  int dim[4];
  const double* e[4];
  double sign[4];
  i64 total;
};
__global__ void __launch_bounds__(256)
k_div_general(double* __restrict__ t, DenomDesc d) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < d.total; idx += stride) {
    i64 rem = idx;
    double den = 0.0;
#pragma unroll
    for (int k = 3; k >= 0; --k) {
      if (k < d.rank) {
        const i64 q = rem / d.dim[k];
        den += d.sign[k] * d.e[k][rem - q * d.dim[k]];
        rem = q;
      }
    }
    t[idx] /= den;
  }
}
}  // namespace mpccb

namespace mpccb {
// out[r*ny + c] = a[idx[r]*lda + idy[c]]            (lib.take_2d, pyscf/lib/numpy_helper.py:478-504: a[idx[:,None], idy])
// lanes along c: the reads are a gather within one row of `a`, the writes contiguous
__global__ void __launch_bounds__(256)
k_take_2d(const double* __restrict__ a, i64 lda, const int* __restrict__ idx, int nx, const int* __restrict__ idy,
          int ny, double* __restrict__ out) {
  const i64 total = (i64)nx * ny;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int r = (int)(t / ny), c = (int)(t - (i64)r * ny);
    out[t] = a[(i64)idx[r] * lda + idy[c]];
  }
}

// out[idx[r]*ldo + idy[c]] += alpha * a[r*ny + c]   (lib.takebak_2d, numpy_helper.py:506-541; alpha = -1 is the
// "t2tril = -t2tril" of ccsd._unpack_4fold, ccsd.py:713-716). The index sets must hold no duplicates (they are
// triangle enumerations): every target element is touched by one thread.
__global__ void __launch_bounds__(256)
k_takebak_2d(double* __restrict__ out, i64 ldo, const double* __restrict__ a, const int* __restrict__ idx, int nx,
             const int* __restrict__ idy, int ny, double alpha) {
  const i64 total = (i64)nx * ny;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 t = (i64)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += stride) {
    const int r = (int)(t / ny), c = (int)(t - (i64)r * ny);
    const i64 off = (i64)idx[r] * ldo + idy[c];
    out[off] = fma(alpha, a[t], out[off]);
  }
}
}  // namespace mpccb
