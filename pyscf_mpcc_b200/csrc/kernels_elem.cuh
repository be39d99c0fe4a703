// Memory-bound kernels of the DF-MPCC iteration: layout changes, the fused
// dressing finalisation, reductions (deterministic: fixed trees, no atomics),
// the T1 update and the DIIS vector algebra. All are HBM-bandwidth bound; they
// are written for coalesced 8-byte (or wider) accesses with lanes along the
// unit-stride index and warp-shuffle reductions.
#pragma once
#include "common.cuh"
#include "gemm_dmma.cuh"   // SymInfo (mirror scheme)

namespace mpccb {

// out[b][c][r] = in[b][r][c]      (Lov -> Lvo, t1 -> t1T)
// Address of flat element f of the LOGICAL [a][b][i][j] t2 tensor in the stored buffer, whose j pitch is nocc rounded
// up to even (pad = 0 or 1): every row of nocc elements is followed by `pad` unused ones.
__device__ __forceinline__ i64 t2_addr(i64 f, int o, int pad) { return pad ? f + (f / o) * pad : f; }

__global__ void __launch_bounds__(256)
k_transpose_batched(const double* __restrict__ in, double* __restrict__ out, int R, int C, int ldo) {
  // ldo >= R: pitch of the output rows (Lvo keeps its j index on an even pitch)
  __shared__ double tile[32][33];
  const i64 boff = (i64)blockIdx.z * R * C, boff_out = (i64)blockIdx.z * ldo * C;
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int dy = threadIdx.y; dy < 32; dy += 8) {
    const int r = r0 + dy, c = c0 + threadIdx.x;
    if (r < R && c < C) tile[dy][threadIdx.x] = in[boff + (i64)r * C + c];
  }
  __syncthreads();
  for (int dy = threadIdx.y; dy < 32; dy += 8) {
    const int c = c0 + dy, r = r0 + threadIdx.x;
    if (r < R && c < C) out[boff_out + (i64)c * ldo + r] = tile[threadIdx.x][dy];
  }
}

// X[L] = sum_i Xoo[L,i,i]   ( == einsum("Lia,ia->L", Lov, t1), reference :121 )
__global__ void __launch_bounds__(128)
k_trace(const double* __restrict__ Xoo, double* __restrict__ X, int L, int o) {
  const int l = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (l >= L) return;
  const int lane = threadIdx.x & 31;
  double s = 0.0;
  for (int i = lane; i < o; i += 32) s += Xoo[((i64)l * o + i) * o + i];
  s = warp_sum(s);
  if (lane == 0) X[l] = s;
}

// Fused end of the T1 dressing (reference :124, :126-127 integrand, :146).
// Inputs, all [L][i][a]:  Xov = (Lvv.t1)^T, Lov, JT = Joo.t1, XT = Xoo.t1
//   J  = Xov + Lov - JT                      -> JK [(a,i)][ldk]   (K-major for the eris GEMM)
//   e1 = XT + X[L]*t1[i,a] + J               -> E1K[(a,i)][ldk]
//   W  = XT - Xov                            -> Wov[L][i][a]
// grid (ceil(v/32), ceil(L/32), o), block (32, 8)
// Where the dressed integrals go. One rank: the local JK buffer. Several ranks, fused all-gather: every
// rank's JK buffer lives in symmetric (peer-mapped) memory and the kernel stores its slab's rows straight
// into all of them over NVLink — either one store per peer pointer, or a single store to the NVSwitch
// multicast address that the switch replicates to every rank (n == 1, p[0] = multicast pointer).
struct JkTargets {
  double* p[16];
  int n;
};

__global__ void __launch_bounds__(256)
k_dress_finalize(const double* __restrict__ Xov, const double* __restrict__ Lov,
                 const double* __restrict__ JT, const double* __restrict__ XT,
                 const double* __restrict__ X, const double* __restrict__ t1,
                 JkTargets jk, double* __restrict__ E1K, double* __restrict__ Wov,
                 int L, int o, int v, int ldk, int a_lo, int a_hi) {
  // only the virtual columns a in [a_lo, a_hi) (this rank's slab) are produced; grid.x covers the slab
  __shared__ double tj[32][33];
  __shared__ double te[32][33];
  const int i = blockIdx.z;
  const int a0 = a_lo + blockIdx.x * 32, l0 = blockIdx.y * 32;
  const int a = a0 + threadIdx.x;
  const double t1ia = (a < a_hi) ? t1[(i64)i * v + a] : 0.0;
  for (int dy = threadIdx.y; dy < 32; dy += 8) {
    const int l = l0 + dy;
    if (l < L && a < a_hi) {
      const i64 off = ((i64)l * o + i) * v + a;
      const double xov = Xov[off], xt = XT[off];
      const double j = xov + Lov[off] - JT[off];
      Wov[off] = xt - xov;
      tj[dy][threadIdx.x] = j;
      te[dy][threadIdx.x] = xt + X[l] * t1ia + j;
    }
  }
  __syncthreads();
  for (int dy = threadIdx.y; dy < 32; dy += 8) {
    const int aa = a0 + dy, l = l0 + threadIdx.x;
    if (aa < a_hi && l < L) {
      const i64 off = ((i64)aa * o + i) * ldk + l;
      const double jv = tj[threadIdx.x][dy];
      for (int t = 0; t < jk.n; ++t) jk.p[t][off] = jv;      // local, peer (NVLink) or multicast store
      E1K[off] = te[threadIdx.x][dy];
    }
  }
}

// y[r] = sum_k M[r*ld + k] * x[k]; one warp per row
__global__ void __launch_bounds__(256)
k_gemv_rows(const double* __restrict__ M, const double* __restrict__ x, double* __restrict__ y,
            int rows, int K, i64 ld) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const double* row = M + (i64)r * ld;
  double s = 0.0;
  for (int k = lane; k < K; k += 32) s += row[k] * x[k];
  s = warp_sum(s);
  if (lane == 0) y[r] = s;
}

// partial[s][n] = sum_{k in slice s} A[k*N + n] * x[k]: y = A^T x for a matrix stored [K][N] (lanes along n:
// coalesced); grid (ceil(N/256), slices), reduced afterwards in fixed order (k_reduce_splits)
__global__ void __launch_bounds__(256)
k_gemv_cols_partial(const double* __restrict__ A, const double* __restrict__ x, double* __restrict__ partial,
                    i64 K, i64 N, i64 kchunk) {
  const i64 n = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const i64 k0 = (i64)blockIdx.y * kchunk, k1 = min(K, k0 + kchunk);
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  i64 k = k0;
  for (; k + 4 <= k1; k += 4) {
    s0 = fma(A[k * N + n], x[k], s0);
    s1 = fma(A[(k + 1) * N + n], x[k + 1], s1);
    s2 = fma(A[(k + 2) * N + n], x[k + 2], s2);
    s3 = fma(A[(k + 3) * N + n], x[k + 3], s3);
  }
  for (; k < k1; ++k) s0 = fma(A[k * N + n], x[k], s0);
  partial[(i64)blockIdx.y * N + n] = (s0 + s1) + (s2 + s3);
}

// w[r0 + r] = sum_k M[(r0 + r)*ld + k] * x[k] and f[r0 + r] = w[r0 + r] - F2[r0 + r] for the slab's rows, both stored
// into every target buffer (peer pointers over NVLink or the multicast address) at offsets w_off / f_off: the slab's
// share of w = Jvo.X (:128) and of Fvo (:133) reaches all ranks with the dressed integrals, so that no rank needs the
// other slabs' JK rows or F2 rows for them. One warp per row, four independent partial sums per lane.
__global__ void __launch_bounds__(256)
k_gemv_rows_bcast(const double* __restrict__ M, const double* __restrict__ x, const double* __restrict__ F2,
                  JkTargets y, i64 w_off, i64 f_off, int r0, int rows, int K, i64 ld) {
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= rows) return;
  const int lane = threadIdx.x & 31;
  const double* row = M + (i64)(r0 + r) * ld;
  double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;
  int k = lane;
  for (; k + 96 < K; k += 128) {
    s0 = fma(row[k], x[k], s0);
    s1 = fma(row[k + 32], x[k + 32], s1);
    s2 = fma(row[k + 64], x[k + 64], s2);
    s3 = fma(row[k + 96], x[k + 96], s3);
  }
  for (; k < K; k += 32) s0 = fma(row[k], x[k], s0);
  const double s = warp_sum((s0 + s1) + (s2 + s3));
  if (lane == 0) {
    const double f = s - F2[r0 + r];
    for (int t = 0; t < y.n; ++t) {
      y.p[t][w_off + r0 + r] = s;
      y.p[t][f_off + r0 + r] = f;
    }
  }
}

// Copies base[off .. off+n) to the same offsets of every target buffer (peer pointers over NVLink, or the multicast
// address, which also rewrites the — identical — local copy): the L slab of Xoo a rank has just computed, for
// everybody. Eight loads in flight per thread.
__global__ void __launch_bounds__(256)
k_bcast_rows(const double* __restrict__ base, JkTargets tg, i64 off, i64 n) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i0 = (i64)blockIdx.x * blockDim.x + threadIdx.x; i0 < n; i0 += 8 * stride) {
    double a[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const i64 i = i0 + u * stride;
      a[u] = i < n ? base[off + i] : 0.0;
    }
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      const i64 i = i0 + u * stride;
      if (i < n)
        for (int t = 0; t < tg.n; ++t) tg.p[t][off + i] = a[u];
    }
  }
}

// out[i] = x[i] + y[i] for i in [0, skip0) and [skip1, n): Joo = Xoo + Loo (:123) for the L slabs that arrived from
// the other ranks (the slab [skip0, skip1) computed here got its Joo rows from the GEMM epilogue)
__global__ void __launch_bounds__(256)
k_add_outside(const double* __restrict__ x, const double* __restrict__ y, double* __restrict__ out,
              i64 skip0, i64 skip1, i64 n) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  const i64 m = n - (skip1 - skip0);
  for (i64 q = (i64)blockIdx.x * blockDim.x + threadIdx.x; q < m; q += stride) {
    const i64 i = q < skip0 ? q : q + (skip1 - skip0);
    out[i] = x[i] + y[i];
  }
}

// dst[i] = beta*dst[i] + alpha * sum_s P[s*stride + i]   (split-K / tile partial reduction, fixed order)
__global__ void __launch_bounds__(256)
k_reduce_splits(double* __restrict__ dst, const double* __restrict__ P, int nsplit, i64 n, i64 stride,
                double alpha, double beta) {
  const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0.0;
  for (int k = 0; k < nsplit; ++k) s += P[(i64)k * stride + i];
  dst[i] = (beta == 0.0 ? 0.0 : beta * dst[i]) + alpha * s;
}

// Fvo[b][j] = w[b*o + j] - F2[b][j]        (reference :133, stored transposed)
__global__ void __launch_bounds__(256)
k_make_fvo(const double* __restrict__ w, const double* __restrict__ F2, double* __restrict__ Fvo, i64 n) {
  const i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) Fvo[i] = w[i] - F2[i];
}

// OmLoc[a][i] += sum_j t1[j][a] * Z[j][i]     (second half of Jvv.Yvo, reference :141-142)
__global__ void __launch_bounds__(256)
k_add_t1T_Z(double* __restrict__ Om, const double* __restrict__ t1, const double* __restrict__ Z,
            int o, int v) {
  const i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (i64)o * v) return;
  const int a = (int)(idx / o), i = (int)(idx % o);
  double s = 0.0;
  for (int j = 0; j < o; ++j) s += t1[(i64)j * v + a] * Z[(i64)j * o + i];
  Om[idx] += s;
}

// partial[block] = sum over a grid-stride range of x[r*ldx + k]*y[r*ldy + k], rows r in [0,rows), k in [0,K)
__global__ void __launch_bounds__(256)
k_dot_rows_partial(const double* __restrict__ x, const double* __restrict__ y, double* __restrict__ partial,
                   int rows, int K, i64 ldx, i64 ldy) {
  __shared__ double scratch[32];
  double s = 0.0;
  const int wpb = blockDim.x >> 5, lane = threadIdx.x & 31;
  for (int r = blockIdx.x * wpb + (threadIdx.x >> 5); r < rows; r += gridDim.x * wpb) {
    const double* xr = x + (i64)r * ldx;
    const double* yr = y + (i64)r * ldy;
    for (int k = lane; k < K; k += 32) s += xr[k] * yr[k];
  }
  s = block_sum(s, scratch);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

// out[q] = sum_b partial[b*nq + q]   (single block; fixed order)
__global__ void __launch_bounds__(256)
k_final_sums(const double* __restrict__ partial, double* __restrict__ out, int nblocks, int nq) {
  __shared__ double scratch[32];
  for (int q = 0; q < nq; ++q) {
    double s = 0.0;
    for (int b = threadIdx.x; b < nblocks; b += blockDim.x) s += partial[(i64)b * nq + q];
    s = block_sum(s, scratch);
    if (threadIdx.x == 0) out[q] = s;
    __syncthreads();
  }
}

// Singles residual assembly + T1 update (reference :128-131, :149-151).
//   Om[a,i]  = [OmRep[a,i] if given] + w[(a,i)] + OmLoc[a,i] + sum_b t1[i,b] fvv[b,a] - sum_k t1[k,a] foo[i,k]
//   t1n[i,a] = -Om[a,i]/(ev[a]-eo[i]);  partial[block] = sum (t1[i,a] - t1n[i,a])^2
// grid (ceil(v/32), o), 256 threads: a CTA makes 32 consecutive a of one i; warp p sums the terms b = p (mod 8) and
// k = p (mod 8) of the two contractions (lanes along a: t1 / fvv rows coalesced), warp 0 adds the eight parts in
// fixed order. The kernel is replicated on every rank — its latency is serial time at 8 GPUs — hence the split of
// each 576-term sum over eight warps rather than one thread per element.
__global__ void __launch_bounds__(256)
k_t1_update(const double* __restrict__ OmRep, const double* __restrict__ w, const double* __restrict__ OmLoc,
            const double* __restrict__ t1, const double* __restrict__ fvv, const double* __restrict__ foo,
            const double* __restrict__ eo, const double* __restrict__ ev,
            double* __restrict__ t1n, double* __restrict__ OmOut, double* __restrict__ partial, int o, int v) {
  __shared__ double part[8][33];
  const int x = threadIdx.x & 31, p = threadIdx.x >> 5;
  const int i = blockIdx.y, a = blockIdx.x * 32 + x;
  double acc = 0.0;
  if (a < v) {
    const double* t1i = t1 + (i64)i * v;
    double s0 = 0.0, s1 = 0.0;
    int b = p;
    for (; b + 8 < v; b += 16) {
      s0 = fma(t1i[b], fvv[(i64)b * v + a], s0);
      s1 = fma(t1i[b + 8], fvv[(i64)(b + 8) * v + a], s1);
    }
    if (b < v) s0 = fma(t1i[b], fvv[(i64)b * v + a], s0);
    double u = 0.0;
    for (int k = p; k < o; k += 8) u = fma(t1[(i64)k * v + a], foo[(i64)i * o + k], u);
    acc = (s0 + s1) - u;
  }
  part[p][x] = acc;
  __syncthreads();
  if (p != 0) return;
  double r2 = 0.0;
  if (a < v) {
    const i64 ai = (i64)a * o + i, idx = (i64)i * v + a;
    double om = (OmRep ? OmRep[ai] : 0.0) + w[ai] + OmLoc[ai];
    double s = 0.0;
#pragma unroll
    for (int q = 0; q < 8; ++q) s += part[q][x];
    om += s;
    const double q = om / (ev[a] - eo[i]);
    t1n[idx] = -q;
    OmOut[ai] = om;
    const double d = t1[idx] + q;
    r2 = d * d;
  }
  r2 = warp_sum(r2);
  if (x == 0) partial[(i64)blockIdx.y * gridDim.x + blockIdx.x] = r2;
}

// ---------------------------------------------------------------------------
// DIIS vector algebra. The vector is x = [ t1 (nov) | rows r of the t2 buffer
// viewed as (nov x nov), columns 0..r ] (reference dfrmpcc_lowlevel.py:162-168
// fed the (v,o,v,o) tensor == our [a,b,i,j] buffer read row-major).
// A rank owning rows [r0, r1) of that matrix holds the matching packed segment
//   seg = x[nov + r0(r0+1)/2 : nov + r1(r1+1)/2].
// ---------------------------------------------------------------------------
struct DiisPtrs {
  const double* e[6];
};

// The previous extrapolated vector, never stored: xprev = sum_k c[k] * x[k] over the n stored trial vectors in slot
// order — the arithmetic of k_diis_combine (diis.py:268-275), recomputed per element inside the push. n == 0: the
// explicit `xprev` pointer is used (second call of a solve: xprev is the first vector). Folding the combination into
// the push removes one write and one read of the whole vector per iteration.
struct DiisPrev {
  const double* x[6];
  double c[6];
  int n;
};

__device__ __forceinline__ double diis_prev_value(const DiisPrev& pv, const double* xprev, i64 p) {
  if (pv.n == 0) return xprev[p];
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < 6; ++k)
    if (k < pv.n) s = fma(pv.x[k][p], pv.c[k], s);
  return s;
}

// Chunked push. The t2 part of the local vector (packed rows r0 .. r0+nrows of the (nov x nov) view of T) is cut into
// G equal contiguous chunks of the PACKED index; CTA b streams chunk b with all vector streams (x_k, e_k, the new x
// and e) contiguous and coalesced, and the matching T row prefixes contiguous up to the row breaks. A thread finds
// (row, column) of its first element with one square root and then walks: column += 256, carrying into the next rows.
// CTA G handles the t1 part (skipped when t1new == nullptr).
//   xs[p] = value;  es[p] = value - xprev[p];  dots[k] += es . e_k  (k < nd; e_slot is `es` itself)
// xs / es may alias pv.x[slot] / eptr.e[slot] (the slot being replaced): every element is read before it is
// written, by the same thread — hence no __restrict__ on them.
// partial layout: [G + 1][7] (slot k -> column k; column 6 unused)
__global__ void __launch_bounds__(256)
k_diis_push(const double* __restrict__ T, const double* __restrict__ t1new,
            const double* xprev, DiisPrev pv, double* xs, double* es,
            DiisPtrs eptr, int nd, int slot, int first_call, int count_t1,
            i64 nov, i64 r0, int nrows, i64 chunk, int G, int o, int pad, double* __restrict__ partial) {
  __shared__ double scratch[32];
  const int blk = blockIdx.x;
  double d[6] = {0, 0, 0, 0, 0, 0};
  auto one = [&](double val, i64 p) {
    if (first_call) {
      xs[p] = val;                       // becomes xprev
    } else {
      const double e = val - diis_prev_value(pv, xprev, p);
      double other[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) other[k] = (k < nd && k != slot) ? eptr.e[k][p] : 0.0;
      xs[p] = val;
      es[p] = e;
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (k < nd) d[k] = fma(e, (k == slot) ? e : other[k], d[k]);
    }
  };
  if (blk < G) {
    const i64 tri0 = (r0 * (r0 + 1)) / 2;
    const i64 total = ((r0 + nrows) * (r0 + nrows + 1)) / 2 - tri0;      // local packed t2 elements
    const i64 q0 = (i64)blk * chunk;
    const i64 q1 = q0 + chunk < total ? q0 + chunk : total;
    i64 q = q0 + threadIdx.x;
    if (q < q1) {
      // global packed position -> (r, c): r = floor((sqrt(8P + 1) - 1) / 2), fixed up for rounding
      const i64 P = tri0 + q;
      i64 r = (i64)((sqrt(8.0 * (double)P + 1.0) - 1.0) * 0.5);
      while ((r * (r + 1)) / 2 > P) --r;
      while (((r + 1) * (r + 2)) / 2 <= P) ++r;
      i64 c = P - (r * (r + 1)) / 2;
      for (; q < q1; q += 256) {
        one(T[t2_addr((r - r0) * nov + c, o, pad)], nov + q);
        c += 256;
        while (c > r) { c -= r + 1; ++r; }
      }
    }
  } else if (t1new != nullptr) {
    for (i64 c = threadIdx.x; c < nov; c += blockDim.x) one(t1new[c], c);
    if (!count_t1) {                     // replicated t1 part counted by one rank only
#pragma unroll
      for (int k = 0; k < 6; ++k) d[k] = 0.0;
    }
  }
  if (!first_call) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      if (k < nd) {
        const double s = block_sum(d[k], scratch);
        if (threadIdx.x == 0) partial[(i64)blk * 7 + k] = s;
      }
    }
  }
}

// The same update for a plain device vector of n elements (the CCSD-type stage packs its own vector,
// ccsd.py:719-726): xs = xnew, es = xnew - xprev, dots[k] += es . e_k. Grid-stride; partial layout [gridDim.x][7].
__global__ void __launch_bounds__(256)
k_vec_diis_push(const double* __restrict__ xnew, const double* xprev, DiisPrev pv, double* xs, double* es,
                DiisPtrs eptr, int nd, int slot, int first_call, i64 n, double* __restrict__ partial) {
  __shared__ double scratch[32];
  double d[6] = {0, 0, 0, 0, 0, 0};
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 p = (i64)blockIdx.x * blockDim.x + threadIdx.x; p < n; p += stride) {
    const double val = xnew[p];
    if (first_call) {
      xs[p] = val;
    } else {
      const double e = val - diis_prev_value(pv, xprev, p);
      double other[6];
#pragma unroll
      for (int k = 0; k < 6; ++k) other[k] = (k < nd && k != slot) ? eptr.e[k][p] : 0.0;
      xs[p] = val;
      es[p] = e;
#pragma unroll
      for (int k = 0; k < 6; ++k)
        if (k < nd) d[k] = fma(e, (k == slot) ? e : other[k], d[k]);
    }
  }
  if (!first_call) {
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      if (k < nd) {
        const double s = block_sum(d[k], scratch);
        if (threadIdx.x == 0) partial[(i64)blockIdx.x * 7 + k] = s;
      }
    }
  }
}

// packed[r(r+1)/2 + c] = M[r*n + c], c <= r: lib.pack_tril of an (n x n) matrix (NPdpack_tril,
// pyscf/lib/np_helper/pack_tril.c:99-107). One block per row.
__global__ void __launch_bounds__(256)
k_pack_tril(const double* __restrict__ M, double* __restrict__ packed, i64 n) {
  const i64 r = blockIdx.x;
  const double* src = M + r * n;
  double* dst = packed + (r * (r + 1)) / 2;
  for (i64 c = threadIdx.x; c <= r; c += blockDim.x) dst[c] = src[c];
}

// ---------------------------------------------------------------------------
// Column-owned form of the same vector (several ranks). Row r of the packed triangle has r+1 elements, so a
// row-sharded vector gives the last of P ranks (2P-1) times the bytes of the first one and everybody waits for it at
// the all-reduce of the dots. But t2[A,B,I,J] = t2[B,A,J,I], and a rank's slab T_loc[al][b][i][j] = t2[a0+al, b, i, j]
// therefore also holds every element whose SECOND virtual index lies in the slab: t2[A=b, B=a0+al, I=j, J=i].
// Ownership by B is uniform in the row index, so every rank gets 1/P of the triangle (within 0.2 % at the
// target shape) from local memory, without communication. The element set, hence every dot product and the
// extrapolation, is that of the reference vector; only the storage order differs:
//   local block k = al * v + b (memory order) holds the flat reference indices f = F0 + t, F0 = (b*v + a0+al) * o^2,
//   t = j*o + i; f is kept iff (f mod nov) <= (f div nov); kept elements are stored in t order at
//   prefix[k] + below(F0 + t) - below(F0), below(f) = r(r+1)/2 + min(c, r+1) = packed elements before flat f.
// ---------------------------------------------------------------------------
__device__ __forceinline__ i64 tril_below(i64 r, i64 c) { return (r * (r + 1)) / 2 + (c < r + 1 ? c : r + 1); }

// grid: G CTAs walk the local blocks k = blockIdx.x, blockIdx.x + G, ..; CTA G handles the t1 part.
// A block is walked in t order, so that all vector streams (xprev, up to 5 older error vectors, x and e) are
// contiguous per warp; the slab block itself is read with stride o (the 4 warps that share a 32-byte sector issue
// together, so it is served by L1), 4 elements per thread in flight. Measured against two alternatives (staging the
// block through shared memory; 8 x 4 thread tiles with sector-exact slab reads) this form was the fastest; it still
// runs about 25 % below the row-owned kernel per byte.
// xprev is always an explicit stored vector here: the on-the-fly combination of the row-owned kernel costs this
// one (bound by its address arithmetic and the strided slab read, not by HBM) more than the combine pass saves
// (two ranks: 8.3 ms against 3.2 + 1.8 ms).
// scatter_to != nullptr: instead of the DIIS update, copy the local vector `xprev` into a full reference-order
// packed vector (host-facing gathers; the caller all-reduces the zero-initialised targets).
__global__ void __launch_bounds__(256)
k_diis_push_cols(const double* __restrict__ T, const double* __restrict__ t1new,
                 const double* __restrict__ xprev, double* __restrict__ xs, double* __restrict__ es,
                 DiisPtrs eptr, int nd, int slot, int first_call, int count_t1,
                 int o, int jp, int v, int a0, int na, const i64* __restrict__ prefix, int G,
                 double* __restrict__ partial, double* __restrict__ scatter_to) {
  __shared__ double scratch[32];
  constexpr int U = 4;
  const i64 nov = (i64)o * v;
  const int oo = o * o;
  double d[6] = {0, 0, 0, 0, 0, 0};
  if ((int)blockIdx.x < G) {
    const i64 nblocks = (i64)na * v;
    for (i64 k = blockIdx.x; k < nblocks; k += G) {
      const i64 cnt = prefix[k + 1] - prefix[k];
      if (cnt == 0) continue;                                    // block entirely above the diagonal
      const int al = (int)(k / v), b = (int)(k - (i64)al * v);
      const i64 F0 = ((i64)b * v + (a0 + al)) * oo;
      const i64 R0 = F0 / nov, C0 = F0 - R0 * nov;
      const i64 base = nov + prefix[k] - tril_below(R0, C0);
      const double* blk = T + k * o * jp;
      for (int t0 = threadIdx.x; t0 < oo; t0 += 256 * U) {
        i64 pos[U], ref[U];
        double val[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int t = t0 + 256 * u;
          ok[u] = false; pos[u] = 0; ref[u] = 0; val[u] = 0.0;
          if (t < oo) {
            i64 c = C0 + t, r = R0;
            while (c >= nov) { c -= nov; ++r; }
            if (c <= r) {
              const int j = t / o, i = t - j * o;                // reference (I, J) = (j, i) of the slab block
              ok[u] = true;
              ref[u] = (r * (r + 1)) / 2 + c;
              pos[u] = base + ref[u];
              val[u] = blk[i * jp + j];
            }
          }
        }
        if (scatter_to) {
#pragma unroll
          for (int u = 0; u < U; ++u) if (ok[u]) scatter_to[ref[u]] = xprev[pos[u]];
        } else if (first_call) {
#pragma unroll
          for (int u = 0; u < U; ++u) if (ok[u]) xs[pos[u]] = val[u];
        } else {
          double e[U];
#pragma unroll
          for (int u = 0; u < U; ++u) e[u] = ok[u] ? val[u] - xprev[pos[u]] : 0.0;
#pragma unroll
          for (int q = 0; q < 6; ++q) {
            if (q < nd) {
#pragma unroll
              for (int u = 0; u < U; ++u) {
                const double other = (q == slot) ? e[u] : (ok[u] ? eptr.e[q][pos[u]] : 0.0);
                d[q] += e[u] * other;
              }
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) if (ok[u]) { xs[pos[u]] = val[u]; es[pos[u]] = e[u]; }
        }
      }
    }
  } else if (t1new != nullptr && !scatter_to) {
    for (i64 c = threadIdx.x; c < nov; c += blockDim.x) {
      const double val = t1new[c];
      if (first_call) {
        xs[c] = val;
      } else {
        const double e = val - xprev[c];
        double other[6];
#pragma unroll
        for (int q = 0; q < 6; ++q) other[q] = (q < nd && q != slot) ? eptr.e[q][c] : 0.0;
        xs[c] = val;
        es[c] = e;
#pragma unroll
        for (int q = 0; q < 6; ++q)
          if (q < nd) d[q] += e * ((q == slot) ? e : other[q]);
      }
    }
    if (!count_t1) {
#pragma unroll
      for (int q = 0; q < 6; ++q) d[q] = 0.0;
    }
  }
  if (!first_call && !scatter_to) {
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      if (q < nd) {
        const double s = block_sum(d[q], scratch);
        if (threadIdx.x == 0) partial[(i64)blockIdx.x * 7 + q] = s;
      }
    }
  }
}

// xnew[i] = sum_k c[k] * x_k[i]  (slot order, like diis.py:268-275)
struct DiisComb {
  const double* x[6];
  double c[6];
};
__global__ void __launch_bounds__(256)
k_diis_combine(DiisComb cb, int nd, double* __restrict__ xnew, i64 n) {
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 i = (i64)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < 6; ++k)
      if (k < nd) s = fma(cb.x[k][i], cb.c[k], s);     // same arithmetic as diis_prev_value
    xnew[i] = s;
  }
}

// t2_out[i][j][a][b] = S[(i*v + a), (j*v + b)],  S = symmetric fill of the packed lower triangle
// (reference vector_to_amplitudes :170-176 + NPdunpack_tril/NPdsymm_triu)
__global__ void __launch_bounds__(256)
k_unpack_oovv(const double* __restrict__ packed, double* __restrict__ out, int o, int v) {
  const i64 nov = (i64)o * v;
  const i64 total = nov * nov;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int b = (int)(idx % v);
    i64 q = idx / v;
    const int a = (int)(q % v); q /= v;
    const int j = (int)(q % o);
    const int i = (int)(q / o);
    i64 r = (i64)i * v + a, c = (i64)j * v + b;
    if (c > r) { const i64 t = r; r = c; c = t; }
    out[idx] = packed[r * (r + 1) / 2 + c];
  }
}

// out[a][i][b][j] = T[a][b][i][j]   (the (v,o,v,o) tensor updated_amps returns, reference :137,:152)
__global__ void __launch_bounds__(256)
k_abij_to_aibj(const double* __restrict__ T, double* __restrict__ out, int o, int jp, int v, int na) {
  const i64 total = (i64)na * v * o * o;
  const i64 stride = (i64)gridDim.x * blockDim.x;
  for (i64 idx = (i64)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += stride) {
    const int j = (int)(idx % o);
    i64 q = idx / o;
    const int b = (int)(q % v); q /= v;
    const int i = (int)(q % o);
    const int a = (int)(q / o);
    out[idx] = T[(((i64)a * v + b) * o + i) * jp + j];
  }
}

// plain row-prefix pack (amplitudes_to_vector without the DIIS bookkeeping)
__global__ void __launch_bounds__(256)
k_pack_rows(const double* __restrict__ T, double* __restrict__ seg, i64 nov, i64 r0, int o, int pad) {
  const i64 r = r0 + blockIdx.x;
  const i64 f0 = (i64)blockIdx.x * nov;
  double* dst = seg + (r * (r + 1)) / 2 - (r0 * (r0 + 1)) / 2;
  for (i64 c = threadIdx.x; c <= r; c += blockDim.x) dst[c] = T[t2_addr(f0 + c, o, pad)];
}

}  // namespace mpccb

namespace mpccb {
// out[q] = sum_b partial[b*ld + q] for q < nq  (one block per q; fixed order). The DIIS dots wait for this on the
// host, so the nq sums run side by side instead of one after the other.
__global__ void __launch_bounds__(256)
k_final_sums_strided(const double* __restrict__ partial, double* __restrict__ out, int nblocks, int ld, int nq) {
  __shared__ double scratch[32];
  const int q = blockIdx.x;
  if (q >= nq) return;
  double s0 = 0.0, s1 = 0.0;
  int b = threadIdx.x;
  for (; b + 256 < nblocks; b += 512) {
    s0 += partial[(i64)b * ld + q];
    s1 += partial[(i64)(b + 256) * ld + q];
  }
  if (b < nblocks) s0 += partial[(i64)b * ld + q];
  const double s = block_sum(s0 + s1, scratch);
  if (threadIdx.x == 0) out[q] = s;
}
}  // namespace mpccb

namespace mpccb {
// Tk[(a,i)][(b,j)] = T[a][b][i][j], both with j pitch jp (pad column copied as it is: zero): the slab with K = (b,j)
// contiguous per row, for the flat-K Yvo GEMM. grid (v, na): one CTA per (a,b) block, read contiguously, written as
// o runs of jp doubles.
__global__ void __launch_bounds__(256)
k_t2_flatten_k(const double* __restrict__ T, double* __restrict__ Tk, int o, int jp, int v) {
  const int b = blockIdx.x, a = blockIdx.y;
  const double* src = T + ((i64)a * v + b) * o * jp;
  double* dst = Tk + (i64)a * o * v * jp + (i64)b * jp;
  const int n = o * jp;
  for (int t = threadIdx.x; t < n; t += blockDim.x) {
    const int i = t / jp, j = t - i * jp;
    dst[(i64)i * v * jp + j] = src[t];
  }
}
}  // namespace mpccb

namespace mpccb {
// In-place t2 build for the unfused path (reference dfrmpcc_lowlevel.py:137,:139): the GEMM wrote
// e[(a,i),(b,j)] into T[a][b][i][j]; one block per (a,b) turns its o x o block into
//   t2[i][j] = (2 e[i][j] - e[j][i]) / (eo[i] + eo[j] - ev[a] - ev[b])
// Each thread owns a pair {(i,j),(j,i)}, i <= j: both inputs are read before either output is
// written, so no staging copy is needed. Then OmP[b][(a,i)] = sum_j t2[i][j] * Fvo[b][j] (row dots,
// fixed order) from the block's own freshly written (L1/L2-hot) values.
__global__ void __launch_bounds__(256)
k_t2_finalize(double* __restrict__ T, const double* __restrict__ eo, const double* __restrict__ ev,
              const double* __restrict__ Fvo, double* __restrict__ OmP, int o, int jp, int v, int a_begin, int na,
              SymInfo sym, const unsigned char* __restrict__ flags) {
  // Mirror scheme (flags != nullptr; one byte per (a,b) block of the slab: bit 0 = this rank computes the block,
  // bit 1 = it also writes t2[b,a,:,:] = t2[a,b,:,:]^T, bits 4.. = rank that owns b): blocks without bit 0 were
  // not stored by the GEMM and are not touched here — they arrive as somebody's mirror, in this rank's buffer or,
  // over NVLink, from the rank that computed the transposed block. Of a diagonal block only the j >= i half of e
  // is read (e is symmetric).
  const int b = blockIdx.x, al = blockIdx.y, a = a_begin + al;
  const unsigned char f = flags ? flags[(i64)al * v + b] : (unsigned char)1;
  if (!(f & 1)) return;
  double* blk = T + ((i64)al * v + b) * o * jp;
  const int s = f >> 4;
  double* mir = (f & 2) ? sym.Tp[s] + ((i64)(b - sym.bounds[s]) * v + a) * o * jp : nullptr;
  const bool diag = flags && b == a;
  const double dab = ev[a] + ev[b];
  const int npair = o * (o + 1) / 2;
  for (int p = threadIdx.x; p < npair; p += blockDim.x) {
    // p -> (j, i) with i <= j: row j of the packed lower triangle
    int j = (int)((sqrt(8.0 * p + 1.0) - 1.0) * 0.5);
    while (j * (j + 1) / 2 > p) --j;
    while ((j + 1) * (j + 2) / 2 <= p) ++j;
    const int i = p - j * (j + 1) / 2;
    const double eij = blk[i * jp + j];
    const double eji = diag ? eij : blk[j * jp + i];
    const double den = eo[i] + eo[j] - dab;
    const double tij = (2.0 * eij - eji) / den, tji = (2.0 * eji - eij) / den;
    blk[i * jp + j] = tij;
    if (i != j) blk[j * jp + i] = tji;
    if (mir) {
      mir[j * jp + i] = tij;
      if (i != j) mir[i * jp + j] = tji;
    }
  }
  __threadfence_block();
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const double* f_b = Fvo + (i64)b * o;
  for (int i = warp; i < o; i += nw) {
    double sum = 0.0;
    for (int j = lane; j < o; j += 32) sum += blk[i * jp + j] * f_b[j];
    sum = warp_sum(sum);
    if (lane == 0) OmP[((i64)b * na + al) * o + i] = sum;
  }
  if (mir) {   // the mirror block's share of sum t2 . Fov: rows (b, j) of whichever slab owns b, from the local block
    const double* f_a = Fvo + (i64)a * o;
    const i64 nov = (i64)o * v;
    for (int j = warp; j < o; j += nw) {
      double sum = 0.0;
      for (int i = lane; i < o; i += 32) sum += blk[i * jp + j] * f_a[i];
      sum = warp_sum(sum);
      if (lane == 0) sym.OmPm[(i64)al * nov + (i64)b * o + j] = sum;
    }
  }
}
}  // namespace mpccb

namespace mpccb {
// full[l][r][c] = packed[l][max(r,c)*(max(r,c)+1)/2 + min(r,c)]: lib.unpack_tril of the Cholesky rows
// with_df.loop() yields (reference dfrmpcc_lowlevel.py:55-57; NPdunpack_tril + NPdsymm_triu,
// pyscf/lib/np_helper/pack_tril.c:24-37,58-69). One block row per (l, r).
__global__ void __launch_bounds__(256)
k_unpack_tril_sym(const double* __restrict__ packed, double* __restrict__ full, int nao, i64 npair) {
  const int r = blockIdx.x, l = blockIdx.y;
  const double* src = packed + (i64)l * npair;
  double* dst = full + ((i64)l * nao + r) * nao;
  const i64 rbase = (i64)r * (r + 1) / 2;
  for (int c = threadIdx.x; c < nao; c += blockDim.x)
    dst[c] = (c <= r) ? src[rbase + c] : src[(i64)c * (c + 1) / 2 + r];
}
}  // namespace mpccb
