// Generic FP64 tensor-core (DMMA) GEMM for the DF-MPCC contractions.
//
//   C[m, n] = sum_k A[m, k] * B[n, k]          (both operands "rows x K")
//
// Every contraction of MPCC_LL.updated_amps (reference
// pyscf/mpcc/dfrmpcc_lowlevel.py:119-147) is an instance of this with a
// different *addressing* of the rows/K of each operand, so operands are
// described by a two-level index map instead of being transposed in memory:
//
//   element(r, k) = base[(r / rdiv) * r_hi + (r % rdiv) * r_lo
//                        + (k / kdiv) * k_hi + (k % kdiv) * k_lo]
//
// Template flag AK/BK says which index is unit-stride in global memory
// (K-major: k_lo == 1, lanes run along k; row-major: r_lo == 1, lanes run
// along rows) — it selects the coalesced load pattern and the conflict-free
// shared-memory layout, it never changes the result.
//
// Pipeline: global -> registers (prefetch of tile t+1 overlaps the DMMA work on
// tile t) -> padded shared memory (double buffered, one __syncthreads per
// K-tile) -> DMMA.8x8x4 fragments. This kernel needs no alignment of sizes or
// pointers beyond 8 bytes, so it also serves odd nocc/nvir; the TMA-fed
// kernel in gemm_tma.cuh takes over the two O(naux*nov^2) products when the
// shapes allow it.
#pragma once
#include "common.cuh"

namespace mpccb {

struct Operand {
  const double* base;
  i64 r_hi, r_lo, k_hi, k_lo;
  int rdiv, kdiv;
  int rows;
};

struct Output {
  double* base;        // base[off] = acc
  double* base2;       // optional: base2[off] = acc + aux[off]
  const double* aux;
  i64 m_hi, m_lo, n_hi, n_lo;
  int mdiv, ndiv;
  i64 split_stride;    // offset between split-K slices (0 if none)
  // optional store mask (mirror scheme of the unfused eris GEMM): element (m, n) belongs to block
  // (m / mdiv, n / ndiv); it is stored only if keep[(m / mdiv) * keep_ld + n / ndiv] & 1, and a tile without any
  // kept block returns before its K loop. nullptr: store everything.
  const unsigned char* keep;
  int keep_ld;
  // base[off] = alpha * acc + beta * base[off]  (beta == 0: the old value is not read). The tensor engine's
  // "out = alpha * einsum + beta * out" lands in the permuted result without a separate pass.
  double alpha, beta;
};

struct GemmDims {
  int M, N, K;
  int tm_eff, tn_eff;  // rows/cols a tile advances by (<= TM/TN); == TM/TN for plain GEMMs
  int tiles_m, tiles_n;
  int kchunk;          // K range per blockIdx.z (multiple of 16)
};

constexpr int TK = 16;
constexpr int KPITCH = TK + 4;   // K-major smem pitch: (4*g + tig) hits 16 distinct 8-byte banks

template <int WM_, int WN_, int MT_, int NT_, bool AK_, bool BK_>
struct GemmCfg {
  static constexpr int WM = WM_, WN = WN_, MT = MT_, NT = NT_;
  static constexpr bool AK = AK_, BK = BK_;
  static constexpr int TM = WM * MT * 8, TN = WN * NT * 8;
  static constexpr int NTH = WM * WN * 32;
  static constexpr int ASZ = AK ? TM * KPITCH : TK * (TM + 4);
  static constexpr int BSZ = BK ? TN * KPITCH : TK * (TN + 4);
  static constexpr int APT = TM * TK / NTH;   // prefetch elements per thread
  static constexpr int BPT = TN * TK / NTH;
  static constexpr size_t MAIN_SMEM = sizeof(double) * 2 * (ASZ + BSZ) + sizeof(i64) * (TM + TN + 4 * TK);
  static_assert((TM * TK) % NTH == 0 && (TN * TK) % NTH == 0, "tile/threads mismatch");
  static_assert(NTH % TK == 0, "a thread keeps one k lane for K-major operands");
  // 32 x 32 warp tiles fit in 128 registers: two CTAs (16 warps) per SM hide the global -> register -> shared
  // staging latency that a single 8-warp CTA exposes
  static constexpr int MIN_CTAS = (MT * NT <= 16 && NTH <= 256) ? 2 : 1;
};

__device__ __forceinline__ i64 row_offset(const Operand& op, int r) {
  if (r >= op.rows) return -1;
  return (i64)(r / op.rdiv) * op.r_hi + (i64)(r % op.rdiv) * op.r_lo;
}
__device__ __forceinline__ i64 k_offset(const Operand& op, int k, int kend) {
  if (k >= kend) return -1;
  return (i64)(k / op.kdiv) * op.k_hi + (i64)(k % op.kdiv) * op.k_lo;
}

// tile id -> (tm, tn): n-tiles grouped 8 wide so that the CTAs resident at one
// time share operand tiles through L2.
__device__ __forceinline__ void tile_coords(int id, int tiles_m, int tiles_n, int& tm, int& tn, int gw = 8) {
  const int per_group = gw * tiles_m;
  const int group = id / per_group;
  const int first = group * gw;
  const int width = min(tiles_n - first, gw);
  const int rem = id - group * per_group;
  tn = first + rem % width;
  tm = rem / width;
}

// Accumulate the (m0.., n0..) tile over k in [kbeg, kend). smem must hold Cfg::MAIN_SMEM bytes.
template <class Cfg>
__device__ __forceinline__ void gemm_mainloop(const Operand& A, const Operand& B, int m0, int n0,
                                              int kbeg, int kend,
                                              double (&acc)[Cfg::MT][Cfg::NT][2], unsigned char* smem_raw) {
  constexpr int TM = Cfg::TM, TN = Cfg::TN, NTH = Cfg::NTH;
  constexpr int MT = Cfg::MT, NT = Cfg::NT;
  double* As = reinterpret_cast<double*>(smem_raw);
  double* Bs = As + 2 * Cfg::ASZ;
  i64* rowA = reinterpret_cast<i64*>(Bs + 2 * Cfg::BSZ);
  i64* rowB = rowA + TM;
  i64* kofA = rowB + TN;        // [2][TK]
  i64* kofB = kofA + 2 * TK;    // [2][TK]

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int g = lane >> 2, tig = lane & 3;
  const int wm = warp / Cfg::WN, wn = warp % Cfg::WN;

#pragma unroll
  for (int i = 0; i < MT; ++i)
#pragma unroll
    for (int j = 0; j < NT; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }

  const int nkt = (kend - kbeg + TK - 1) / TK;
  for (int r = t; r < TM; r += NTH) rowA[r] = row_offset(A, m0 + r);
  for (int r = t; r < TN; r += NTH) rowB[r] = row_offset(B, n0 + r);
  if (t < 2 * TK) {
    const int kt = t / TK, kk = t % TK;
    kofA[kt * TK + kk] = k_offset(A, kbeg + kt * TK + kk, kend);
    kofB[kt * TK + kk] = k_offset(B, kbeg + kt * TK + kk, kend);
  }
  __syncthreads();

  double ra[Cfg::APT], rb[Cfg::BPT];

  auto fetch = [&](int kt) {
    const i64* ka = kofA + (kt & 1) * TK;
    const i64* kb = kofB + (kt & 1) * TK;
#pragma unroll
    for (int p = 0; p < Cfg::APT; ++p) {
      const int e = t + p * NTH;
      const int k = Cfg::AK ? (e % TK) : (e / TM);
      const int r = Cfg::AK ? (e / TK) : (e % TM);
      const i64 ro = rowA[r], ko = ka[k];
      ra[p] = (ro >= 0 && ko >= 0) ? __ldg(A.base + ro + ko) : 0.0;
    }
#pragma unroll
    for (int p = 0; p < Cfg::BPT; ++p) {
      const int e = t + p * NTH;
      const int k = Cfg::BK ? (e % TK) : (e / TN);
      const int r = Cfg::BK ? (e / TK) : (e % TN);
      const i64 ro = rowB[r], ko = kb[k];
      rb[p] = (ro >= 0 && ko >= 0) ? __ldg(B.base + ro + ko) : 0.0;
    }
  };
  auto stash = [&](int buf) {
    double* as = As + buf * Cfg::ASZ;
    double* bs = Bs + buf * Cfg::BSZ;
#pragma unroll
    for (int p = 0; p < Cfg::APT; ++p) {
      const int e = t + p * NTH;
      if (Cfg::AK) as[(e / TK) * KPITCH + (e % TK)] = ra[p];
      else         as[(e / TM) * (TM + 4) + (e % TM)] = ra[p];
    }
#pragma unroll
    for (int p = 0; p < Cfg::BPT; ++p) {
      const int e = t + p * NTH;
      if (Cfg::BK) bs[(e / TK) * KPITCH + (e % TK)] = rb[p];
      else         bs[(e / TN) * (TN + 4) + (e % TN)] = rb[p];
    }
  };

  fetch(0);
  stash(0);
  __syncthreads();

  for (int kt = 0; kt < nkt; ++kt) {
    const bool more = (kt + 1 < nkt);
    if (more) fetch(kt + 1);
    const double* as = As + (kt & 1) * Cfg::ASZ;
    const double* bs = Bs + (kt & 1) * Cfg::BSZ;
#pragma unroll
    for (int ks = 0; ks < TK / 4; ++ks) {
      double a[MT], b[NT];
#pragma unroll
      for (int i = 0; i < MT; ++i) {
        const int r = wm * MT * 8 + i * 8 + g;
        a[i] = Cfg::AK ? as[r * KPITCH + ks * 4 + tig] : as[(ks * 4 + tig) * (TM + 4) + r];
      }
#pragma unroll
      for (int j = 0; j < NT; ++j) {
        const int r = wn * NT * 8 + j * 8 + g;
        b[j] = Cfg::BK ? bs[r * KPITCH + ks * 4 + tig] : bs[(ks * 4 + tig) * (TN + 4) + r];
      }
#pragma unroll
      for (int i = 0; i < MT; ++i)
#pragma unroll
        for (int j = 0; j < NT; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    if (more) stash((kt + 1) & 1);
    if (kt + 2 < nkt && t < TK) {
      kofA[(kt & 1) * TK + t] = k_offset(A, kbeg + (kt + 2) * TK + t, kend);
      kofB[(kt & 1) * TK + t] = k_offset(B, kbeg + (kt + 2) * TK + t, kend);
    }
    __syncthreads();
  }
}

// ---------------------------------------------------------------------------
// plain GEMM with strided / transposing store epilogue (optionally split-K:
// blockIdx.z selects the K chunk and the output slice).
// ---------------------------------------------------------------------------
template <class Cfg>
__global__ void __launch_bounds__(Cfg::NTH, Cfg::MIN_CTAS)
gemm_store_kernel(Operand A, Operand B, Output C, GemmDims d) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int tm, tn;
  tile_coords(blockIdx.x, d.tiles_m, d.tiles_n, tm, tn);
  const int m0 = tm * d.tm_eff, n0 = tn * d.tn_eff;
  if (C.keep) {
    bool any = false;
    const int r1 = (min(m0 + d.tm_eff, d.M) - 1) / C.mdiv, c1 = (min(n0 + d.tn_eff, d.N) - 1) / C.ndiv;
    for (int r = m0 / C.mdiv; r <= r1 && !any; ++r)
      for (int cidx = n0 / C.ndiv; cidx <= c1; ++cidx)
        if (C.keep[(i64)r * C.keep_ld + cidx] & 1) { any = true; break; }
    if (!any) return;
  }
  const int kbeg = blockIdx.z * d.kchunk;
  const int kend = min(d.K, kbeg + d.kchunk);
  double acc[Cfg::MT][Cfg::NT][2];
  gemm_mainloop<Cfg>(A, B, m0, n0, kbeg, kend, acc, smem_raw);

  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, tig = lane & 3;
  const int wm = warp / Cfg::WN, wn = warp % Cfg::WN;
  const i64 zoff = (i64)blockIdx.z * C.split_stride;
#pragma unroll
  for (int i = 0; i < Cfg::MT; ++i) {
    const int ml = wm * Cfg::MT * 8 + i * 8 + g;
    const int m = m0 + ml;
    if (ml >= d.tm_eff || m >= d.M) continue;
    const i64 moff = (i64)(m / C.mdiv) * C.m_hi + (i64)(m % C.mdiv) * C.m_lo + zoff;
#pragma unroll
    for (int j = 0; j < Cfg::NT; ++j) {
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        const int nl = wn * Cfg::NT * 8 + j * 8 + 2 * tig + c;
        const int n = n0 + nl;
        if (nl >= d.tn_eff || n >= d.N) continue;
        if (C.keep && !(C.keep[(i64)(m / C.mdiv) * C.keep_ld + n / C.ndiv] & 1)) continue;
        const i64 off = moff + (i64)(n / C.ndiv) * C.n_hi + (i64)(n % C.ndiv) * C.n_lo;
        const double v = acc[i][j][c];
        C.base[off] = (C.alpha == 1.0 && C.beta == 0.0) ? v : fma(C.alpha, v, C.beta != 0.0 ? C.beta * C.base[off] : 0.0);
        if (C.base2) C.base2[off] = v + C.aux[off];
      }
    }
  }
}

// ---------------------------------------------------------------------------
// Dressed-integral GEMM with the fused T2 epilogue
//   (reference dfrmpcc_lowlevel.py:135-137,139):
//   e[(a,i),(b,j)] = sum_L J[(a,i),L] J[(b,j),L]
//   t2[a,b,i,j]    = (2 e[(a,i),(b,j)] - e[(a,j),(b,i)]) / (-(ev[a]-eo[i]) - (ev[b]-eo[j]))
//   OmP[tn][(a,i)] = sum_{b in tile, j} t2[a,b,i,j] * Fvo[b,j]
// A tile holds n_a whole `a` values x n_b whole `b` values (all i, all j), so
// the exchange partner e[(a,j),(b,i)] sits in the same CTA: the accumulators
// are staged once through shared memory and t2 is written exactly once, in
// [a,b,i,j] order (contiguous o x o blocks: coalesced, DIIS-pack order).
// ---------------------------------------------------------------------------
struct T2Params {
  double* T;            // [(a - a_begin), b, i, j]
  const double* eo;     // [o]
  const double* ev;     // [v]
  const double* Fvo;    // [b, j]
  double* OmP;          // [tiles_n][(a - a_begin) * o + i]
  int o, v, a_begin, a_end, n_a, n_b;
  int jp;               // pitch of j in T (nocc rounded up to even)
};

// Symmetry of the dressed integrals: e[(a,i),(b,j)] = e[(b,j),(a,i)] (a Gram matrix), hence
// t2[a,b,i,j] = t2[b,a,j,i]. Only one block of every pair {(a,b),(b,a)} is computed; its CTA also writes the
// mirror block t2[b,a,:,:] = t2[a,b,:,:]^T — into the local t2 buffer when b lies in the own slab, or straight into
// the owning rank's buffer over NVLink when the t2 buffers are peer-mapped (halves the SYRK-shaped GEMM at every
// rank count). Which block of a pair is computed (q = rank owning a, s = rank owning b, P ranks):
//   s == q          : b >= a
//   no peer mapping : every block with s != q is computed by q (nothing to mirror into)
//   d = (s-q) mod P : 2d < P -> q computes;  2d > P -> s computes;  2d == P (antipodal slabs): the rank with the
//                     smaller index computes the lower half of the other slab's b range, the larger one mirrors the rest
struct SymInfo {
  int enabled, nranks, rank, peers;
  int bounds[17];            // slab begin of every rank; bounds[nranks] = nvir
  double* Tp[16];            // t2 buffer of every rank ([a - bounds[r], b, i, j]); Tp[rank] is the local one
  const int2* tiles;         // launch list of (tm, tn) pairs, or nullptr for the full grid
  double* OmPm;              // [tiles_m][nov]: partial sums of t2 . Fov (:139) contributed through mirror blocks
};

__host__ __device__ __forceinline__ int sym_owner(const SymInfo& y, int b) {
  int s = 0;
  while (s + 1 < y.nranks && b >= y.bounds[s + 1]) ++s;
  return s;
}
// does the rank that owns `a` compute block (a,b)?  mirror: must it also write block (b,a)?  s: owner of b
__host__ __device__ __forceinline__ bool sym_compute(const SymInfo& y, int a, int b, bool& mirror, int& s) {
  mirror = false;
  s = y.rank;
  if (!y.enabled) return true;
  const int q = y.rank;
  s = sym_owner(y, b);
  if (s == q) { mirror = b > a; return b >= a; }
  if (!y.peers) return true;
  const int P = y.nranks;
  const int d = (s - q + P) % P;
  bool mine;
  if (2 * d < P) mine = true;
  else if (2 * d > P) mine = false;
  else if (q < s) mine = b < (y.bounds[s] + y.bounds[s + 1]) / 2;
  else mine = a >= (y.bounds[q] + y.bounds[q + 1]) / 2;
  mirror = mine;
  return mine;
}

template <class Cfg>
struct T2Smem {
  static constexpr int PITCH = Cfg::TN + 1;   // odd pitch: column reads of the partner are conflict-free
  static constexpr size_t STAGE = sizeof(double) * Cfg::TM * PITCH;
  static constexpr size_t BYTES = STAGE > Cfg::MAIN_SMEM ? STAGE : Cfg::MAIN_SMEM;
};

template <class Cfg>
__device__ __forceinline__ void t2_epilogue(double (&acc)[Cfg::MT][Cfg::NT][2], unsigned char* smem_raw,
                                            const T2Params& p, const GemmDims& d, int tm, int tn) {
  constexpr int PITCH = T2Smem<Cfg>::PITCH;
  double* S = reinterpret_cast<double*>(smem_raw);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, tig = lane & 3;
  const int wm = warp / Cfg::WN, wn = warp % Cfg::WN;
  // caller guarantees all warps are past the main loop's last __syncthreads
#pragma unroll
  for (int i = 0; i < Cfg::MT; ++i)
#pragma unroll
    for (int j = 0; j < Cfg::NT; ++j) {
      const int r = wm * Cfg::MT * 8 + i * 8 + g;
      const int c = wn * Cfg::NT * 8 + j * 8 + 2 * tig;
      S[r * PITCH + c] = acc[i][j][0];
      S[r * PITCH + c + 1] = acc[i][j][1];
    }
  __syncthreads();
  const int o = p.o;
  const int nwarps = Cfg::NTH / 32;
  const int mslab = (p.a_end - p.a_begin) * o;
  for (int r = warp; r < d.tm_eff; r += nwarps) {
    const int al = r / o, i = r - al * o;
    const int a = p.a_begin + tm * p.n_a + al;
    if (a >= p.a_end) break;
    const double dai = p.ev[a] - p.eo[i];
    double fsum = 0.0;
    for (int bl = 0; bl < p.n_b; ++bl) {
      const int b = tn * p.n_b + bl;
      if (b >= p.v) break;
      const double evb = p.ev[b];
      double* trow = p.T + (((i64)(a - p.a_begin) * p.v + b) * o + i) * p.jp;
      const double* frow = p.Fvo + (i64)b * o;
      for (int j = lane; j < o; j += 32) {
        const double e1 = S[r * PITCH + bl * o + j];
        const double e2 = S[(al * o + j) * PITCH + bl * o + i];
        const double den = -dai - (evb - p.eo[j]);
        const double tv = (2.0 * e1 - e2) / den;
        trow[j] = tv;
        fsum += tv * frow[j];
      }
    }
    fsum = warp_sum(fsum);
    if (lane == 0) p.OmP[(i64)tn * mslab + (i64)(a - p.a_begin) * o + i] = fsum;
  }
}

template <class Cfg>
__global__ void __launch_bounds__(Cfg::NTH, 1)
gemm_t2_kernel(Operand A, Operand B, T2Params p, GemmDims d) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  int tm, tn;
  tile_coords(blockIdx.x, d.tiles_m, d.tiles_n, tm, tn);
  double acc[Cfg::MT][Cfg::NT][2];
  gemm_mainloop<Cfg>(A, B, tm * d.tm_eff, tn * d.tn_eff, 0, d.K, acc, smem_raw);
  t2_epilogue<Cfg>(acc, smem_raw, p, d, tm, tn);
}

}  // namespace mpccb
