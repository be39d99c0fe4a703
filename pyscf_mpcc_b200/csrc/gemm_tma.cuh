// TMA-fed FP64 DMMA GEMM for the two O(naux * nov^2) products of the DF-MPCC
// iteration (reference pyscf/mpcc/dfrmpcc_lowlevel.py:135 "eris" and :138 "Yvo").
//
//   C[m, n] = sum_k A[m, k] * B[n, k],  A and B K-major in global memory.
//
// Blackwell has no tcgen05 kind for FP64, so the math is DMMA.8x8x4 from registers;
// what is sm_100a-native here is the data path: cp.async.bulk.tensor (TMA) loads of 128-row x 16-k
// operand boxes into a 4-stage shared-memory ring (SWIZZLE_128B), completion signalled on mbarriers;
// every warp (4 per SM sub-partition, 32x32 accumulators each) is a consumer that never touches global
// memory or block-wide barriers inside the K loop, and lane 0 of warp 0 doubles as the TMA producer
// (a separate producer warp would be the 5th warp of one sub-partition and cap the kernel at 96 registers).
// Epilogues: plain / split-K store (Yvo, generic products) and the fused eris -> t2 epilogue with the
// mirror scheme (SymInfo, gemm_dmma.cuh).
//
// Bank conflicts: a fragment load reads element (row, k = 4*ks + tig) for the 8 rows `g` of a group.
// With the 128-byte swizzle the 16-byte chunk index is (k/2) ^ (row % 8); mapping fragment row g to
// tile row rho(g) = {0,2,4,6,1,3,5,7}[g] makes the 16 lanes of each half-warp hit 16 distinct 8-byte
// banks. rho is a permutation inside every 8-row group, applied to both operands; the epilogues
// simply use the permuted coordinates.
#pragma once
#include <cuda.h>

#include "gemm_dmma.cuh"

namespace mpccb {

constexpr int TMA_TM = 128, TMA_TK = 16;
constexpr int TMA_STAGES = 4;
constexpr int TMA_BAR_BYTES = 128;

// WN = consumer warps along N (32 columns each): 4 -> 128 x 128 tile, 16 warps, one CTA per SM;
// 2 -> 128 x 64 tile, 8 warps, two CTAs per SM (one CTA's epilogue overlaps the other's K loop).
template <int WN>
struct TmaCfg {
  static constexpr int TN = 32 * WN;
  static constexpr int WARPS = 4 * WN;
  static constexpr int THREADS = WARPS * 32;
  static constexpr int MIN_CTAS = (WN == 2) ? 2 : 1;
  static constexpr int STAGE_BYTES = (TMA_TM + TN) * TMA_TK * 8;
  static constexpr int RING_BYTES = TMA_STAGES * STAGE_BYTES;
  static constexpr int STAGING_BYTES = 8 * TMA_TM * (TN + 1);      // fused-t2 epilogue staging (aliases the ring)
  static constexpr int FLAG_BYTES = TMA_TM * TN;                   // one byte per (a,b) block of the tile (nocc = 1)
  static constexpr int AREA_EPI1 = STAGING_BYTES + FLAG_BYTES > RING_BYTES ? STAGING_BYTES + FLAG_BYTES : RING_BYTES;
  // plain / index-mapped store epilogues stage the tile through shared memory too (coalesced global stores)
  static constexpr int AREA_STORE = STAGING_BYTES > RING_BYTES ? STAGING_BYTES : RING_BYTES;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1,
                                            int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
      ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}

// fragment row g (0..7) -> tile row inside its 8-row group
__device__ __forceinline__ int rho(int g) { return ((g & 3) << 1) | (g >> 2); }

// 8-byte element (row, k) of a [rows][16] swizzle-128B tile, as a double index
__device__ __forceinline__ int swz(int row, int k) { return row * 16 + ((((k >> 1) ^ (row & 7)) << 1) | (k & 1)); }

// x / y without the compiler's special-case branch: MUFU.RCP64H seed (about 20 bits), Newton steps, one
// residual correction of the quotient. Within 1 ulp of the IEEE quotient for normal y (orbital-energy
// denominators); branch-free, so that many independent divisions interleave in one thread.
__device__ __forceinline__ double fast_div(double x, double y) {
  double r;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(y));
  double e = fma(-y, r, 1.0);
  r = fma(r, e, r);
  e = fma(-y, r, 1.0);
  r = fma(r, e, r);
  e = fma(-y, r, 1.0);
  r = fma(r, e, r);
  double q = x * r;
  const double rem = fma(-q, y, x);
  return fma(rem, r, q);
}

// N independent quotients with every Newton step issued for all of them before the next one
template <int N>
__device__ __forceinline__ void fast_div_batch(const double* x, const double* y, double* q) {
  double r[N], e[N];
#pragma unroll
  for (int u = 0; u < N; ++u) asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r[u]) : "d"(y[u]));
#pragma unroll
  for (int it = 0; it < 2; ++it) {
#pragma unroll
    for (int u = 0; u < N; ++u) e[u] = fma(-y[u], r[u], 1.0);
#pragma unroll
    for (int u = 0; u < N; ++u) r[u] = fma(r[u], e[u], r[u]);
  }
#pragma unroll
  for (int u = 0; u < N; ++u) q[u] = x[u] * r[u];
#pragma unroll
  for (int u = 0; u < N; ++u) e[u] = fma(-q[u], y[u], x[u]);
#pragma unroll
  for (int u = 0; u < N; ++u) q[u] = fma(e[u], r[u], q[u]);
}

struct TmaGemmParams {
  int M, N;              // valid rows of A / B for this launch
  int tm_eff, tn_eff;    // rows a tile advances by
  int tiles_m, tiles_n;
  int nk;                // number of 16-wide K steps
  // A-operand addressing: mode 0 = 2-D map (k, row); mode 1 = 4-D map over T[a][b][i][j]:
  //   k step kt -> (b = kt / jsteps, j0 = 16 * (kt % jsteps)), rows = (a, i)
  // mode 2 = 3-D map (k, r, q) over a matrix whose rows m = q * rows_per_q + r are NOT equally spaced (a slab of the
  //   middle index of a 3-index tensor, e.g. Lvv[L][a in slab][b]): rows_per_q = o (reused), tile row m0 -> box origin
  //   (r = m0 % rows_per_q, q = m0 / rows_per_q); the box is (16, min(rows_per_q, 128), 128 / that)
  int a_mode, jsteps, o, n_a;
  int i0;                // a_mode 1: first occupied index i of the rows this launch covers (box = 16 j x w i x 1 b x n_a a)
  uint32_t a_bytes, b_bytes;   // bytes one stage receives for A / B (box sizes)
  // plain store epilogue
  double* C;
  i64 ldc;
  // split-K (EPI 0 only): blockIdx.x = split * tiles + tile; split s covers K steps [s*nk_split, ...) and
  // stores to C + s*split_stride (partials reduced afterwards in fixed order). ksplit == 1: none.
  int ksplit, nk_split, tiles_total;
  i64 split_stride;
  int gw;                // raster: column tiles per group of tile_coords (0 = its default of 8)
  // EPI 2: store through a two-level index map with an optional block mask (gemm_dmma.cuh, Output): the unfused
  // eris GEMM scatters e[(a,i),(b,j)] into T[a][b][i][j] and skips the blocks that arrive as somebody's mirror;
  // the tensor engine's contractions store straight into the permuted result (out.alpha / out.beta: axpby)
  Output out;
};

// EPI: 0 = plain row-major store (Yvo), 1 = fused eris -> t2 epilogue, 2 = index-mapped, masked store (p.out).
// Every warp is a consumer; lane 0 of warp 0 doubles as the TMA producer: at the top of K step kt it
// refills the stage that all warps released at the end of step kt-1 with the operands of step
// kt + STAGES-1. No 17th warp, so a sub-partition holds 4 (or 2 x 2) warps and ptxas gets 128 registers.
template <int EPI, int WN>
__global__ void __launch_bounds__(TmaCfg<WN>::THREADS, TmaCfg<WN>::MIN_CTAS)
gemm_tma_kernel(const __grid_constant__ CUtensorMap mapA, const __grid_constant__ CUtensorMap mapB,
                TmaGemmParams p, T2Params t2p, SymInfo sym) {
  typedef TmaCfg<WN> Cfg;
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  double* ring = reinterpret_cast<double*>(smem_raw);
  constexpr size_t STAGE_AREA = (EPI == 1) ? (size_t)Cfg::AREA_EPI1 : (size_t)Cfg::AREA_STORE;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + STAGE_AREA);
  uint64_t* empty = full + TMA_STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int tm, tn;
  const int ntiles = p.tiles_m * p.tiles_n;
  const int split = (EPI == 0 && p.ksplit > 1) ? blockIdx.x / ntiles : 0;
  if (EPI == 1 && sym.tiles != nullptr) {
    const int2 t = sym.tiles[blockIdx.x];
    tm = t.x; tn = t.y;
  } else {
    tile_coords(blockIdx.x - split * ntiles, p.tiles_m, p.tiles_n, tm, tn, p.gw > 0 ? p.gw : 8);
  }
  const int m0 = tm * p.tm_eff, n0 = tn * p.tn_eff;
  if (EPI == 2 && p.out.keep) {           // no kept block in this tile: nothing to compute (uniform over the CTA)
    bool any = false;
    const int r1 = (min(m0 + p.tm_eff, p.M) - 1) / p.out.mdiv, c1 = (min(n0 + p.tn_eff, p.N) - 1) / p.out.ndiv;
    for (int r = m0 / p.out.mdiv; r <= r1 && !any; ++r)
      for (int cidx = n0 / p.out.ndiv; cidx <= c1; ++cidx)
        if (p.out.keep[(i64)r * p.out.keep_ld + cidx] & 1) { any = true; break; }
    if (!any) return;
  }
  const bool producer = (threadIdx.x == 0);
  const int k_first = split * p.nk_split;                        // first K step of this CTA
  const int nk = (EPI == 0 && p.ksplit > 1) ? min(p.nk_split, p.nk - k_first) : p.nk;

  auto issue = [&](int kl) {                                     // kl: K step local to this CTA
    const int s = kl % TMA_STAGES;
    const int kt = k_first + kl;
    double* a_dst = ring + s * (Cfg::STAGE_BYTES / 8);
    double* b_dst = a_dst + TMA_TM * TMA_TK;
    mbar_expect_tx(&full[s], p.a_bytes + p.b_bytes);
    if (p.a_mode == 0) {
      tma_load_2d(a_dst, &mapA, &full[s], kt * TMA_TK, m0);
    } else if (p.a_mode == 2) {
      const int q = m0 / p.o;
      tma_load_3d(a_dst, &mapA, &full[s], kt * TMA_TK, m0 - q * p.o, q);
    } else {
      // K runs over (b, j); j in chunks of 16 — for nocc not a multiple of 16 the last chunk of every b reads past
      // the j extent of the 4-D map and TMA zero-fills it, which also cancels B's columns of the next b
      const int b = kt / p.jsteps, j0 = (kt - b * p.jsteps) * TMA_TK;
      tma_load_4d(a_dst, &mapA, &full[s], j0, p.i0, b, tm * p.n_a);
      tma_load_2d(b_dst, &mapB, &full[s], b * p.o + j0, n0);
      return;
    }
    tma_load_2d(b_dst, &mapB, &full[s], kt * TMA_TK, n0);
  };

  if (producer) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&mapB) : "memory");
    for (int s = 0; s < TMA_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], Cfg::WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    for (int kt = 0; kt < TMA_STAGES - 1 && kt < nk; ++kt) issue(kt);
  }
  __syncthreads();

  double acc[4][4][2];
  const int g = lane >> 2, tig = lane & 3;
  const int wm = warp / WN, wn = warp % WN;     // 4 x WN warps, 32 x 32 each
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { acc[i][j][0] = 0.0; acc[i][j][1] = 0.0; }
  const int rg = rho(g);
  const int arow = (wm * 32 + rg) * 16, brow = (wn * 32 + rg) * 16;
  for (int kt = 0; kt < nk; ++kt) {
    if (producer) {
      const int kp = kt + TMA_STAGES - 1;
      if (kp < nk) {
        if (kt > 0) mbar_wait(&empty[kp % TMA_STAGES], ((kp / TMA_STAGES) - 1) & 1);
        issue(kp);
      }
    }
    __syncwarp();
    const int s = kt % TMA_STAGES;
    mbar_wait(&full[s], (kt / TMA_STAGES) & 1);
    const double* as = ring + s * (Cfg::STAGE_BYTES / 8) + arow;
    const double* bs = ring + s * (Cfg::STAGE_BYTES / 8) + TMA_TM * TMA_TK + brow;
#pragma unroll
    for (int ks = 0; ks < TMA_TK / 4; ++ks) {
      // swizzle-128B: 16-byte chunk (k/2) ^ (row % 8); row % 8 == rg for every fragment row of this lane
      const int kk = ks * 4 + tig;
      const int col = ((((kk >> 1) ^ rg) << 1) | (kk & 1));
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = as[i * 128 + col];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = bs[j * 128 + col];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma884(acc[i][j][0], acc[i][j][1], a[i], b[j]);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[s]);
  }

  if (EPI == 0 || EPI == 2) {
    // Stage the accumulators through shared memory (the ring is drained: every warp waited on every full barrier,
    // the block barrier orders the last reads before the overwrite), then store with lanes along the unit-stride
    // index of the target: 256-byte runs instead of single doubles at rho-permuted columns.
    __syncthreads();
    constexpr int PITCH = Cfg::TN + 1;            // odd: row- and column-wise reads are both conflict-free
    double* S = reinterpret_cast<double*>(smem_raw);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
#pragma unroll
        for (int c = 0; c < 2; ++c)
          S[(wm * 32 + i * 8 + rho(g)) * PITCH + wn * 32 + j * 8 + rho(2 * tig + c)] = acc[i][j][c];
    __syncthreads();
    constexpr int nwarps = Cfg::WARPS;
    const int rows = min(p.tm_eff, p.M - m0), cols = min(p.tn_eff, p.N - n0);
    if (EPI == 0) {
      double* Cb = p.C + split * p.split_stride;
      for (int r = warp; r < rows; r += nwarps) {
        double* crow = Cb + (i64)(m0 + r) * p.ldc + n0;
        for (int cl = lane; cl < cols; cl += 32) crow[cl] = S[r * PITCH + cl];
      }
    } else {
      const Output& C = p.out;
      const bool axpby = (C.alpha != 1.0) || (C.beta != 0.0);
      auto put = [&](i64 off, double val) {
        C.base[off] = axpby ? fma(C.alpha, val, C.beta != 0.0 ? C.beta * C.base[off] : 0.0) : val;
      };
      if (C.n_lo == 1 || C.m_lo != 1) {           // lanes along n
        for (int r = warp; r < rows; r += nwarps) {
          const int m = m0 + r;
          const int mq = m / C.mdiv;
          const i64 moff = (i64)mq * C.m_hi + (i64)(m - mq * C.mdiv) * C.m_lo;
          for (int cl = lane; cl < cols; cl += 32) {
            const int n = n0 + cl;
            const int nq = n / C.ndiv;
            if (C.keep && !(C.keep[(i64)mq * C.keep_ld + nq] & 1)) continue;
            put(moff + (i64)nq * C.n_hi + (i64)(n - nq * C.ndiv) * C.n_lo, S[r * PITCH + cl]);
          }
        }
      } else {                                    // the target is contiguous along m: lanes along m
        for (int cl = warp; cl < cols; cl += nwarps) {
          const int n = n0 + cl;
          const int nq = n / C.ndiv;
          const i64 noff = (i64)nq * C.n_hi + (i64)(n - nq * C.ndiv) * C.n_lo;
          for (int r = lane; r < rows; r += 32) {
            const int m = m0 + r;
            const int mq = m / C.mdiv;
            if (C.keep && !(C.keep[(i64)mq * C.keep_ld + nq] & 1)) continue;
            put(noff + (i64)mq * C.m_hi + (i64)(m - mq * C.mdiv) * C.m_lo, S[r * PITCH + cl]);
          }
        }
      }
    }
  } else {
    // all consumers have drained the ring (every full barrier was waited on) before the staging
    // buffer, which aliases it, is written
    __syncthreads();
    constexpr int PITCH = Cfg::TN + 1;
    double* S = reinterpret_cast<double*>(smem_raw);
    {
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
          for (int c = 0; c < 2; ++c)
            S[(wm * 32 + i * 8 + rho(g)) * PITCH + wn * 32 + j * 8 + rho(2 * tig + c)] = acc[i][j][c];
    }
    // which (a,b) blocks of the tile this CTA produces (bit 0) and mirrors (bit 1): the mirror rule, once per block
    unsigned char* flags = smem_raw + Cfg::STAGING_BYTES;
    for (int q = threadIdx.x; q < t2p.n_a * t2p.n_b; q += Cfg::THREADS) {
      const int al = q / t2p.n_b, bl = q - al * t2p.n_b;
      const int a = t2p.a_begin + tm * t2p.n_a + al, b = tn * t2p.n_b + bl;
      unsigned char f = 0;
      if (a < t2p.a_end && b < t2p.v) {
        bool mirror; int s;
        const bool comp = sym_compute(sym, a, b, mirror, s);
        f = (comp ? 1 : 0) | (mirror ? 2 : 0);
      }
      flags[q] = f;
    }
    __syncthreads();
    const int o = t2p.o;
    constexpr int nwarps = Cfg::WARPS;
    constexpr int RPW = TMA_TM / nwarps;          // tile rows per warp
    constexpr int CPL = Cfg::TN / 32;             // tile columns per lane
    const int mslab = (t2p.a_end - t2p.a_begin) * o;
    const i64 nov = (i64)t2p.v * o;

    // phase 1: t2 = (2 e - e^x) / D for the whole tile, branch-free, RPW x CPL independent divisions per thread.
    // The FP64 pipe is shared with the sibling CTA's DMMAs, so every dependent FP64 instruction waits long;
    // a row-at-a-time loop (one division chain per thread) made this epilogue as long as the K loop.
    // x / o for x < 2^16 without a division (no control flow inside the batches): umulhi(x, floor(2^32/o) + 1)
    const unsigned magic = o > 1 ? 0xFFFFFFFFu / (unsigned)o + 1u : 0u;
    auto div_o = [&](int x) -> int { return o > 1 ? (int)__umulhi((unsigned)x, magic) : x; };
    int cbl[CPL], cj[CPL], cok[CPL];
    double cden[CPL], cfvo[CPL];
    i64 coloff[CPL];
#pragma unroll
    for (int cc = 0; cc < CPL; ++cc) {
      const int c = lane + cc * 32;
      const bool in = c < p.tn_eff;
      const int bl = in ? div_o(c) : 0;
      const int j = in ? c - bl * o : 0;
      const int b = tn * t2p.n_b + bl;
      cok[cc] = in && b < t2p.v;
      cbl[cc] = bl; cj[cc] = j;
      cden[cc] = t2p.ev[cok[cc] ? b : 0] - t2p.eo[j];
      coloff[cc] = (i64)b * o * t2p.jp + j;
      cfvo[cc] = cok[cc] ? t2p.Fvo[(i64)b * o + j] : 0.0;
    }
    double tv[RPW][CPL];
    constexpr int RB = 8 / CPL;                   // rows per batch: 8 divisions in lockstep
#pragma unroll
    for (int r0 = 0; r0 < RPW; r0 += RB) {
      double num[RB][CPL], den[RB][CPL];
      bool mine[RB][CPL];
      i64 rowoff[RB];
      int omp[RB];
#pragma unroll
      for (int u = 0; u < RB; ++u) {
        const int r = warp + (r0 + u) * nwarps;
        const bool in = r < p.tm_eff;
        const int al = in ? div_o(r) : 0;
        const int i = in ? r - al * o : 0;
        const int a = t2p.a_begin + tm * t2p.n_a + al;
        const bool rok = in && a < t2p.a_end;
        const double dai = t2p.ev[rok ? a : t2p.a_begin] - t2p.eo[i];
        rowoff[u] = ((i64)(a - t2p.a_begin) * t2p.v * o + i) * t2p.jp;
        omp[u] = rok ? (a - t2p.a_begin) * o + i : -1;
#pragma unroll
        for (int cc = 0; cc < CPL; ++cc) {
          const double e1 = S[r * PITCH + lane + cc * 32];
          const double e2 = S[(al * o + cj[cc]) * PITCH + cbl[cc] * o + i];
          num[u][cc] = 2.0 * e1 - e2;
          den[u][cc] = -dai - cden[cc];
          mine[u][cc] = rok && cok[cc] && (flags[al * t2p.n_b + cbl[cc]] & 1);
        }
      }
      double q[RB][CPL];
      fast_div_batch<RB * CPL>(&num[0][0], &den[0][0], &q[0][0]);
      // rows (a,i) -> t2[a,b,i,:] straight from the registers (lanes run along j: coalesced) and the row's share
      // of sum t2 . Fov (:139); blocks produced by the CTA that owns the mirror block stay untouched
#pragma unroll
      for (int u = 0; u < RB; ++u) {
        double fsum = 0.0;
#pragma unroll
        for (int cc = 0; cc < CPL; ++cc) {
          const double x = mine[u][cc] ? q[u][cc] : 0.0;
          tv[r0 + u][cc] = x;
          if (mine[u][cc]) t2p.T[rowoff[u] + coloff[cc]] = x;
          fsum = fma(x, cfvo[cc], fsum);
        }
        fsum = warp_sum(fsum);
        if (lane == 0 && omp[u] >= 0) t2p.OmP[(i64)tn * mslab + omp[u]] = fsum;
      }
    }
    // phase 2 (mirror scheme): rows (b,j) of t2[b,a,:,:] = t2[a,b,:,:]^T, lanes along i (coalesced in the target,
    // which is the local buffer or, over NVLink, the buffer of the rank that owns b). The amplitudes go back into
    // the staging tile (in place, once every thread has read its exchange partners) to be read column-wise.
    if (sym.enabled) {
      __syncthreads();
#pragma unroll
      for (int rr = 0; rr < RPW; ++rr)
#pragma unroll
        for (int cc = 0; cc < CPL; ++cc) S[(warp + rr * nwarps) * PITCH + lane + cc * 32] = tv[rr][cc];
      __syncthreads();
      // thread-fixed mapping again: tile columns rr = warp + k * nwarps are the target rows (b,j); lanes (+32 m)
      // run over the tile rows (a,i), which are contiguous along i in the target
      constexpr int MR = Cfg::TN / nwarps;        // target rows per warp
      constexpr int MC = TMA_TM / 32;             // (a,i) positions per lane
      i64 mcol[MC];
      double mfvo[MC];
      int mal[MC];
      bool mok[MC];
#pragma unroll
      for (int m = 0; m < MC; ++m) {
        const int c = lane + 32 * m;
        const bool in = c < p.tm_eff;
        const int al = in ? div_o(c) : 0;
        const int i = in ? c - al * o : 0;
        const int a = t2p.a_begin + tm * t2p.n_a + al;
        mok[m] = in && a < t2p.a_end;
        mal[m] = al;
        mcol[m] = (i64)a * o * t2p.jp + i;
        mfvo[m] = mok[m] ? t2p.Fvo[(i64)a * o + i] : 0.0;
      }
#pragma unroll
      for (int k = 0; k < MR; ++k) {
        const int rr = warp + k * nwarps;
        const bool in = rr < p.tn_eff;
        const int bl = in ? div_o(rr) : 0;
        const int j = in ? rr - bl * o : 0;
        const int b = tn * t2p.n_b + bl;
        const bool rok = in && b < t2p.v;
        const int s = sym_owner(sym, rok ? b : 0);
        double* tgt = sym.Tp[s] + ((i64)(b - sym.bounds[s]) * t2p.v * o + j) * t2p.jp;
        double fsum = 0.0;
        bool any = false;
#pragma unroll
        for (int m = 0; m < MC; ++m) {
          const bool w = rok && mok[m] && (flags[mal[m] * t2p.n_b + bl] & 2);
          const double x = S[(lane + 32 * m) * PITCH + rr];
          if (w) tgt[mcol[m]] = x;
          fsum = fma(w ? x : 0.0, mfvo[m], fsum);
          any |= w;
        }
        any = __any_sync(0xffffffffu, any);
        fsum = warp_sum(fsum);
        if (lane == 0 && any) sym.OmPm[(i64)tm * nov + (i64)b * o + j] = fsum;
      }
    }
  }
}

}  // namespace mpccb
