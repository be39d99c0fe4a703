// C ABI of libmpcc_b200 (see include/mpcc_b200.h). Host orchestration of one
// DF-MPCC residual evaluation = reference MPCC_LL.updated_amps
// (pyscf/mpcc/dfrmpcc_lowlevel.py:103-152) as a fixed sequence of DMMA GEMMs and
// HBM-bound kernels on one stream.
#include "../../include/mpcc_b200.h"
#include "gemm_dmma.cuh"
#include "gemm_tma.cuh"
#include "kernels_elem.cuh"
#include "kernels_tensor.cuh"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdarg>
#include <cstring>
#include <string>
#include <vector>

using namespace mpccb;

namespace {
constexpr int BIG = 1 << 30;
constexpr int NSM = 148;

// tile configurations of the generic DMMA GEMM (warps m x n, fragments m x n, A K-major?, B K-major?)
typedef GemmCfg<2, 4, 8, 4, true, true> CfgBigKK;     // 128 x 128, 8 warps x (64 x 32): fewest LDS per DMMA, for long K
typedef GemmCfg<4, 4, 4, 4, true, true> CfgBig16KK;   // 128 x 128, 16 warps x (32 x 32): short K (latency-bound)
typedef GemmCfg<4, 2, 4, 4, true, true> CfgN64KK;     // 128 x 64
typedef GemmCfg<4, 2, 4, 4, false, false> CfgN64RR;
typedef GemmCfg<4, 2, 4, 4, true, false> CfgN64KR;
typedef GemmCfg<4, 2, 4, 4, false, true> CfgN64RK;
typedef GemmCfg<2, 4, 8, 4, true, false> CfgBigKR;    // 128 x 128, B row-major
typedef GemmCfg<2, 4, 8, 4, false, false> CfgBigRR;   // 128 x 128, both row-major
typedef GemmCfg<2, 4, 4, 4, false, false> CfgM64RR;   // 64 x 128
}  // namespace

constexpr int DIIS_COLS_GRID = 148 * 16;
constexpr int DIIS_ROWS_GRID = 148 * 32;     // chunks of the packed vector (k_diis_push): 5 CTAs resident per SM (48 registers), 6-7 waves

struct mpccb_ctx {
  int o, v, L, device, a_begin, a_end, na, ldk;
  int jp;                          // pitch of the innermost (j) index of T[a][b][i][j] and Lvo[L][b][j]: nocc rounded up
                                   // to even, so that the 4-D / 2-D tensor maps have 16-byte strides for odd nocc too
  i64 nov;
  cudaStream_t stream;
  long long launches;
  int use_tma;
  int mapped_narrow_k;       // mpccb_gemm_nt_mapped uses 128 x 64 tiles (two CTAs per SM) up to this K (0: never)
  int xvo_generic;           // 1: keep the Xvo GEMM (:119) on the generic kernel (A/B measurements)
  char err[512];
  // bound tensors
  const double *Loo, *Lov, *Lvv, *foo, *fvv, *eo, *ev;
  double* T;
  cudaStream_t side;              // second stream: the F2 GEMM runs beside the bandwidth-bound dressing kernel
  cudaEvent_t ev_fork, ev_join;
  double* P2;                     // split-K scratch of the side stream
  i64 P2_count;
  const long long* diis_prefix;   // column-owned DIIS vector: packed elements before local block k (na*v + 1 entries)
  unsigned char* ws;
  size_t ws_bytes;
  // workspace carve-up (doubles)
  double *Lvo, *t1, *t1T, *t1n, *Xov, *Xoo, *Joo, *X, *XT, *JT, *JK, *E1K, *Wov, *w, *F2, *Fvo, *OmP, *Y,
      *P, *OmRep, *OmLoc, *OmTot, *Z, *scal, *red, *YP;
  int yvo_split, yvo_split_force;
  int yvo_flat, yvo_flat_force;  // flat-K form of the Yvo GEMM (below); force: 0 automatic, 1 on, -1 off
  int raster_yvo, raster_eris;   // column tiles per raster group of the Yvo GEMM / the eris tile list (0 = default)
  double* Tk;                    // t2 slab re-laid as [(a,i)][(b,j)] for it
  JkTargets jk_targets;      // n == 0: plain local JK
  JkTargets w_targets;       // n > 0: the slab's rows of w = Jvo.X are broadcast by mpccb_dress (subset JK exchange)
  i64 w_off;                 // offset of w inside the symmetric JK allocation (doubles)
  double* w_ws;              // w in the private workspace (used unless the subset exchange is bound)
  double* Fvo_ws;            // Fvo likewise
  double* Xoo_ws;            // [Xoo | Joo] in the private workspace (used unless the Xoo exchange is bound)
  JkTargets xoo_targets;     // n > 0: Xoo is computed for an L slab per rank and stored into the peers' buffers
  int xoo_l0, xoo_l1;        // this rank's slab of the auxiliary index
  SymInfo sym;               // symmetric eris/t2 scheme (host copy; tiles/OmPm are device pointers)
  int sym_off, n_tiles;
  const unsigned char* blkflags;   // unfused t2 path under the mirror scheme (see build_tile_list)
  int2* tiles_dev;
  i64 tiles_cap;
  i64 P_count, OmP_count;
  int t2_na, t2_nb, t2_tiles_m, t2_tiles_n, t2_wn, t2_unfused, force_unfused;
  // optional per-stage CUDA-event timing (mpccb_set_option("profile", 1))
  int profile;
  std::vector<cudaEvent_t>* prof_ev;   // [stage] -> start/stop pairs
  int prof_used[16];
  // kernels whose dynamic-shared-memory opt-in was already set for THIS context's device. The attribute is per
  // device, so a process-wide flag would leave a second GPU of the same process without it.
  std::vector<const void*>* attr_set;
};

enum Stage { ST_DRESS = 0, ST_DRESS_POST, ST_ERIS_T2, ST_YVO, ST_SINGLES_FROM_Y, ST_FINISH, ST_DIIS_PUSH, ST_DIIS_COMBINE, ST_COUNT };
static const char* kStageNames[ST_COUNT] = {"dress", "dress_post", "eris_t2", "yvo", "singles_from_y", "finish", "diis_push",
                                            "diis_combine"};

namespace {

int fail(mpccb_ctx* c, const char* fmt, ...) {
  if (c) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(c->err, sizeof(c->err), fmt, ap);
    va_end(ap);
  }
  return 1;
}

// cudaFuncAttributeMaxDynamicSharedMemorySize (and optionally the carve-out preference) once per context and kernel
cudaError_t ensure_smem_attr(mpccb_ctx* c, const void* func, size_t smem, bool prefer_smem) {
  for (const void* f : *c->attr_set)
    if (f == func) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  if (prefer_smem) {
    e = cudaFuncSetAttribute(func, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
    if (e != cudaSuccess) return e;
  }
  c->attr_set->push_back(func);
  return cudaSuccess;
}

// every entry point that launches work runs on the context's device, whatever the caller's current device is
cudaError_t enter_device(const mpccb_ctx* c) {
  int cur = -1;
  cudaError_t e = cudaGetDevice(&cur);
  if (e != cudaSuccess) return e;
  return cur == c->device ? cudaSuccess : cudaSetDevice(c->device);
}

struct StageTimer {
  mpccb_ctx* c;
  int stage;
  bool on;
  StageTimer(mpccb_ctx* ctx, int st) : c(ctx), stage(st), on(false) {
    if (!c->profile) return;
    std::vector<cudaEvent_t>& v = c->prof_ev[stage];
    const int need = c->prof_used[stage] + 2;
    while ((int)v.size() < need) {
      cudaEvent_t e;
      if (cudaEventCreate(&e) != cudaSuccess) return;
      v.push_back(e);
    }
    cudaEventRecord(v[c->prof_used[stage]], c->stream);
    on = true;
  }
  ~StageTimer() {
    if (!on) return;
    cudaEventRecord(c->prof_ev[stage][c->prof_used[stage] + 1], c->stream);
    c->prof_used[stage] += 2;
  }
};

#define CUDA_OK(c, call)                                                                   \
  do {                                                                                     \
    cudaError_t e_ = (call);                                                               \
    if (e_ != cudaSuccess)                                                                 \
      return fail(c, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

#define LAUNCH_OK(c, what)                                                                 \
  do {                                                                                     \
    (c)->launches++;                                                                       \
    cudaError_t e_ = cudaGetLastError();                                                   \
    if (e_ != cudaSuccess) return fail(c, "launch of %s failed: %s", what, cudaGetErrorString(e_)); \
  } while (0)

inline i64 align_up(i64 x, i64 a) { return (x + a - 1) / a * a; }
inline int cdiv(i64 a, i64 b) { return (int)((a + b - 1) / b); }

Operand flat_operand(const double* base, int rows, i64 row_stride, i64 k_stride) {
  Operand op;
  op.base = base; op.rows = rows;
  op.r_hi = 0; op.r_lo = row_stride; op.rdiv = BIG;
  op.k_hi = 0; op.k_lo = k_stride; op.kdiv = BIG;
  return op;
}

Output flat_output(double* base, i64 m_stride, i64 n_stride) {
  Output c;
  c.base = base; c.base2 = nullptr; c.aux = nullptr;
  c.m_hi = 0; c.m_lo = m_stride; c.mdiv = BIG;
  c.n_hi = 0; c.n_lo = n_stride; c.ndiv = BIG;
  c.split_stride = 0;
  c.keep = nullptr; c.keep_ld = 0;
  c.alpha = 1.0; c.beta = 0.0;
  return c;
}

// number of K splits so that a small-output GEMM still fills the machine
int choose_splits(int tiles, int K) {
  if (tiles >= NSM) return 1;
  int s = cdiv(2 * NSM, tiles);
  const int max_s = K / 64 > 0 ? K / 64 : 1;
  if (s > max_s) s = max_s;
  return s < 1 ? 1 : s;
}

template <class Cfg>
int launch_gemm(mpccb_ctx* c, const char* what, const Operand& A, const Operand& B, Output C,
                int M, int N, int K, int splits, double* partial) {
  GemmDims d;
  d.M = M; d.N = N; d.K = K;
  d.tm_eff = Cfg::TM; d.tn_eff = Cfg::TN;
  d.tiles_m = cdiv(M, Cfg::TM); d.tiles_n = cdiv(N, Cfg::TN);
  d.kchunk = (int)align_up(cdiv(K, splits), TK);
  splits = cdiv(K, d.kchunk);
  if (splits > 1) {
    // partial[s][m][n] row-major, reduced by the caller
    C = flat_output(partial, N, 1);
    C.split_stride = (i64)M * N;
  }
  CUDA_OK(c, ensure_smem_attr(c, reinterpret_cast<const void*>(&gemm_store_kernel<Cfg>), Cfg::MAIN_SMEM, false));
  dim3 grid(d.tiles_m * d.tiles_n, 1, splits);
  gemm_store_kernel<Cfg><<<grid, Cfg::NTH, Cfg::MAIN_SMEM, c->stream>>>(A, B, C, d);
  LAUNCH_OK(c, what);
  return 0;
}

int reduce_splits(mpccb_ctx* c, double* dst, const double* P, int nsplit, i64 n, i64 stride,
                  double alpha, double beta) {
  k_reduce_splits<<<cdiv(n, 256), 256, 0, c->stream>>>(dst, P, nsplit, n, stride, alpha, beta);
  LAUNCH_OK(c, "k_reduce_splits");
  return 0;
}

// split-K GEMM with tiny M x N: partials then a fixed-order reduction into dst (row-major M x N)
template <class Cfg>
int splitk_gemm(mpccb_ctx* c, const char* what, const Operand& A, const Operand& B, double* dst,
                int M, int N, int K, double alpha, double beta) {
  const int tiles = cdiv(M, Cfg::TM) * cdiv(N, Cfg::TN);
  int splits = choose_splits(tiles, K);
  const int kchunk = (int)align_up(cdiv(K, splits), TK);
  splits = cdiv(K, kchunk);
  if ((i64)splits * M * N > c->P_count) return fail(c, "%s: split-K scratch too small", what);
  if (splits == 1) {
    // still go through the partial buffer so alpha/beta apply uniformly
    Output C = flat_output(c->P, N, 1);
    if (launch_gemm<Cfg>(c, what, A, B, C, M, N, K, 1, c->P)) return 1;
  } else {
    Output C = flat_output(c->P, N, 1);
    if (launch_gemm<Cfg>(c, what, A, B, C, M, N, K, splits, c->P)) return 1;
  }
  return reduce_splits(c, dst, c->P, splits, (i64)M * N, (i64)M * N, alpha, beta);
}

i64 splitk_scratch(int M, int N, int K, int TM, int TN) {
  const int tiles = cdiv(M, TM) * cdiv(N, TN);
  const int splits = choose_splits(tiles, K);
  return (i64)(splits + 1) * M * N;
}

}  // namespace

// ===========================================================================
// lifetime / binding
// ===========================================================================
namespace {

struct Carve {
  unsigned char* base;
  i64 off;
  double* take(i64 count) {
    double* p = base ? reinterpret_cast<double*>(base + off) : nullptr;
    off += align_up(count * (i64)sizeof(double), 256);
    return p;
  }
};

// the mirror scheme applies to the fused TMA kernel only
// fused kernel: only the TMA form has the mirror passes; unfused path: any GEMM (store mask + k_t2_finalize)
bool t2_sym_active(const mpccb_ctx* c) { return !c->sym_off && (c->use_tma || c->t2_unfused); }

// Launch list of the fused eris -> t2 kernel under the mirror scheme: every tile that holds at least one (a,b)
// block this rank computes (sym_compute), ordered like tile_coords (groups of column tiles x all row tiles).
int build_tile_list(mpccb_ctx* c) {
  c->sym.enabled = t2_sym_active(c) ? 1 : 0;
  c->sym.tiles = nullptr;
  c->n_tiles = c->t2_tiles_m * c->t2_tiles_n;
  if (!c->sym.enabled) return 0;
  if (!c->sym.peers) {               // no peer-mapped buffers: [0,a_begin) | own slab | [a_end,v), only the own slab mirrors
    c->sym.nranks = 3; c->sym.rank = 1;
    c->sym.bounds[0] = 0; c->sym.bounds[1] = c->a_begin; c->sym.bounds[2] = c->a_end; c->sym.bounds[3] = c->v;
  }
  c->sym.Tp[c->sym.rank] = c->T;
  if (c->t2_unfused) {
    // one byte per (a,b) block of the slab: bit 0 computed here, bit 1 mirrored from here, bits 4.. owner of b
    // (store mask of the GEMM and work list of k_t2_finalize); lives in the tile-list workspace
    std::vector<unsigned char> flags((size_t)c->na * c->v);
    int computed = 0;
    for (int al = 0; al < c->na; ++al)
      for (int b = 0; b < c->v; ++b) {
        bool mirror; int s;
        const bool comp = sym_compute(c->sym, c->a_begin + al, b, mirror, s);
        flags[(size_t)al * c->v + b] = (unsigned char)((comp ? 1 : 0) | (mirror ? 2 : 0) | (s << 4));
        computed += comp;
      }
    if ((i64)flags.size() > c->tiles_cap * (i64)sizeof(int2)) return fail(c, "block flags exceed their workspace");
    CUDA_OK(c, cudaMemcpyAsync(c->tiles_dev, flags.data(), flags.size(), cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(c, cudaStreamSynchronize(c->stream));
    c->n_tiles = computed;
    c->blkflags = reinterpret_cast<const unsigned char*>(c->tiles_dev);
    return 0;
  }
  c->blkflags = nullptr;
  std::vector<int2> list;
  list.reserve((size_t)c->n_tiles / 2 + 16);
  // group width of the raster: the CTAs resident at one time (2 per SM with 128 x 64 tiles) should cover a patch of
  // the tile grid that streams the fewest JK rows. Measured (ncu, (64,512,1536)): DRAM reads of the launch 12.9 GB with
  // 8 column tiles per group, 6.7 with 16, 5.0 with 24, 4.2 with 32 (0.4 GB algorithmic); time unchanged (47.9 ms)
  const int gw = c->raster_eris > 0 ? c->raster_eris : (c->t2_wn == 2 ? 32 : 12);
  for (int g0 = 0; g0 < c->t2_tiles_n; g0 += gw)
    for (int tm = 0; tm < c->t2_tiles_m; ++tm)
      for (int tn = g0; tn < std::min(g0 + gw, c->t2_tiles_n); ++tn) {
        bool need = false;
        for (int al = 0; al < c->t2_na && !need; ++al) {
          const int a = c->a_begin + tm * c->t2_na + al;
          if (a >= c->a_end) break;
          for (int bl = 0; bl < c->t2_nb; ++bl) {
            const int b = tn * c->t2_nb + bl;
            if (b >= c->v) break;
            bool mirror; int s;
            if (sym_compute(c->sym, a, b, mirror, s)) { need = true; break; }
          }
        }
        if (need) list.push_back(make_int2(tm, tn));
      }
  if ((i64)list.size() > c->tiles_cap) return fail(c, "tile list exceeds its workspace");
  CUDA_OK(c, cudaMemcpyAsync(c->tiles_dev, list.data(), sizeof(int2) * list.size(), cudaMemcpyHostToDevice, c->stream));
  CUDA_OK(c, cudaStreamSynchronize(c->stream));      // `list` is pageable host memory
  c->n_tiles = (int)list.size();
  c->sym.tiles = c->tiles_dev;
  return 0;
}

// Yvo through one TMA launch whose 128-row tiles hold 128/nocc whole values of a: when those fill >= 85 % of the
// tile (nocc = 64, 32, 22, 42, 110..128, ...). Other even nocc: power-of-two pieces (launch_yvo_tma_pieces).
bool yvo_single_launch(int o) { return o <= 128 && (128 / o) * o >= 109; }

// Flat-K form of the Yvo GEMM. K runs over (b, j) and the 4-D tensor map over T[a][b][i][j] takes 16 columns j of one b
// per step, so an occupied count that is not a multiple of 16 pays ceil(o/16)*16 columns per b: 32 for nocc = 21
// (34 % of the DMMA work wasted), 48 for 42. (Packing 16 / jw different b times jw columns into one step would fix it
// at the source, but a TMA box narrower than its swizzle span is padded to the span in shared memory, and without the
// 128-byte swizzle the fragment loads conflict 4-way.) Instead the slab is copied once per iteration into
// Tk[(a,i)][(b,j)] — j pitch jp, pad columns zero — where K is a plain contiguous index of length v*jp: two more
// HBM passes over t2 (< 4 % of the GEMM's time) against the wasted columns. Used when it saves more than 10 % of the
// K steps and the copy fits 24 GB.
bool yvo_flat_wanted(const mpccb_ctx* c) {
  if (!c->use_tma || c->yvo_flat_force < 0) return false;
  const double steps16 = (double)c->v * cdiv(c->o, TMA_TK);
  const double steps_flat = (double)cdiv((i64)c->v * c->jp, TMA_TK);
  const double bytes = 8.0 * c->na * c->o * (double)c->v * c->jp;
  if (c->yvo_flat_force > 0) return true;
  return steps_flat < 0.9 * steps16 && bytes <= 24e9;
}

// Split-K factor of the TMA Yvo GEMM: with few output tiles per SM (a-slab sharding) the last wave is
// mostly idle; s K-slices turn T tiles into T*s shorter CTAs. Pick s <= 8 minimising ceil(T*s/148)/s.
int choose_yvo_splits(const mpccb_ctx* c) {
  const int o = c->o;
  if (!c->use_tma || !(c->yvo_flat || yvo_single_launch(o))) return 1;      // split-K only on the one-launch paths
  const int n_a = 128 / o;
  const i64 T = (c->yvo_flat ? (i64)cdiv(c->na * o, TMA_TM) : (i64)cdiv(c->na, n_a)) * cdiv(c->L, 128);
  const i64 nk = c->yvo_flat ? (i64)cdiv((i64)c->v * c->jp, TMA_TK) : (i64)c->v * cdiv(o, TMA_TK);
  if (c->yvo_split_force > 0) return (int)std::min<i64>(c->yvo_split_force, nk);
  double best = (double)cdiv(T, NSM);
  int best_s = 1;
  for (int s = 2; s <= 8; ++s) {
    if (nk / s < 64) break;
    const double w = (double)cdiv(T * s, NSM) / s + 0.01 * s;      // small bias against needless splitting
    if (w < best * 0.98) { best = w; best_s = s; }
  }
  return best_s;
}

// lays the scratch tensors out in the workspace; with base == nullptr only sizes it
i64 carve_workspace(mpccb_ctx* c, unsigned char* base) {
  const i64 o = c->o, v = c->v, L = c->L, nov = c->nov, ldk = c->ldk, na = c->na;
  Carve w{base, 0};
  c->Lvo = w.take(L * v * c->jp);
  c->t1 = w.take(nov);
  c->t1T = w.take(nov);
  c->t1n = w.take(nov);
  c->Xov = w.take(L * o * v);
  c->Xoo = w.take(2 * L * o * o);           // [Xoo | Joo] adjacent: one GEMM dresses both with t1
  c->Joo = c->Xoo ? c->Xoo + L * o * o : nullptr;
  c->Xoo_ws = c->Xoo;
  c->X = w.take(L);
  c->XT = w.take(2 * L * o * v);            // [XT | JT]
  c->JT = c->XT ? c->XT + L * o * v : nullptr;
  c->JK = w.take(nov * ldk);
  c->E1K = w.take(nov * ldk);
  c->Wov = w.take(L * o * v);
  c->w = w.take(nov);
  c->w_ws = c->w;
  c->F2 = w.take(nov);
  c->Fvo = w.take(nov);
  c->Fvo_ws = c->Fvo;
  c->OmP_count = (i64)c->t2_tiles_n * na * o;
  c->OmP = w.take(c->OmP_count);
  const bool symm = t2_sym_active(c);
  c->tiles_cap = symm ? (i64)c->t2_tiles_m * c->t2_tiles_n : 0;
  c->tiles_dev = reinterpret_cast<int2*>(w.take(c->tiles_cap));          // int2 = one double each
  c->sym.OmPm = w.take(symm ? (i64)c->t2_tiles_m * nov : 0);
  c->Y = w.take(nov * L);
  c->yvo_flat = yvo_flat_wanted(c) ? 1 : 0;
  c->Tk = w.take(c->yvo_flat ? na * o * v * c->jp : 0);
  c->yvo_split = choose_yvo_splits(c);
  c->YP = w.take(c->yvo_split > 1 ? (i64)c->yvo_split * na * o * L : 0);
  i64 p = 0;
  auto mx = [&](i64 x) { if (x > p) p = x; };
  mx(splitk_scratch((int)v, (int)o, (int)(L * o), 128, 64));   // S5, S6
  mx(splitk_scratch((int)na, (int)o, (int)(o * L), 128, 64));  // S7
  mx(splitk_scratch((int)v, (int)o, (int)(na * L), 128, 64));  // G2
  mx(splitk_scratch((int)o, (int)o, (int)(na * L), 128, 64));  // S8
  mx((DIIS_ROWS_GRID + 1) * 7);                                // DIIS dot partials (row-owned vector)
  mx((i64)2 * NSM * 128 * 128);                                // split-K partials of mpccb_gemm_nt (tensor engine): <= 2 waves of tiles
  mx((DIIS_COLS_GRID + 1) * 7);                                //   "   (column-owned vector)
  c->P_count = p;
  c->P = w.take(p);
  c->P2_count = std::max(splitk_scratch((int)v, (int)o, (int)(L * o), 128, 64),      // F2 (mpccb_dress_rest)
                         splitk_scratch((int)o, (int)o, (int)(L * na), 128, 64));    // Lov.Yvo (mpccb_doubles_end)
  c->P2 = w.take(c->P2_count);
  c->OmRep = w.take(nov);
  c->OmLoc = w.take(nov);
  c->OmTot = w.take(nov);
  c->Z = w.take(o * o);
  c->red = w.take(nov + 8);
  c->scal = w.take(8192);
  return w.off;
}

}  // namespace

namespace {
// tile geometry of the fused eris -> t2 kernel: n_a x n_b whole (i,j) blocks per 128 x (32*wn) tile.
// wn = 0 picks the better-filled of the two tile shapes. When whole blocks fill the tile badly
// (e.g. nocc = 80: one 80 x 80 block in a 128 x 128 tile = 39 %) or nocc > 128, the path is unfused:
// generic-tile GEMM scattering e into T[a][b][i][j], then the in-place k_t2_finalize pass.
double t2_fill(int o, int wn) {
  if (o > 128 || 32 * wn < o) return 0.0;
  return (double)((128 / o) * o) / 128.0 * (double)(((32 * wn) / o) * o) / (32.0 * wn);
}
void set_t2_geometry(mpccb_ctx* c, int wn) {
  if (wn == 0) wn = (t2_fill(c->o, 2) >= 0.95 * t2_fill(c->o, 4) && t2_fill(c->o, 2) > 0.0) ? 2 : 4;
  if (t2_fill(c->o, wn) == 0.0) wn = 4;
  const double fill = t2_fill(c->o, wn);
  c->t2_unfused = (c->force_unfused || fill < 0.70) ? 1 : 0;
  c->t2_wn = wn;
  if (c->t2_unfused) {
    c->t2_na = 1; c->t2_nb = 1;
    c->t2_tiles_m = c->na; c->t2_tiles_n = c->v;     // OmP[b][(a,i)]
  } else {
    c->t2_na = 128 / c->o; c->t2_nb = (32 * wn) / c->o;
    c->t2_tiles_m = cdiv(c->na, c->t2_na);
    c->t2_tiles_n = cdiv(c->v, c->t2_nb);
  }
}
}  // namespace

extern "C" {

const char* mpccb_version(void) { return "mpcc_b200 0.2 (sm_100a, DMMA.8x8x4, TMA)"; }

int mpccb_create(mpccb_ctx** out, int nocc, int nvir, int naux, int a_begin, int a_end, int device) {
  if (!out) return 1;
  *out = nullptr;
  mpccb_ctx* c = new mpccb_ctx();
  memset(c, 0, sizeof(*c));
  *out = c;   // returned even on failure so the caller can read the message
  if (nocc <= 0 || nvir <= 0 || naux <= 0) return fail(c, "sizes must be positive");
  if (a_begin < 0 || a_end > nvir || a_begin >= a_end) return fail(c, "bad slab [%d,%d) of %d", a_begin, a_end, nvir);
  // explicit limits of the index arithmetic (rows and tile ids are 32-bit, element offsets 64-bit):
  //   nov <= 2^21            32-bit GEMM row indices, (row, L) products below 2^31 elements per operand view
  //   nocc <= 2^15           umulhi index split of the fused epilogue
  //   nov * ldk, nov * nov   < 2^62 as byte counts (always true under the nov limit; kept for the record)
  //   (na*nvir) block count and (tiles_m * tiles_n) tile count of the eris launch < 2^31 (checked after the geometry)
  const i64 nov64 = (i64)nocc * nvir;
  if (nov64 > ((i64)1 << 21)) return fail(c, "nocc*nvir = %lld exceeds the supported 2^21", (long long)nov64);
  if (nocc > (1 << 15)) return fail(c, "nocc = %d exceeds the supported 2^15", nocc);
  if ((i64)naux > ((i64)1 << 24)) return fail(c, "naux = %d exceeds the supported 2^24", naux);
  c->o = nocc; c->v = nvir; c->L = naux; c->device = device;
  c->a_begin = a_begin; c->a_end = a_end; c->na = a_end - a_begin;
  c->nov = (i64)nocc * nvir;
  c->ldk = (int)align_up(naux, 2);
  c->jp = (int)align_up(nocc, 2);
  c->use_tma = 1;
  c->mapped_narrow_k = 1024;
  c->profile = 0;
  c->prof_ev = new std::vector<cudaEvent_t>[ST_COUNT];
  c->attr_set = new std::vector<const void*>();
  set_t2_geometry(c, 0);
  if ((i64)c->t2_tiles_m * c->t2_tiles_n > (i64)INT_MAX || (i64)c->na * nvir > (i64)INT_MAX)
    return fail(c, "eris launch would exceed 2^31 tiles");
  CUDA_OK(c, cudaSetDevice(device));
  c->ws_bytes = 0;
  return 0;
}

void mpccb_destroy(mpccb_ctx* ctx) {
  if (!ctx) return;
  if (ctx->prof_ev) {
    for (int s = 0; s < ST_COUNT; ++s)
      for (cudaEvent_t e : ctx->prof_ev[s]) cudaEventDestroy(e);
    delete[] ctx->prof_ev;
  }
  delete ctx->attr_set;
  if (ctx->side) {
    cudaStreamDestroy(ctx->side);
    cudaEventDestroy(ctx->ev_fork);
    cudaEventDestroy(ctx->ev_join);
  }
  delete ctx;
}

int mpccb_profile_stages(void) { return ST_COUNT; }
const char* mpccb_profile_stage_name(int stage) { return (stage >= 0 && stage < ST_COUNT) ? kStageNames[stage] : ""; }

// synchronises, then returns per-stage summed milliseconds and launch-group counts since the last read
int mpccb_profile_read(mpccb_ctx* c, double* ms_out, long long* count_out) {
  CUDA_OK(c, enter_device(c));
  CUDA_OK(c, cudaStreamSynchronize(c->stream));
  for (int s = 0; s < ST_COUNT; ++s) {
    double tot = 0.0;
    for (int k = 0; k + 1 < c->prof_used[s]; k += 2) {
      float ms = 0.f;
      CUDA_OK(c, cudaEventElapsedTime(&ms, c->prof_ev[s][k], c->prof_ev[s][k + 1]));
      tot += ms;
    }
    ms_out[s] = tot;
    count_out[s] = c->prof_used[s] / 2;
    c->prof_used[s] = 0;
  }
  return 0;
}

const char* mpccb_last_error(const mpccb_ctx* ctx) { return ctx ? ctx->err : "null context"; }

int mpccb_set_stream(mpccb_ctx* c, void* s) { c->stream = reinterpret_cast<cudaStream_t>(s); return 0; }

int mpccb_synchronize(mpccb_ctx* c) { CUDA_OK(c, cudaStreamSynchronize(c->stream)); return 0; }

long long mpccb_launch_count(const mpccb_ctx* c) { return c->launches; }

int mpccb_set_option(mpccb_ctx* c, const char* name, long long value) {
  if (!strcmp(name, "tma")) {
    c->use_tma = value != 0;
    if (!c->ws) set_t2_geometry(c, c->use_tma ? 0 : 4);   // the generic fallback kernel has 128 x 128 tiles
    return 0;
  }
  if (!strcmp(name, "profile")) { c->profile = value != 0; return 0; }
  if (!strcmp(name, "xvo_generic")) { c->xvo_generic = value != 0; return 0; }
  if (!strcmp(name, "mapped_narrow_k")) { c->mapped_narrow_k = (int)value; return 0; }
  if (!strcmp(name, "yvo_flat")) {        // 1 / -1: force the flat-K Yvo form on / off (0 = automatic; before bind_workspace)
    if (c->ws) return fail(c, "set yvo_flat before binding the workspace");
    c->yvo_flat_force = (int)value;
    return 0;
  }
  if (!strcmp(name, "t2_sym")) {          // 0: compute every (a,b) block (no mirror writes); before bind_workspace
    if (c->ws) return fail(c, "set t2_sym before binding the workspace");
    c->sym_off = value == 0;
    return 0;
  }
  if (!strcmp(name, "raster_yvo") || !strcmp(name, "raster_eris")) {   // A/B switch: raster group width (0 = default)
    if (value < 0 || value > 4096) return fail(c, "%s = %lld out of range", name, (long long)value);
    if (name[7] == 'y') c->raster_yvo = (int)value;
    else {
      if (c->ws) return fail(c, "set raster_eris before binding the workspace");
      c->raster_eris = (int)value;
    }
    return 0;
  }
  if (!strcmp(name, "yvo_split")) {       // force the split-K factor of the TMA Yvo GEMM (0 = automatic; before bind_workspace)
    if (c->ws) return fail(c, "set yvo_split before binding the workspace");
    c->yvo_split_force = (int)value;
    return 0;
  }
  if (!strcmp(name, "t2_unfused")) {      // 1: always GEMM -> T, then in-place k_t2_finalize (before bind_workspace)
    if (c->ws) return fail(c, "set t2_unfused before binding the workspace");
    c->force_unfused = value != 0;
    set_t2_geometry(c, c->use_tma ? 0 : 4);
    return 0;
  }
  if (!strcmp(name, "t2_wide_tiles")) {   // 1: 128x128 tiles, one CTA/SM; 0: 128x64, two CTAs/SM (before bind_workspace)
    if (c->ws) return fail(c, "set t2_wide_tiles before binding the workspace");
    set_t2_geometry(c, value ? 4 : 2);
    return 0;
  }
  return fail(c, "unknown option %s", name);
}

size_t mpccb_workspace_bytes(const mpccb_ctx* c) {
  mpccb_ctx tmp = *c;
  return (size_t)carve_workspace(&tmp, nullptr);
}

int mpccb_bind_workspace(mpccb_ctx* c, void* ws, size_t bytes) {
  CUDA_OK(c, enter_device(c));
  const size_t need = mpccb_workspace_bytes(c);
  if (bytes < need) return fail(c, "workspace too small: %zu < %zu", bytes, need);
  if ((reinterpret_cast<uintptr_t>(ws) & 255) != 0) return fail(c, "workspace must be 256-byte aligned");
  c->ws = static_cast<unsigned char*>(ws);
  c->ws_bytes = bytes;
  carve_workspace(c, c->ws);
  // K padding columns of JK / E1K (ldk > L) must read as zero for the TMA path
  CUDA_OK(c, cudaMemsetAsync(c->JK, 0, sizeof(double) * c->nov * c->ldk, c->stream));
  CUDA_OK(c, cudaMemsetAsync(c->E1K, 0, sizeof(double) * c->nov * c->ldk, c->stream));
  return build_tile_list(c);
}

int mpccb_bind_integrals(mpccb_ctx* c, const double* Loo, const double* Lov, const double* Lvv,
                         const double* foo, const double* fvv, const double* eo, const double* ev) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws) return fail(c, "bind the workspace first");
  c->Loo = Loo; c->Lov = Lov; c->Lvv = Lvv; c->foo = foo; c->fvv = fvv; c->eo = eo; c->ev = ev;
  // Lvo[L][b][j] = Lov[L][j][b], j pitch jp (the pad column, odd nocc only, stays zero)
  if (c->jp != c->o) CUDA_OK(c, cudaMemsetAsync(c->Lvo, 0, sizeof(double) * c->L * c->v * c->jp, c->stream));
  dim3 grid(cdiv(c->v, 32), cdiv(c->o, 32), c->L), block(32, 8);
  k_transpose_batched<<<grid, block, 0, c->stream>>>(Lov, c->Lvo, c->o, c->v, c->jp);
  LAUNCH_OK(c, "k_transpose_batched(Lov)");
  return 0;
}

int mpccb_bind_t2(mpccb_ctx* c, double* T) {
  c->T = T;
  if (c->ws) return build_tile_list(c);
  return 0;
}

// Several ranks with peer-mapped t2 buffers (symmetric memory): rank `rank` of `nranks` owns a in
// [slab_begin[rank], slab_begin[rank+1]); T_peers_dev[r] is rank r's t2 buffer as seen from this device. Enables the
// cross-rank half of the mirror scheme: blocks (a,b) with b in another rank's slab are computed by one of the two
// ranks only, which stores the mirror block straight into the other's buffer. The caller must run a cross-rank
// barrier on the stream between mpccb_doubles_begin and mpccb_doubles_end.
int mpccb_bind_t2_peers(mpccb_ctx* c, int nranks, int rank, const int* slab_begin, double* const* T_peers_dev) {
  if (nranks < 1 || nranks > 16 || rank < 0 || rank >= nranks) return fail(c, "bad rank layout");
  if (slab_begin[rank] != c->a_begin || slab_begin[rank + 1] != c->a_end) return fail(c, "slab bounds do not match the context");
  c->sym.peers = nranks > 1 ? 1 : 0;
  c->sym.nranks = nranks; c->sym.rank = rank;
  for (int r = 0; r <= nranks; ++r) c->sym.bounds[r] = slab_begin[r];
  for (int r = 0; r < nranks; ++r) c->sym.Tp[r] = T_peers_dev[r];
  c->T = T_peers_dev[rank];
  if (c->ws) return build_tile_list(c);
  return 0;
}

int mpccb_bind_jk_exchange(mpccb_ctx* c, double* jk_local_dev, double* const* peer_jk_dev, int npeers,
                           double* jk_multicast_dev) {
  if (!c->ws) return fail(c, "bind the workspace first");
  if (!jk_local_dev) return fail(c, "null JK buffer");
  if (npeers < 0 || npeers > 16) return fail(c, "at most 16 peers");
  c->JK = jk_local_dev;
  c->w_targets.n = 0;        // full exchange: every rank has every JK row and forms w itself
  c->w = c->w_ws;
  c->Fvo = c->Fvo_ws;
  if (jk_multicast_dev) {
    c->jk_targets.n = 1;
    c->jk_targets.p[0] = jk_multicast_dev;
  } else {
    c->jk_targets.n = npeers;
    for (int r = 0; r < npeers; ++r) c->jk_targets.p[r] = peer_jk_dev[r];
    if (npeers == 0) { c->jk_targets.n = 1; c->jk_targets.p[0] = jk_local_dev; }
  }
  return 0;
}

// Subset form of the fused exchange. The mirror scheme lets rank q compute only the slab pairs (q, s) with
// (s - q) mod P <= P/2, so rank r's JK rows are needed by the P/2 + 1 ranks q = r - d, d = 0..P/2, only: one NVLink
// store per needed peer instead of a multicast to everybody (8 ranks: 5 x 50 MB leave and 4 x 50 MB enter each GPU
// instead of 7 x 50 MB entering). The one consumer that wanted every row on every rank, w = Jvo.X (:128, :133), is
// computed for the own slab inside mpccb_dress and broadcast, together with the slab's rows of Fvo = w - F2 (so that
// F2, :133, is a GEMM over the slab's rows only): the symmetric allocation is nov*ldk (JK) + nov (w) + nov (Fvo)
// doubles. peer_jk_dev: all nranks buffers in rank order; w goes through w_multicast_dev (the
// multicast address of the allocation) when given, else through the peer pointers.
// Pure host rule of the subset exchange: does rank `dst` need the JK rows of rank `src`'s slab? It does iff it computes
// a slab pair (dst, src) under the mirror scheme, i.e. (src - dst) mod P <= P/2 (checked against mpccb_sym_query in
// tests/test_host_logic.py).
int mpccb_jk_subset_needs(int nranks, int src, int dst) {
  if (nranks < 1 || src < 0 || dst < 0 || src >= nranks || dst >= nranks) return 0;
  return ((src - dst + nranks) % nranks) <= nranks / 2 ? 1 : 0;
}

int mpccb_bind_jk_subset(mpccb_ctx* c, double* jk_local_dev, double* const* peer_jk_dev, int nranks, int rank,
                         double* w_multicast_dev) {
  if (!c->ws) return fail(c, "bind the workspace first");
  if (!jk_local_dev || nranks < 2 || nranks > 16 || rank < 0 || rank >= nranks) return fail(c, "bad rank layout");
  c->JK = jk_local_dev;
  c->w_off = c->nov * c->ldk;
  c->w = jk_local_dev + c->w_off;
  c->Fvo = c->w + c->nov;          // Fvo = w - F2 follows w in the symmetric allocation
  c->jk_targets.n = 0;
  c->jk_targets.p[c->jk_targets.n++] = peer_jk_dev[rank];              // own buffer first
  for (int q = 0; q < nranks; ++q)
    if (q != rank && mpccb_jk_subset_needs(nranks, rank, q)) c->jk_targets.p[c->jk_targets.n++] = peer_jk_dev[q];
  if (w_multicast_dev) {
    c->w_targets.n = 1;
    c->w_targets.p[0] = w_multicast_dev;
  } else {
    c->w_targets.n = nranks;
    for (int r = 0; r < nranks; ++r) c->w_targets.p[r] = peer_jk_dev[r];
  }
  return 0;
}

// Xoo = Lov.t1^T (:120) and Joo = Xoo + Loo (:123) do not depend on the virtual slab: replicated they cost every rank
// the same GEMM however many ranks there are. Sharded form: rank r computes the rows of its slab of the AUXILIARY
// index, l in [L r / P, L (r+1) / P), and stores its Xoo rows into every rank's [Xoo | Joo] buffer (2 L o^2 doubles
// in symmetric memory; only the Xoo half travels, Joo is completed locally) — through the multicast address when
// given, else one store per peer. mpccb_dress then stops
// after the exchange has been issued (the slab's Xvo GEMM hides its latency); the caller runs a cross-rank barrier
// on the stream and calls mpccb_dress_rest. xoo_local_dev == nullptr: back to the replicated form.
int mpccb_bind_xoo_exchange(mpccb_ctx* c, double* xoo_local_dev, double* const* peer_xoo_dev, int nranks, int rank,
                            double* xoo_multicast_dev) {
  if (!c->ws) return fail(c, "bind the workspace first");
  if (!xoo_local_dev) {
    c->Xoo = c->Xoo_ws; c->Joo = c->Xoo + (i64)c->L * c->o * c->o;
    c->xoo_targets.n = 0;
    return 0;
  }
  if (nranks < 2 || nranks > 16 || rank < 0 || rank >= nranks) return fail(c, "bad rank layout");
  if (c->L < nranks) return fail(c, "fewer auxiliary functions than ranks");
  c->Xoo = xoo_local_dev; c->Joo = c->Xoo + (i64)c->L * c->o * c->o;
  c->xoo_l0 = (int)((i64)c->L * rank / nranks);
  c->xoo_l1 = (int)((i64)c->L * (rank + 1) / nranks);
  c->xoo_targets.n = 0;
  if (xoo_multicast_dev) {
    c->xoo_targets.p[c->xoo_targets.n++] = xoo_multicast_dev;
  } else {
    for (int r = 0; r < nranks; ++r)
      if (peer_xoo_dev[r] != xoo_local_dev) c->xoo_targets.p[c->xoo_targets.n++] = peer_xoo_dev[r];
  }
  return 0;
}

}  // extern "C"

// ===========================================================================
// phase 1: T1 dressing + singles intermediates (reference :119-133)
// ===========================================================================
extern "C" int mpccb_dress_post(mpccb_ctx* c);

// Slab-local part: everything that carries the virtual index a is computed for a in [a_begin, a_end)
// only (JK/E1K rows, W columns, the W.Joo singles term); Xoo/Joo/X/F2 have no virtual index and are
// replicated (2 x 2 L o^2 v flop). With several ranks the caller all-gathers the JK rows and then
// calls mpccb_dress_post; on one rank mpccb_dress runs the post part itself.
extern "C" int mpccb_dress_rest(mpccb_ctx* c);
namespace {
int ensure_side(mpccb_ctx* c) {       // side stream + fork/join events, created on first use
  if (!c->side) {
    CUDA_OK(c, cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
    CUDA_OK(c, cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    CUDA_OK(c, cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming));
  }
  return 0;
}
}  // namespace
namespace { int launch_xvo_tma(mpccb_ctx* c); int launch_xtjt_tma(mpccb_ctx* c); }

extern "C" int mpccb_dress(mpccb_ctx* c, const double* t1_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws || !c->Lov) return fail(c, "workspace/integrals not bound");
  const int o = c->o, v = c->v, L = c->L, a0 = c->a_begin, na = c->na;
  const i64 nov = c->nov;
  const bool xoo_sharded = c->xoo_targets.n > 0;
  {
  StageTimer timer(c, ST_DRESS);
  if (t1_dev != c->t1)
    CUDA_OK(c, cudaMemcpyAsync(c->t1, t1_dev, sizeof(double) * nov, cudaMemcpyDeviceToDevice, c->stream));
  {  // t1T[a][i] = t1[i][a]
    dim3 grid(cdiv(v, 32), cdiv(o, 32), 1), block(32, 8);
    k_transpose_batched<<<grid, block, 0, c->stream>>>(c->t1, c->t1T, o, v, o);
    LAUNCH_OK(c, "k_transpose_batched(t1)");
  }
  {  // :120,:123  Xoo[(L,i)][j] = sum_a Lov[(L,i)][a] t1[j][a];  Joo = Xoo + Loo
     // replicated, or (Xoo exchange bound) the rows of this rank's L slab, stored into every rank's buffer
    const int l0 = xoo_sharded ? c->xoo_l0 : 0, l1 = xoo_sharded ? c->xoo_l1 : L;
    const i64 roff = (i64)l0 * o * o;
    Operand A = flat_operand(c->Lov + (i64)l0 * o * v, (l1 - l0) * o, v, 1);
    Operand B = flat_operand(c->t1, o, v, 1);
    Output C = flat_output(c->Xoo + roff, o, 1);
    C.base2 = c->Joo + roff; C.aux = c->Loo + roff;
    if (launch_gemm<CfgN64KK>(c, "Xoo", A, B, C, (l1 - l0) * o, o, v, 1, nullptr)) return 1;
    if (xoo_sharded) {
      const i64 n = (i64)(l1 - l0) * o * o;
      k_bcast_rows<<<(int)std::min<i64>(cdiv(n, 2048), 4 * NSM), 256, 0, c->stream>>>(c->Xoo, c->xoo_targets, roff, n);
      LAUNCH_OK(c, "k_bcast_rows");
    }
  }
  {  // :119  Xov[L][i][a] = sum_b Lvv[L][a][b] t1[i][b]      M=(L, a in slab) N=i K=b
    const int rc = launch_xvo_tma(c);
    if (rc < 0) return 1;
    if (rc == 0) {
    Operand A = flat_operand(c->Lvv + (i64)a0 * v, L * na, v, 1);
    A.rdiv = na; A.r_hi = (i64)v * v; A.r_lo = v;
    Operand B = flat_operand(c->t1, o, v, 1);
    Output C = flat_output(c->Xov + a0, 0, v);
    C.mdiv = na; C.m_hi = (i64)o * v; C.m_lo = 1;
    if (launch_gemm<CfgN64KK>(c, "Xvo", A, B, C, L * na, o, v, 1, nullptr)) return 1;
    }
  }
  }
  if (xoo_sharded) return 0;        // caller: cross-rank barrier on the stream, then mpccb_dress_rest
  return mpccb_dress_rest(c);
}

// Second half of the dressing: everything that reads Xoo / Joo for every auxiliary index.
extern "C" int mpccb_dress_rest(mpccb_ctx* c) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws || !c->Lov) return fail(c, "workspace/integrals not bound");
  const int o = c->o, v = c->v, L = c->L, a0 = c->a_begin, na = c->na;
  const i64 nov = c->nov;
  {
  StageTimer timer(c, ST_DRESS);
  if (c->xoo_targets.n > 0) {   // :123 for the L slabs computed elsewhere (only Xoo travels)
    const i64 n = (i64)L * o * o;
    k_add_outside<<<4 * NSM, 256, 0, c->stream>>>(c->Xoo, c->Loo, c->Joo, (i64)c->xoo_l0 * o * o,
                                                  (i64)c->xoo_l1 * o * o, n);
    LAUNCH_OK(c, "k_add_outside");
  }
  // :121  X[L] = trace_i Xoo[L]
  k_trace<<<cdiv(L, 4), 128, 0, c->stream>>>(c->Xoo, c->X, L, o);
  LAUNCH_OK(c, "k_trace");
  // :133 second half   F2[b][j] = sum_{(L,i)} Lov[(L,i)][b] Xoo[(L,i)][j]      (replicated)
  // Needs only Xoo and is compute-bound, while [XT | JT] is bound by its output and k_dress_finalize by HBM (one rank)
  // or by the NVLink stores of the JK exchange (several ranks): it runs on a side stream next to them.
  if (ensure_side(c)) return 1;
  CUDA_OK(c, cudaEventRecord(c->ev_fork, c->stream));
  CUDA_OK(c, cudaStreamWaitEvent(c->side, c->ev_fork, 0));
  {
    cudaStream_t main_stream = c->stream;
    double* main_P = c->P;
    const i64 main_P_count = c->P_count;
    c->stream = c->side; c->P = c->P2; c->P_count = c->P2_count;
    // subset exchange: only the slab's rows b of F2 (Fvo = w - F2 is broadcast with w below), else all of them
    const bool slab_only = c->w_targets.n > 0;
    const int fb0 = slab_only ? a0 : 0, fnb = slab_only ? na : v;
    Operand A = flat_operand(c->Lov + fb0, fnb, 1, v);
    Operand B = flat_operand(c->Xoo, o, 1, o);
    const int rc = splitk_gemm<CfgN64RR>(c, "Xoo.Lov", A, B, c->F2 + (i64)fb0 * o, fnb, o, L * o, 1.0, 0.0);
    c->stream = main_stream; c->P = main_P; c->P_count = main_P_count;
    if (rc) return 1;
    CUDA_OK(c, cudaEventRecord(c->ev_join, c->side));
  }
  {  // [XT | JT][(L,i)][a] = sum_k [Xoo | Joo][(L,i)][k] t1[k][a], a in slab    (:124 last term, :127, :146)
    Operand A = flat_operand(c->Xoo, 2 * L * o, o, 1);
    Operand B = flat_operand(c->t1T + (i64)a0 * o, na, o, 1);
    Output C = flat_output(c->XT + a0, v, 1);
    const int rc = launch_xtjt_tma(c);
    if (rc < 0) return 1;
    if (rc == 0) {
    if (na >= 96) { if (launch_gemm<CfgBig16KK>(c, "XT|JT", A, B, C, 2 * L * o, na, o, 1, nullptr)) return 1; }
    else          { if (launch_gemm<CfgN64KK>(c, "XT|JT", A, B, C, 2 * L * o, na, o, 1, nullptr)) return 1; }
    }
  }
  {  // :124, :146 integrand, W = XT - Xov   (columns a of the slab -> JK/E1K rows of the slab)
    dim3 grid(cdiv(na, 32), cdiv(L, 32), o), block(32, 8);
    JkTargets jk = c->jk_targets;
    if (jk.n == 0) { jk.n = 1; jk.p[0] = c->JK; }
    k_dress_finalize<<<grid, block, 0, c->stream>>>(c->Xov, c->Lov, c->JT, c->XT, c->X, c->t1, jk, c->E1K,
                                                    c->Wov, L, o, v, c->ldk, a0, a0 + na);
    LAUNCH_OK(c, "k_dress_finalize");
    if (c->w_targets.n > 0) {
      // subset exchange: the slab's rows of w = Jvo.X (:128) and of Fvo = w - F2 (:133) go to every rank with the
      // integrals — nobody needs the other slabs' JK rows or a replicated F2 GEMM for them
      CUDA_OK(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
      k_gemv_rows_bcast<<<cdiv((i64)na * o, 8), 256, 0, c->stream>>>(c->JK, c->X, c->F2, c->w_targets, c->w_off,
                                                                     c->w_off + nov, a0 * o, na * o, L, c->ldk);
      LAUNCH_OK(c, "k_gemv_rows_bcast");
    }
  }
  {  // :126-127  OmRep[a][i] = sum_{(L,j)} W[(L,j)][a] Joo[(L,j)][i],  a in slab
    Operand A = flat_operand(c->Wov + a0, na, 1, v);
    Operand B = flat_operand(c->Joo, o, 1, o);
    if (splitk_gemm<CfgN64RR>(c, "W.Joo", A, B, c->OmRep + (i64)a0 * o, na, o, L * o, 1.0, 0.0)) return 1;
  }
  CUDA_OK(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
  }
  if (na == v) return mpccb_dress_post(c);
  return 0;
}

// Needs every row of JK (all-gathered over the ranks): w = Jvo.X for all (b,j), Fvo = w - F2.
extern "C" int mpccb_dress_post(mpccb_ctx* c) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws || !c->Lov) return fail(c, "workspace/integrals not bound");
  StageTimer timer(c, ST_DRESS_POST);
  const i64 nov = c->nov;
  // :128 and first half of :133   w[(a,i)] = sum_L J[(a,i)][L] X[L]   (already exchanged in the subset form)
  if (c->w_targets.n > 0) return 0;     // subset exchange: w and Fvo arrived with the integrals
  k_gemv_rows<<<cdiv(nov, 8), 256, 0, c->stream>>>(c->JK, c->X, c->w, (int)nov, c->L, c->ldk);
  LAUNCH_OK(c, "k_gemv_rows");
  k_make_fvo<<<cdiv(nov, 256), 256, 0, c->stream>>>(c->w, c->F2, c->Fvo, nov);
  LAUNCH_OK(c, "k_make_fvo");
  return 0;
}

// ===========================================================================
// phase 2: doubles for the local slab (reference :135-147)
// ===========================================================================
namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// cuTensorMapEncodeTiled through the runtime's driver entry point: no link-time dependency on libcuda,
// so the library still loads on a machine without a driver (symbol/ABI tests).
EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// rank-`rank` FP64 tensor, dims/box innermost first, strides in bytes for dims 1.., 128-byte swizzle
int make_tensor_map(mpccb_ctx* c, CUtensorMap* map, const double* base, int rank, const cuuint64_t* dims,
                    const cuuint64_t* strides_bytes, const cuuint32_t* box) {
  EncodeTiledFn fn = encode_tiled_fn();
  if (!fn) return fail(c, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint32_t ones[5] = {1, 1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, (cuuint32_t)rank, const_cast<double*>(base), dims,
                  strides_bytes, box, ones, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(c, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return 0;
}

int make_map_2d(mpccb_ctx* c, CUtensorMap* map, const double* base, i64 K, i64 rows, i64 ld, int box_rows) {
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 8};
  cuuint32_t box[2] = {(cuuint32_t)TMA_TK, (cuuint32_t)box_rows};
  return make_tensor_map(c, map, base, 2, dims, strides, box);
}

template <int EPI, int WN>
int launch_tma(mpccb_ctx* c, const char* what, const CUtensorMap& mA, const CUtensorMap& mB,
               const TmaGemmParams& p, const T2Params& t2p) {
  typedef TmaCfg<WN> Cfg;
  const size_t smem = (EPI == 1 ? (size_t)Cfg::AREA_EPI1 : (size_t)Cfg::AREA_STORE) + TMA_BAR_BYTES;
  CUDA_OK(c, ensure_smem_attr(c, reinterpret_cast<const void*>(&gemm_tma_kernel<EPI, WN>), smem, true));
  SymInfo sym;
  memset(&sym, 0, sizeof(sym));
  int grid = p.tiles_total > 0 ? p.tiles_total : p.tiles_m * p.tiles_n;
  if (EPI == 1 && c->sym.enabled) { sym = c->sym; sym.Tp[sym.rank] = c->T; grid = c->n_tiles; }
  gemm_tma_kernel<EPI, WN><<<grid, Cfg::THREADS, smem, c->stream>>>(mA, mB, p, t2p, sym);
  LAUNCH_OK(c, what);
  return 0;
}

// fused eris -> t2: 128 x (32*WN) tiles; WN = 2 runs two CTAs per SM so that one CTA's epilogue
// overlaps the other's K loop (needs n_b = 64 / o >= 1)
template <int WN>
int launch_eris_t2_tma_wn(mpccb_ctx* c) {
  typedef TmaCfg<WN> Cfg;
  const int o = c->o, v = c->v, L = c->L;
  CUtensorMap mA, mB;
  if (make_map_2d(c, &mA, c->JK + (i64)c->a_begin * o * c->ldk, L, (i64)c->na * o, c->ldk, TMA_TM)) return 1;
  if (make_map_2d(c, &mB, c->JK, L, (i64)v * o, c->ldk, Cfg::TN)) return 1;
  T2Params t;
  t.T = c->T; t.eo = c->eo; t.ev = c->ev; t.Fvo = c->Fvo; t.OmP = c->OmP;
  t.o = o; t.v = v; t.a_begin = c->a_begin; t.a_end = c->a_end; t.n_a = c->t2_na; t.n_b = c->t2_nb; t.jp = c->jp;
  TmaGemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = c->na * o; p.N = v * o;
  p.tm_eff = c->t2_na * o; p.tn_eff = c->t2_nb * o;
  p.tiles_m = c->t2_tiles_m; p.tiles_n = c->t2_tiles_n;
  p.nk = cdiv(L, TMA_TK);
  p.a_mode = 0;
  p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = Cfg::TN * TMA_TK * 8;
  return launch_tma<1, WN>(c, "gemm_tma_kernel<t2>", mA, mB, p, t);
}

int launch_eris_t2_tma(mpccb_ctx* c) {
  return c->t2_wn == 2 ? launch_eris_t2_tma_wn<2>(c) : launch_eris_t2_tma_wn<4>(c);
}

int launch_yvo_tma(mpccb_ctx* c) {
  constexpr int TN = TmaCfg<4>::TN;
  const int o = c->o, v = c->v, L = c->L, jp = c->jp;
  const int n_a = 128 / o;
  CUtensorMap mA, mB;
  {  // T[a][b][i][j], box = (16 j, o i, 1 b, n_a a) -> smem rows (a,i) x 16 k
    // the j extent of the map is the LOGICAL nocc (columns past it are zero-filled by TMA); only the pitch is padded
    cuuint64_t dims[4] = {(cuuint64_t)o, (cuuint64_t)o, (cuuint64_t)v, (cuuint64_t)c->na};
    cuuint64_t strides[3] = {(cuuint64_t)jp * 8, (cuuint64_t)o * jp * 8, (cuuint64_t)v * o * jp * 8};
    cuuint32_t box[4] = {(cuuint32_t)TMA_TK, (cuuint32_t)o, 1, (cuuint32_t)n_a};
    if (make_tensor_map(c, &mA, c->T, 4, dims, strides, box)) return 1;
  }
  if (make_map_2d(c, &mB, c->Lvo, (i64)v * jp, L, (i64)v * jp, TN)) return 1;
  TmaGemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = c->na * o; p.N = L;
  p.tm_eff = n_a * o; p.tn_eff = TN;
  p.tiles_m = cdiv(c->na, n_a); p.tiles_n = cdiv(L, TN);
  p.nk = v * cdiv(o, TMA_TK);
  p.a_mode = 1; p.jsteps = cdiv(o, TMA_TK); p.o = jp; p.n_a = n_a;      // p.o: K columns of Lvo per b (the j pitch)
  p.a_bytes = (uint32_t)(TMA_TK * o * n_a * 8); p.b_bytes = TN * TMA_TK * 8;
  double* Yslab = c->Y + (i64)c->a_begin * o * L;
  p.C = Yslab; p.ldc = L;
  p.gw = c->raster_yvo > 0 ? c->raster_yvo : 12;   // 148 resident CTAs cover 12 x 12.3 tiles (22.9 -> 18.4 GB per launch)
  p.ksplit = 1;
  const int splits = c->yvo_split;
  if (splits > 1) {
    p.ksplit = splits; p.nk_split = cdiv(p.nk, splits);
    p.ksplit = cdiv(p.nk, p.nk_split);
    p.C = c->YP; p.split_stride = (i64)c->na * o * L;
  }
  T2Params t;
  memset(&t, 0, sizeof(t));
  const int ntiles = p.tiles_m * p.tiles_n;
  p.tiles_total = ntiles * p.ksplit;
  if (launch_tma<0, 4>(c, "gemm_tma_kernel<store>", mA, mB, p, t)) return 1;
  if (p.ksplit > 1)
    return reduce_splits(c, Yslab, c->YP, p.ksplit, (i64)c->na * o * L, p.split_stride, 1.0, 0.0);
  return 0;
}

// unfused form of :135-137,:139 for occupied counts that tile badly: e -> T[a][b][i][j] with the
// generic DMMA GEMM (any tile alignment), then t2 = (2e - e^T)/D in place per (a,b) block
// Yvo for an even nocc that does not divide 128: the occupied index of the rows is cut into power-of-two pieces
// (80 = 64 + 16, 22 = 16 + 4 + 2, ...); a piece of width w is one launch whose 128-row tiles hold 128/w values of a
// (4-D box 16 j x w i x 1 b x 128/w a), so every tile is full whatever nocc is. Rows go back to Y[(a,i)][L]
// through the index-mapped store (EPI 2). TMA needs 16-byte strides, hence even nocc; odd nocc stays on the generic
// kernel.
int launch_yvo_tma_pieces(mpccb_ctx* c) {
  constexpr int TN = TmaCfg<4>::TN;
  const int o = c->o, v = c->v, L = c->L, jp = c->jp;
  double* Yslab = c->Y + (i64)c->a_begin * o * L;
  CUtensorMap mB;
  if (make_map_2d(c, &mB, c->Lvo, (i64)v * jp, L, (i64)v * jp, TN)) return 1;
  int i0 = 0;
  for (int w = 128; w >= 1; w >>= 1) {
    while (o - i0 >= w) {
      const int n_a = 128 / w;
      CUtensorMap mA;
      cuuint64_t dims[4] = {(cuuint64_t)o, (cuuint64_t)o, (cuuint64_t)v, (cuuint64_t)c->na};
      cuuint64_t strides[3] = {(cuuint64_t)jp * 8, (cuuint64_t)o * jp * 8, (cuuint64_t)v * o * jp * 8};
      cuuint32_t box[4] = {(cuuint32_t)TMA_TK, (cuuint32_t)w, 1, (cuuint32_t)n_a};
      if (make_tensor_map(c, &mA, c->T, 4, dims, strides, box)) return 1;
      TmaGemmParams p;
      memset(&p, 0, sizeof(p));
      p.M = c->na * w; p.N = L;
      p.tm_eff = TMA_TM; p.tn_eff = TN;
      p.tiles_m = cdiv(c->na, n_a); p.tiles_n = cdiv(L, TN);
      p.nk = v * cdiv(o, TMA_TK);
      p.a_mode = 1; p.jsteps = cdiv(o, TMA_TK); p.o = jp; p.n_a = n_a; p.i0 = i0;
      p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = TN * TMA_TK * 8;
      p.ksplit = 1;
      p.out = flat_output(Yslab + (i64)i0 * L, 0, 1);
      p.out.mdiv = w; p.out.m_hi = (i64)o * L; p.out.m_lo = L;      // tile row m -> a = m / w, i = i0 + m % w
      T2Params t;
      memset(&t, 0, sizeof(t));
      if (launch_tma<2, 4>(c, "gemm_tma_kernel<yvo piece>", mA, mB, p, t)) return 1;
      i0 += w;
    }
  }
  return 0;
}

// :119 through the TMA-fed kernel: A = Lvv[L][a in slab][b] as a 3-D map (rows (L,a) of the slab are not equally
// spaced), B = t1[i][b], stored into Xov[L][i][a] through the index map (lanes along a). Needs 16-byte strides (even
// nvir) and a slab that tiles 128 rows (na a multiple of 128, or a divisor of it).
// Returns 1 if launched, 0 if the shape is not eligible (caller uses the generic kernel), -1 on error.
int launch_xvo_tma(mpccb_ctx* c) {
  const int o = c->o, v = c->v, L = c->L, na = c->na, a0 = c->a_begin;
  if (!c->use_tma || c->xvo_generic || (v & 1) || v < 16) return 0;
  if (!(na % TMA_TM == 0 || TMA_TM % na == 0)) return 0;
  const int ba = std::min(na, TMA_TM), bl = TMA_TM / ba;
  CUtensorMap mA, mB;
  {
    cuuint64_t dims[3] = {(cuuint64_t)v, (cuuint64_t)na, (cuuint64_t)L};
    cuuint64_t strides[2] = {(cuuint64_t)v * 8, (cuuint64_t)v * v * 8};
    cuuint32_t box[3] = {(cuuint32_t)TMA_TK, (cuuint32_t)ba, (cuuint32_t)bl};
    if (make_tensor_map(c, &mA, c->Lvv + (i64)a0 * v, 3, dims, strides, box)) return -1;
  }
  const bool narrow = o <= 64;
  const int TN = narrow ? TmaCfg<2>::TN : TmaCfg<4>::TN;
  if (make_map_2d(c, &mB, c->t1, v, o, v, TN)) return -1;
  TmaGemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = L * na; p.N = o;
  p.tm_eff = TMA_TM; p.tn_eff = TN;
  p.tiles_m = cdiv(p.M, TMA_TM); p.tiles_n = cdiv(o, TN);
  p.nk = cdiv(v, TMA_TK);
  p.a_mode = 2; p.o = na;
  p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = (uint32_t)(TN * TMA_TK * 8);
  p.ksplit = 1;
  p.out = flat_output(c->Xov + a0, 0, v);
  p.out.mdiv = na; p.out.m_hi = (i64)o * v; p.out.m_lo = 1;
  T2Params t;
  memset(&t, 0, sizeof(t));
  const int rc = narrow ? launch_tma<2, 2>(c, "gemm_tma_kernel<Xvo>", mA, mB, p, t)
                        : launch_tma<2, 4>(c, "gemm_tma_kernel<Xvo>", mA, mB, p, t);
  return rc ? -1 : 1;
}

// [XT | JT] = [Xoo | Joo] . t1 (:124, :127, :146) through the TMA-fed kernel. K = nocc only: the product is bound by
// its 2 L o na output, which this kernel stores in full rows. Needs even nocc (16-byte strides).
// Returns 1 if launched, 0 if not eligible, -1 on error.
int launch_xtjt_tma(mpccb_ctx* c) {
  const int o = c->o, v = c->v, L = c->L, na = c->na, a0 = c->a_begin;
  const double* Bp = c->t1T + (i64)a0 * o;
  if (!c->use_tma || c->xvo_generic || (o & 1) || o < 16 || na < 64) return 0;
  if (((reinterpret_cast<uintptr_t>(c->Xoo) | reinterpret_cast<uintptr_t>(Bp)) & 15) != 0) return 0;
  const bool narrow = na <= 64;
  const int TN = narrow ? TmaCfg<2>::TN : TmaCfg<4>::TN;
  CUtensorMap mA, mB;
  if (make_map_2d(c, &mA, c->Xoo, o, (i64)2 * L * o, o, TMA_TM)) return -1;
  if (make_map_2d(c, &mB, Bp, o, na, o, TN)) return -1;
  TmaGemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = 2 * L * o; p.N = na;
  p.tm_eff = TMA_TM; p.tn_eff = TN;
  p.tiles_m = cdiv(p.M, TMA_TM); p.tiles_n = cdiv(na, TN);
  p.nk = cdiv(o, TMA_TK);
  p.a_mode = 0;
  p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = (uint32_t)(TN * TMA_TK * 8);
  p.ksplit = 1;
  p.C = c->XT + a0; p.ldc = v;
  T2Params t;
  memset(&t, 0, sizeof(t));
  const int rc = narrow ? launch_tma<0, 2>(c, "gemm_tma_kernel<XT|JT>", mA, mB, p, t)
                        : launch_tma<0, 4>(c, "gemm_tma_kernel<XT|JT>", mA, mB, p, t);
  return rc ? -1 : 1;
}

int launch_eris_t2_unfused(mpccb_ctx* c) {
  const int o = c->o, v = c->v, L = c->L;
  Operand A = flat_operand(c->JK + (i64)c->a_begin * o * c->ldk, c->na * o, c->ldk, 1);
  Operand B = flat_operand(c->JK, v * o, c->ldk, 1);
  Output C = flat_output(c->T, 0, 0);
  C.mdiv = o; C.m_hi = (i64)v * o * c->jp; C.m_lo = c->jp;   // m = (a_l, i)
  C.ndiv = o; C.n_hi = (i64)o * c->jp;     C.n_lo = 1;       // n = (b, j)
  const unsigned char* flags = c->sym.enabled ? c->blkflags : nullptr;     // mirror scheme (build_tile_list)
  C.keep = flags; C.keep_ld = v;
  if (c->use_tma) {
    // JK rows are a plain K-major matrix whatever nocc is: the TMA-fed kernel with the index-mapped, masked store
    constexpr int TN = TmaCfg<4>::TN;
    CUtensorMap mA, mB;
    if (make_map_2d(c, &mA, c->JK + (i64)c->a_begin * o * c->ldk, L, (i64)c->na * o, c->ldk, TMA_TM)) return 1;
    if (make_map_2d(c, &mB, c->JK, L, (i64)v * o, c->ldk, TN)) return 1;
    TmaGemmParams p;
    memset(&p, 0, sizeof(p));
    p.M = c->na * o; p.N = v * o;
    p.tm_eff = TMA_TM; p.tn_eff = TN;
    p.tiles_m = cdiv(p.M, TMA_TM); p.tiles_n = cdiv(p.N, TN);
    p.nk = cdiv(L, TMA_TK);
    p.a_mode = 0;
    p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = TN * TMA_TK * 8;
    p.ksplit = 1;
    p.out = C;
    T2Params t;
    memset(&t, 0, sizeof(t));
    if (launch_tma<2, 4>(c, "gemm_tma_kernel<scatter>", mA, mB, p, t)) return 1;
  } else if (launch_gemm<CfgBigKK>(c, "eris(unfused)", A, B, C, c->na * o, v * o, L, 1, nullptr)) {
    return 1;
  }
  dim3 grid(v, c->na);
  SymInfo sym = c->sym;
  sym.Tp[sym.rank] = c->T;
  k_t2_finalize<<<grid, 256, 0, c->stream>>>(c->T, c->eo, c->ev, c->Fvo, c->OmP, o, c->jp, v, c->a_begin, c->na, sym, flags);
  LAUNCH_OK(c, "k_t2_finalize");
  return 0;
}

int launch_eris_t2(mpccb_ctx* c) {
  if (c->t2_unfused) return launch_eris_t2_unfused(c);
  if (c->use_tma) return launch_eris_t2_tma(c);
  typedef CfgBigKK Cfg;
  const int o = c->o, v = c->v, L = c->L;
  Operand A = flat_operand(c->JK + (i64)c->a_begin * o * c->ldk, c->na * o, c->ldk, 1);
  Operand B = flat_operand(c->JK, v * o, c->ldk, 1);
  T2Params p;
  p.T = c->T; p.eo = c->eo; p.ev = c->ev; p.Fvo = c->Fvo; p.OmP = c->OmP;
  p.o = o; p.v = v; p.a_begin = c->a_begin; p.a_end = c->a_end; p.n_a = c->t2_na; p.n_b = c->t2_nb; p.jp = c->jp;
  GemmDims d;
  d.M = c->na * o; d.N = v * o; d.K = L;
  d.tm_eff = c->t2_na * o; d.tn_eff = c->t2_nb * o;
  d.tiles_m = c->t2_tiles_m; d.tiles_n = c->t2_tiles_n;
  d.kchunk = L;
  CUDA_OK(c, ensure_smem_attr(c, reinterpret_cast<const void*>(&gemm_t2_kernel<Cfg>), T2Smem<Cfg>::BYTES, false));
  gemm_t2_kernel<Cfg><<<d.tiles_m * d.tiles_n, Cfg::NTH, T2Smem<Cfg>::BYTES, c->stream>>>(A, B, p, d);
  LAUNCH_OK(c, "gemm_t2_kernel");
  return 0;
}

// Yvo with K flattened (yvo_flat_wanted): copy the slab into Tk[(a,i)][(b,j)], then a plain K-major product
int launch_yvo_flat(mpccb_ctx* c) {
  constexpr int TN = TmaCfg<4>::TN;
  const int o = c->o, v = c->v, L = c->L, jp = c->jp, na = c->na;
  const i64 K = (i64)v * jp;
  k_t2_flatten_k<<<dim3(v, na), 256, 0, c->stream>>>(c->T, c->Tk, o, jp, v);
  LAUNCH_OK(c, "k_t2_flatten_k");
  CUtensorMap mA, mB;
  if (make_map_2d(c, &mA, c->Tk, K, (i64)na * o, K, TMA_TM)) return 1;
  if (make_map_2d(c, &mB, c->Lvo, K, L, K, TN)) return 1;
  TmaGemmParams p;
  memset(&p, 0, sizeof(p));
  p.M = na * o; p.N = L;
  p.tm_eff = TMA_TM; p.tn_eff = TN;
  p.tiles_m = cdiv(p.M, TMA_TM); p.tiles_n = cdiv(L, TN);
  p.nk = (int)cdiv(K, TMA_TK);
  p.a_mode = 0;
  p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = TN * TMA_TK * 8;
  double* Yslab = c->Y + (i64)c->a_begin * o * L;
  p.C = Yslab; p.ldc = L;
  p.gw = c->raster_yvo > 0 ? c->raster_yvo : 12;   // 148 resident CTAs cover 12 x 12.3 tiles (22.9 -> 18.4 GB per launch)
  p.ksplit = 1;
  const int splits = c->yvo_split;
  if (splits > 1) {
    p.ksplit = splits; p.nk_split = cdiv(p.nk, splits);
    p.ksplit = cdiv(p.nk, p.nk_split);
    p.C = c->YP; p.split_stride = (i64)na * o * L;
  }
  T2Params t;
  memset(&t, 0, sizeof(t));
  p.tiles_total = p.tiles_m * p.tiles_n * p.ksplit;
  if (launch_tma<0, 4>(c, "gemm_tma_kernel<yvo flat>", mA, mB, p, t)) return 1;
  if (p.ksplit > 1)
    return reduce_splits(c, Yslab, c->YP, p.ksplit, (i64)na * o * L, p.split_stride, 1.0, 0.0);
  return 0;
}

int launch_yvo(mpccb_ctx* c) {
  const int o = c->o, v = c->v, L = c->L;
  if (c->yvo_flat && c->Tk) return launch_yvo_flat(c);
  // TMA strides must be multiples of 16 bytes: the j pitch of T and Lvo is nocc rounded up to even
  if (c->use_tma && yvo_single_launch(o)) return launch_yvo_tma(c);
  if (c->use_tma) return launch_yvo_tma_pieces(c);
  // :138  Y[(a,i)][L] = sum_{(b,j)} T[a][b][i][j] Lvo[L][(b,j)]
  Operand A;
  A.base = c->T; A.rows = c->na * o;
  A.rdiv = o; A.r_hi = (i64)v * o * c->jp; A.r_lo = c->jp;
  A.kdiv = o; A.k_hi = (i64)o * c->jp; A.k_lo = 1;
  Operand B = flat_operand(c->Lvo, L, (i64)v * c->jp, 1);
  B.kdiv = o; B.k_hi = c->jp; B.k_lo = 1;
  Output C = flat_output(c->Y + (i64)c->a_begin * o * L, L, 1);
  return launch_gemm<CfgBigKK>(c, "Yvo", A, B, C, c->na * o, L, v * o, 1, nullptr);
}

}  // namespace

extern "C" int mpccb_doubles_begin(mpccb_ctx* c, double* red_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->T) return fail(c, "t2 buffer not bound");
  const int o = c->o, na = c->na;
  const i64 nov = c->nov;
  double* Om = red_dev;   // [a][i] partial sums
  CUDA_OK(c, cudaMemsetAsync(red_dev, 0, sizeof(double) * (nov + 1), c->stream));
  // the slab's W.Joo singles term from mpccb_dress (:126-127) joins the partial sums
  CUDA_OK(c, cudaMemcpyAsync(Om + (i64)c->a_begin * o, c->OmRep + (i64)c->a_begin * o, sizeof(double) * na * o,
                             cudaMemcpyDeviceToDevice, c->stream));
  const bool symm = c->sym.enabled;
  if (symm) {   // tiles that are not launched / rows that are not mirrored must read as zero in the reductions
    CUDA_OK(c, cudaMemsetAsync(c->OmP, 0, sizeof(double) * c->OmP_count, c->stream));
    CUDA_OK(c, cudaMemsetAsync(c->sym.OmPm, 0, sizeof(double) * c->t2_tiles_m * nov, c->stream));
  }
  {
    StageTimer timer(c, ST_ERIS_T2);
    if (launch_eris_t2(c)) return 1;                                 // :135-137 (+ :139 partials)
  }
  // :139  Om[a_loc][i] += sum_tiles OmP;  mirror blocks contribute to rows (b,j) of any slab
  if (reduce_splits(c, Om + (i64)c->a_begin * o, c->OmP, c->t2_tiles_n, (i64)na * o, (i64)na * o, 1.0, 1.0))
    return 1;
  if (symm && reduce_splits(c, Om, c->sym.OmPm, c->t2_tiles_m, nov, nov, 1.0, 1.0)) return 1;
  return 0;
}

extern "C" int mpccb_doubles_end(mpccb_ctx* c, double* red_dev);

extern "C" int mpccb_doubles(mpccb_ctx* c, double* red_dev) {
  if (c->sym.enabled && c->sym.peers)
    return fail(c, "peer-mapped t2: call mpccb_doubles_begin, a cross-rank barrier, then mpccb_doubles_end");
  if (mpccb_doubles_begin(c, red_dev)) return 1;
  return mpccb_doubles_end(c, red_dev);
}

extern "C" int mpccb_doubles_end(mpccb_ctx* c, double* red_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->T) return fail(c, "t2 buffer not bound");
  const int o = c->o, v = c->v, L = c->L, na = c->na;
  const i64 nov = c->nov;
  double* Om = red_dev;
  {
    StageTimer timer(c, ST_YVO);
    if (launch_yvo(c)) return 1;                                     // :138
  }
  StageTimer timer(c, ST_SINGLES_FROM_Y);
  const double* Yslab = c->Y + (i64)c->a_begin * o * L;
  Operand By;   // Y[(b,i)][L] read as rows i, K = (b in slab, L)
  By.base = Yslab; By.rows = o; By.rdiv = BIG; By.r_hi = 0; By.r_lo = L;
  By.kdiv = L; By.k_hi = (i64)o * L; By.k_lo = 1;
  // :141-142 (Lov.t1 part)  Z[j][i] = sum_{(b,L)} Lvo[L][b][j] Y[(b,i)][L]  — its own output buffer, so it runs on the
  // side stream next to the two products that accumulate into Om (all three have small outputs and long K: alone,
  // none of them fills the machine at 8 ranks)
  if (ensure_side(c)) return 1;
  CUDA_OK(c, cudaEventRecord(c->ev_fork, c->stream));
  CUDA_OK(c, cudaStreamWaitEvent(c->side, c->ev_fork, 0));
  {
    cudaStream_t main_stream = c->stream;
    double* main_P = c->P;
    const i64 main_P_count = c->P_count;
    c->stream = c->side; c->P = c->P2; c->P_count = c->P2_count;
    Operand A;
    A.base = c->Lvo + (i64)c->a_begin * c->jp; A.rows = o; A.rdiv = BIG; A.r_hi = 0; A.r_lo = 1;
    A.kdiv = L; A.k_hi = c->jp; A.k_lo = (i64)v * c->jp;
    const int rc = splitk_gemm<CfgN64RK>(c, "Lov.Yvo", A, By, c->Z, o, o, na * L, 1.0, 0.0);
    c->stream = main_stream; c->P = main_P; c->P_count = main_P_count;
    if (rc) return 1;
    CUDA_OK(c, cudaEventRecord(c->ev_join, c->side));
  }
  {  // :143  Om[a][i] -= sum_{(j,L)} Y[(a,j)][L] Joo[L][j][i]          (a in slab)
    Operand A = flat_operand(Yslab, na, (i64)o * L, 1);
    Operand B;
    B.base = c->Joo; B.rows = o; B.rdiv = BIG; B.r_hi = 0; B.r_lo = 1;
    B.kdiv = L; B.k_hi = o; B.k_lo = (i64)o * o;
    if (splitk_gemm<CfgN64KR>(c, "Joo.Yvo", A, B, Om + (i64)c->a_begin * o, na, o, o * L, -1.0, 1.0)) return 1;
  }
  {  // :141-142 (Lvv part)  Om[a][i] += sum_{(b,L)} Lvv[L][b][a] Y[(b,i)][L]     (b in slab, all a)
    Operand A;
    A.base = c->Lvv + (i64)c->a_begin * v; A.rows = v; A.rdiv = BIG; A.r_hi = 0; A.r_lo = 1;
    A.kdiv = L; A.k_hi = v; A.k_lo = (i64)v * v;
    if (splitk_gemm<CfgN64RK>(c, "Lvv.Yvo", A, By, Om, v, o, na * L, 1.0, 1.0)) return 1;
  }
  {  // Om[a][i] += sum_j t1[j][a] Z[j][i]
    CUDA_OK(c, cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    k_add_t1T_Z<<<cdiv(nov, 256), 256, 0, c->stream>>>(Om, c->t1, c->Z, o, v);
    LAUNCH_OK(c, "k_add_t1T_Z");
  }
  {  // :146-147  dE part = sum_{(a,i) in slab, L} e1 * Y
    const int nb = 592;
    k_dot_rows_partial<<<nb, 256, 0, c->stream>>>(c->E1K + (i64)c->a_begin * o * c->ldk, Yslab, c->scal, na * o, L,
                                                  c->ldk, L);
    LAUNCH_OK(c, "k_dot_rows_partial");
    k_final_sums<<<1, 256, 0, c->stream>>>(c->scal, red_dev + nov, nb, 1);
    LAUNCH_OK(c, "k_final_sums");
  }
  return 0;
}

// ===========================================================================
// phase 3: T1 update (reference :128-131, :149-151)
// ===========================================================================
extern "C" int mpccb_finish(mpccb_ctx* c, const double* red_dev, double* t1_new_dev, double* scal_dev) {
  CUDA_OK(c, enter_device(c));
  StageTimer timer(c, ST_FINISH);
  const i64 nov = c->nov;
  const int nb = cdiv(c->v, 32) * c->o;       // residual-norm partials go to the split-K scratch (idle here)
  if ((i64)nb > c->P_count) return fail(c, "residual scratch too small");
  k_t1_update<<<dim3(cdiv(c->v, 32), c->o), 256, 0, c->stream>>>(nullptr, c->w, red_dev, c->t1, c->fvv, c->foo, c->eo,
                                                                 c->ev, t1_new_dev, c->OmTot, c->P, c->o, c->v);
  LAUNCH_OK(c, "k_t1_update");
  k_final_sums<<<1, 256, 0, c->stream>>>(c->P, scal_dev, nb, 1);
  LAUNCH_OK(c, "k_final_sums");
  CUDA_OK(c, cudaMemcpyAsync(scal_dev + 1, red_dev + nov, sizeof(double), cudaMemcpyDeviceToDevice, c->stream));
  return 0;
}

extern "C" int mpccb_updated_amps_host(mpccb_ctx* c, const double* t1_host, double* t1_new_host,
                                       double* res_host, double* dE_host) {
  CUDA_OK(c, enter_device(c));
  const i64 nov = c->nov;
  CUDA_OK(c, cudaMemcpyAsync(c->t1, t1_host, sizeof(double) * nov, cudaMemcpyHostToDevice, c->stream));
  if (c->na != c->v) return fail(c, "mpccb_updated_amps_host works on a whole-range context (one rank)");
  if (mpccb_dress(c, c->t1)) return 1;
  if (mpccb_doubles(c, c->red)) return 1;
  double* sc = c->scal + 8000;
  if (mpccb_finish(c, c->red, c->t1n, sc)) return 1;
  double host_sc[2];
  CUDA_OK(c, cudaMemcpyAsync(t1_new_host, c->t1n, sizeof(double) * nov, cudaMemcpyDeviceToHost, c->stream));
  CUDA_OK(c, cudaMemcpyAsync(host_sc, sc, sizeof(double) * 2, cudaMemcpyDeviceToHost, c->stream));
  CUDA_OK(c, cudaStreamSynchronize(c->stream));
  *res_host = sqrt(host_sc[0]);
  *dE_host = host_sc[1];
  return 0;
}

// ===========================================================================
// amplitude <-> vector layouts, DIIS algebra
// ===========================================================================
extern "C" int mpccb_t2_to_aibj(mpccb_ctx* c, double* out_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->T) return fail(c, "t2 buffer not bound");
  const i64 total = (i64)c->na * c->v * c->o * c->o;
  const int nb = (int)std::min<i64>(cdiv(total, 256), 148 * 16);
  k_abij_to_aibj<<<nb, 256, 0, c->stream>>>(c->T, out_dev, c->o, c->jp, c->v, c->na);
  LAUNCH_OK(c, "k_abij_to_aibj");
  return 0;
}

namespace {
int launch_diis_cols(mpccb_ctx* c, const double* t1_new_dev, const double* xprev_dev,
                     double* x_slot_dev, double* e_slot_dev, const DiisPtrs& ep, int nd, int slot, int first_call,
                     int count_t1, double* scatter_to, int* grid_out) {
  const i64 nblocks = (i64)c->na * c->v;
  const int G = (int)std::min<i64>(nblocks, DIIS_COLS_GRID);
  if ((i64)(G + 1) * 7 > c->P_count) return fail(c, "DIIS partial scratch too small");
  k_diis_push_cols<<<G + 1, 256, 0, c->stream>>>(c->T, t1_new_dev, xprev_dev, x_slot_dev, e_slot_dev, ep, nd, slot,
                                                 first_call, count_t1, c->o, c->jp, c->v, c->a_begin, c->na,
                                                 reinterpret_cast<const i64*>(c->diis_prefix), G, c->P, scatter_to);
  LAUNCH_OK(c, "k_diis_push_cols");
  if (grid_out) *grid_out = G + 1;
  return 0;
}
}  // namespace

extern "C" int mpccb_bind_diis_columns(mpccb_ctx* c, const long long* prefix_dev) {
  c->diis_prefix = prefix_dev;
  return 0;
}

extern "C" int mpccb_pack_tril_rows(mpccb_ctx* c, double* seg_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->T) return fail(c, "t2 buffer not bound");
  if (c->diis_prefix) {                      // column-owned order; seg_dev is the t2 part (vector + nov)
    DiisPtrs ep = {};
    return launch_diis_cols(c, nullptr, nullptr, seg_dev - c->nov, nullptr, ep, 0, 0, 1, 0, nullptr, nullptr);
  }
  k_pack_rows<<<c->na * c->o, 256, 0, c->stream>>>(c->T, seg_dev, c->nov, (i64)c->a_begin * c->o, c->o, c->jp - c->o);
  LAUNCH_OK(c, "k_pack_rows");
  return 0;
}

extern "C" int mpccb_diis_scatter_tril(mpccb_ctx* c, const double* seg_dev, double* full_packed_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->diis_prefix) return fail(c, "the DIIS vector is row-owned: its segments are contiguous, no scatter needed");
  DiisPtrs ep = {};
  return launch_diis_cols(c, nullptr, seg_dev - c->nov, nullptr, nullptr, ep, 0, 0, 0, 0, full_packed_dev, nullptr);
}

extern "C" int mpccb_unpack_oovv(mpccb_ctx* c, const double* packed_dev, double* out_dev) {
  CUDA_OK(c, enter_device(c));
  const i64 total = c->nov * c->nov;
  const int nb = (int)std::min<i64>(cdiv(total, 256), 148 * 16);
  k_unpack_oovv<<<nb, 256, 0, c->stream>>>(packed_dev, out_dev, c->o, c->v);
  LAUNCH_OK(c, "k_unpack_oovv");
  return 0;
}

extern "C" int mpccb_diis_push(mpccb_ctx* c, const double* t1_new_dev, const double* xprev_dev,
                               const double* const* x_prev_list_dev, const double* coeff_prev_host, int n_prev,
                               double* x_slot_dev, double* e_slot_dev, const double* const* e_list_dev,
                               int nd, int slot, int first_call, int count_t1, double* dots_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->T) return fail(c, "t2 buffer not bound");
  if (nd < 0 || nd > 6 || slot < 0 || slot > 5) return fail(c, "bad DIIS slot/nd");
  if (n_prev < 0 || n_prev > 6) return fail(c, "bad DIIS n_prev");
  if (!first_call && n_prev == 0 && !xprev_dev) return fail(c, "DIIS push needs xprev (explicit or as a combination)");
  StageTimer timer(c, ST_DIIS_PUSH);
  DiisPtrs ep;
  for (int k = 0; k < 6; ++k) ep.e[k] = (k < nd && e_list_dev) ? e_list_dev[k] : nullptr;
  DiisPrev pv;
  pv.n = n_prev;
  for (int k = 0; k < 6; ++k) {
    pv.x[k] = k < n_prev ? x_prev_list_dev[k] : nullptr;
    pv.c[k] = k < n_prev ? coeff_prev_host[k] : 0.0;
  }
  int grid = 0;
  if (c->diis_prefix) {
    if (n_prev > 0) return fail(c, "the column-owned DIIS push takes an explicit xprev (n_prev = 0)");
    if (launch_diis_cols(c, t1_new_dev, xprev_dev, x_slot_dev, e_slot_dev, ep, nd, slot, first_call, count_t1,
                         nullptr, &grid)) return 1;
  } else {
    const i64 r0 = (i64)c->a_begin * c->o;
    const int nrows = c->na * c->o;
    const i64 total = ((r0 + nrows) * (r0 + nrows + 1)) / 2 - (r0 * (r0 + 1)) / 2;
    const int G = (int)std::min<i64>(DIIS_ROWS_GRID, std::max<i64>(1, cdiv(total, 2048)));
    const i64 chunk = align_up(cdiv(total, G) > 0 ? (total + G - 1) / G : 1, 256);
    grid = G + 1;
    if ((i64)grid * 7 > c->P_count) return fail(c, "DIIS partial scratch too small");
    k_diis_push<<<grid, 256, 0, c->stream>>>(c->T, t1_new_dev, xprev_dev, pv, x_slot_dev, e_slot_dev, ep, nd, slot,
                                             first_call, count_t1, c->nov, r0, nrows, chunk, G, c->o, c->jp - c->o, c->P);
    LAUNCH_OK(c, "k_diis_push");
  }
  if (!first_call && nd > 0) {
    k_final_sums_strided<<<nd, 256, 0, c->stream>>>(c->P, dots_dev, grid, 7, nd);
    LAUNCH_OK(c, "k_final_sums_strided");
  }
  return 0;
}

// DIIS push for a plain device vector (CCSD-type stage): same arithmetic as mpccb_diis_push without the t2 packing
extern "C" int mpccb_vec_diis_push(mpccb_ctx* c, const double* x_new_dev, long long n, const double* xprev_dev,
                                   const double* const* x_prev_list_dev, const double* coeff_prev_host, int n_prev,
                                   double* x_slot_dev, double* e_slot_dev, const double* const* e_list_dev,
                                   int nd, int slot, int first_call, double* dots_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws) return fail(c, "bind the workspace first");
  if (nd < 0 || nd > 6 || slot < 0 || slot > 5 || n_prev < 0 || n_prev > 6 || n <= 0) return fail(c, "bad DIIS arguments");
  if (!first_call && n_prev == 0 && !xprev_dev) return fail(c, "DIIS push needs xprev (explicit or as a combination)");
  DiisPtrs ep;
  for (int k = 0; k < 6; ++k) ep.e[k] = (k < nd && e_list_dev) ? e_list_dev[k] : nullptr;
  DiisPrev pv;
  pv.n = n_prev;
  for (int k = 0; k < 6; ++k) {
    pv.x[k] = k < n_prev ? x_prev_list_dev[k] : nullptr;
    pv.c[k] = k < n_prev ? coeff_prev_host[k] : 0.0;
  }
  const int grid = (int)std::min<i64>(std::max<i64>(cdiv(n, 2048), 1), std::min<i64>(NSM * 8, c->P_count / 7));
  k_vec_diis_push<<<grid, 256, 0, c->stream>>>(x_new_dev, xprev_dev, pv, x_slot_dev, e_slot_dev, ep, nd, slot,
                                               first_call, n, c->P);
  LAUNCH_OK(c, "k_vec_diis_push");
  if (!first_call && nd > 0) {
    k_final_sums_strided<<<nd, 256, 0, c->stream>>>(c->P, dots_dev, grid, 7, nd);
    LAUNCH_OK(c, "k_final_sums_strided");
  }
  return 0;
}

// lib.pack_tril / lib.unpack_tril(filltriu=SYMMETRIC) of one (n x n) matrix on the device
// (pyscf/lib/numpy_helper.py:256-358; the CCSD-type stage's amplitudes_to_vector / vector_to_amplitudes, ccsd.py:719-735)
extern "C" int mpccb_pack_tril(mpccb_ctx* c, const double* M_dev, long long n, double* packed_dev) {
  CUDA_OK(c, enter_device(c));
  if (n <= 0 || n > INT_MAX) return fail(c, "pack_tril: bad n");
  k_pack_tril<<<(int)n, 256, 0, c->stream>>>(M_dev, packed_dev, n);
  LAUNCH_OK(c, "k_pack_tril");
  return 0;
}

extern "C" int mpccb_unpack_tril(mpccb_ctx* c, const double* packed_dev, long long n, double* full_dev) {
  CUDA_OK(c, enter_device(c));
  if (n <= 0 || n > INT_MAX) return fail(c, "unpack_tril: bad n");
  k_unpack_tril_sym<<<dim3((int)n, 1), 256, 0, c->stream>>>(packed_dev, full_dev, (int)n, n * (n + 1) / 2);
  LAUNCH_OK(c, "k_unpack_tril_sym");
  return 0;
}

extern "C" int mpccb_diis_combine(mpccb_ctx* c, const double* const* x_list_dev, const double* coeff_host,
                                  int nd, double* xnew_dev, long long n) {
  CUDA_OK(c, enter_device(c));
  if (nd < 1 || nd > 6) return fail(c, "bad DIIS nd");
  StageTimer timer(c, ST_DIIS_COMBINE);
  DiisComb cb;
  for (int k = 0; k < 6; ++k) {
    cb.x[k] = k < nd ? x_list_dev[k] : nullptr;
    cb.c[k] = k < nd ? coeff_host[k] : 0.0;
  }
  const int nb = (int)std::min<i64>(cdiv(n, 256), 148 * 16);
  k_diis_combine<<<nb, 256, 0, c->stream>>>(cb, nd, xnew_dev, n);
  LAUNCH_OK(c, "k_diis_combine");
  return 0;
}

// ===========================================================================
// AO -> MO transform of the Cholesky tensor (reference compute_three_center_ints, :46-69)
// ===========================================================================
extern "C" size_t mpccb_ao2mo_scratch_bytes(const mpccb_ctx* c, int nb) {
  const i64 nao = c->o + c->v;
  return (size_t)(2 * (i64)nb * nao * nao * sizeof(double));
}

// packed_dev: nb rows of nao(nao+1)/2 (lower triangle, row-major) = one block of with_df.loop();
// C_dev: mo_coeff (nao x nao, row-major, columns = MOs). Writes Loo/Lov/Lvv[l0 : l0+nb].
//   full = unpack_tril(packed);  half[(l,mu)][q] = sum_nu full[(l,mu)][nu] C[nu][q];
//   Lxy[l][p][q] = sum_mu C[mu][p] half[l][mu][q]   for the oo, ov and vv blocks (the vo block is skipped)
extern "C" int mpccb_ao2mo_block(mpccb_ctx* c, const double* packed_dev, int nb, int l0, const double* C_dev,
                                 double* Loo_dev, double* Lov_dev, double* Lvv_dev, void* scratch_dev,
                                 size_t scratch_bytes) {
  CUDA_OK(c, enter_device(c));
  const int o = c->o, v = c->v, nao = o + v;
  if (nb <= 0 || l0 < 0 || l0 + nb > c->L) return fail(c, "bad auxiliary block [%d,%d)", l0, l0 + nb);
  if (scratch_bytes < mpccb_ao2mo_scratch_bytes(c, nb)) return fail(c, "ao2mo scratch too small");
  if ((i64)nb * nao > (i64)INT_MAX) return fail(c, "auxiliary block too large");
  double* full = static_cast<double*>(scratch_dev);
  double* half = full + (i64)nb * nao * nao;
  k_unpack_tril_sym<<<dim3(nao, nb), 256, 0, c->stream>>>(packed_dev, full, nao, (i64)nao * (nao + 1) / 2);
  LAUNCH_OK(c, "k_unpack_tril_sym");
  {
    Operand A = flat_operand(full, nb * nao, nao, 1);
    Operand B = flat_operand(C_dev, nao, 1, nao);
    Output C = flat_output(half, nao, 1);
    if (launch_gemm<CfgBigKR>(c, "ao2mo.half", A, B, C, nb * nao, nao, nao, 1, nullptr)) return 1;
  }
  struct Blk { int p0, np, q0, nq; double* out; };
  const Blk blks[3] = {{0, o, 0, o, Loo_dev + (i64)l0 * o * o}, {0, o, o, v, Lov_dev + (i64)l0 * o * v},
                       {o, v, o, v, Lvv_dev + (i64)l0 * v * v}};
  for (const Blk& b : blks) {
    Operand A = flat_operand(C_dev + b.p0, b.np, 1, nao);              // (p, mu) -> C[mu][p0 + p]
    Operand B;                                                        // ((l,q), mu) -> half[l][mu][q0 + q]
    B.base = half + b.q0; B.rows = nb * b.nq;
    B.rdiv = b.nq; B.r_hi = (i64)nao * nao; B.r_lo = 1;
    B.kdiv = BIG; B.k_hi = 0; B.k_lo = nao;
    Output C = flat_output(b.out, b.nq, 0);                           // out[l][p][q]
    C.ndiv = b.nq; C.n_hi = (i64)b.np * b.nq; C.n_lo = 1;
    if (b.np > 64) { if (launch_gemm<CfgBigRR>(c, "ao2mo.full", A, B, C, b.np, nb * b.nq, nao, 1, nullptr)) return 1; }
    else           { if (launch_gemm<CfgM64RR>(c, "ao2mo.full", A, B, C, b.np, nb * b.nq, nao, 1, nullptr)) return 1; }
  }
  return 0;
}

// tiles of the fused eris -> t2 kernel: launched under the mirror scheme vs the full grid
extern "C" int mpccb_t2_tiles(const mpccb_ctx* c, long long* launched, long long* total) {
  *total = (long long)c->t2_tiles_m * c->t2_tiles_n;
  *launched = c->sym.enabled ? c->n_tiles : *total;
  return 0;
}

// pure host query of the mirror-scheme rule (no context, no GPU): bit 0 = the rank owning `a` computes block
// (a,b), bit 1 = it also writes the mirror block (b,a), bits 8.. = rank that owns b
extern "C" int mpccb_sym_query(int nranks, int rank, const int* slab_begin, int peers, int a, int b) {
  SymInfo y;
  memset(&y, 0, sizeof(y));
  y.enabled = 1; y.nranks = nranks; y.rank = rank; y.peers = peers;
  for (int r = 0; r <= nranks && r < 17; ++r) y.bounds[r] = slab_begin[r];
  bool mirror; int s;
  const bool comp = sym_compute(y, a, b, mirror, s);
  return (comp ? 1 : 0) | (mirror ? 2 : 0) | (s << 8);
}

// ===========================================================================
// introspection / test hooks
// ===========================================================================
extern "C" int mpccb_get_buffer(mpccb_ctx* c, const char* name, double** ptr, long long* count) {
  const i64 o = c->o, v = c->v, L = c->L, nov = c->nov;
  struct Ent { const char* n; double* p; i64 cnt; };
  const Ent tab[] = {
      {"Lvo", c->Lvo, L * v * c->jp},   {"t1", c->t1, nov},          {"t1n", c->t1n, nov},
      {"Xov", c->Xov, L * o * v},   {"Xoo", c->Xoo, L * o * o},   {"Joo", c->Joo, L * o * o},
      {"X", c->X, L},               {"XT", c->XT, L * o * v},     {"JT", c->JT, L * o * v},
      {"JK", c->JK, nov * c->ldk},  {"E1K", c->E1K, nov * c->ldk}, {"Wov", c->Wov, L * o * v},
      {"w", c->w, nov},             {"Fvo", c->Fvo, nov},         {"Y", c->Y, nov * L},
      {"OmRep", c->OmRep, nov},     {"OmTot", c->OmTot, nov},     {"Z", c->Z, o * o},
      {"red", c->red, nov + 1},     {"T", c->T, (i64)c->na * v * o * c->jp},
  };
  for (const Ent& e : tab)
    if (!strcmp(e.n, name)) { *ptr = e.p; *count = e.cnt; return 0; }
  return fail(c, "unknown buffer %s", name);
}

extern "C" int mpccb_copy_to_host(mpccb_ctx* c, const double* src, double* dst, long long count) {
  CUDA_OK(c, enter_device(c));
  CUDA_OK(c, cudaMemcpyAsync(dst, src, sizeof(double) * count, cudaMemcpyDeviceToHost, c->stream));
  CUDA_OK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

extern "C" int mpccb_copy_to_device(mpccb_ctx* c, const double* src, double* dst, long long count) {
  CUDA_OK(c, enter_device(c));
  CUDA_OK(c, cudaMemcpyAsync(dst, src, sizeof(double) * count, cudaMemcpyHostToDevice, c->stream));
  CUDA_OK(c, cudaStreamSynchronize(c->stream));
  return 0;
}

// C[m][n] = sum_k A[m][k] B[n][k], all row-major contiguous. Large products whose rows are 16-byte aligned go
// through the TMA-fed kernel (same code as the Yvo GEMM), everything else through the generic DMMA kernel.
extern "C" int mpccb_gemm_nt(mpccb_ctx* c, const double* A_dev, const double* B_dev, double* C_dev,
                             int M, int N, int K) {
  CUDA_OK(c, enter_device(c));
  if (M <= 0 || N <= 0 || K <= 0) return fail(c, "gemm_nt: empty product");
  const bool aligned = (K % 2 == 0) && ((reinterpret_cast<uintptr_t>(A_dev) | reinterpret_cast<uintptr_t>(B_dev)) & 15) == 0;
  if (c->use_tma && aligned && K >= 16 && (double)M * N * K >= 5e8 && M >= 64 && N >= 64) {
    constexpr int TN = TmaCfg<4>::TN;
    CUtensorMap mA, mB;
    if (make_map_2d(c, &mA, A_dev, K, M, K, TMA_TM)) return 1;
    if (make_map_2d(c, &mB, B_dev, K, N, K, TN)) return 1;
    TmaGemmParams p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N;
    p.tm_eff = TMA_TM; p.tn_eff = TN;
    p.tiles_m = cdiv(M, TMA_TM); p.tiles_n = cdiv(N, TN);
    p.nk = cdiv(K, TMA_TK);
    p.a_mode = 0;
    p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = TN * TMA_TK * 8;
    p.C = C_dev; p.ldc = N; p.ksplit = 1;
    // fewer tiles than two waves and a long contraction index: K slices (fixed-order reduction afterwards), the
    // factor that minimises ceil(T * s / 148) / s
    const i64 T = (i64)p.tiles_m * p.tiles_n;
    int best_s = 1;
    if (c->P && T < 2 * NSM) {
      double best = (double)cdiv(T, NSM);
      const i64 cap = c->P_count / ((i64)M * N);
      for (int sp = 2; sp <= 64; ++sp) {
        if (p.nk / sp < 32 || sp > cap) break;
        const double w = (double)cdiv(T * sp, NSM) / sp + 0.002 * sp;
        if (w < best * 0.97) { best = w; best_s = sp; }
      }
    }
    T2Params t;
    memset(&t, 0, sizeof(t));
    if (best_s > 1) {
      p.nk_split = cdiv(p.nk, best_s);
      p.ksplit = cdiv(p.nk, p.nk_split);
      p.C = c->P; p.split_stride = (i64)M * N;
      p.tiles_total = (int)T * p.ksplit;
      if (launch_tma<0, 4>(c, "gemm_tma_kernel<nt split-K>", mA, mB, p, t)) return 1;
      return reduce_splits(c, C_dev, c->P, p.ksplit, (i64)M * N, (i64)M * N, 1.0, 0.0);
    }
    return launch_tma<0, 4>(c, "gemm_tma_kernel<nt>", mA, mB, p, t);
  }
  Operand A = flat_operand(A_dev, M, K, 1);
  Operand B = flat_operand(B_dev, N, K, 1);
  Output C = flat_output(C_dev, N, 1);
  // few output tiles but a long contraction index (the "->ia", "->ki", "->ac" terms: K = o*v*v): split K over
  // the machine into the context's partial buffer, then a fixed-order reduction
  const int tiles = cdiv(M, 128) * cdiv(N, N <= 64 ? 64 : 128);
  if (c->P && tiles * 2 <= NSM && K >= 2048) {
    int splits = std::min(cdiv(2 * NSM, tiles), K / 256);
    const i64 cap = c->P_count / ((i64)M * N);
    if (splits > cap) splits = (int)cap;
    if (splits >= 2) {
      const int kchunk = (int)align_up(cdiv(K, splits), TK);
      splits = cdiv(K, kchunk);
      int rc = (N <= 64) ? launch_gemm<CfgN64KK>(c, "gemm_nt(split-K)", A, B, C, M, N, K, splits, c->P)
                         : launch_gemm<CfgBigKK>(c, "gemm_nt(split-K)", A, B, C, M, N, K, splits, c->P);
      if (rc) return rc;
      return reduce_splits(c, C_dev, c->P, splits, (i64)M * N, (i64)M * N, 1.0, 0.0);
    }
  }
  if (N <= 64) return launch_gemm<CfgN64KK>(c, "gemm_nt", A, B, C, M, N, K, 1, nullptr);
  return launch_gemm<CfgBigKK>(c, "gemm_nt", A, B, C, M, N, K, 1, nullptr);
}

// y = alpha * op(A) x + beta * y for the contractions in which one operand is summed away entirely ("kdac,kd->ac":
// matrix-vector products, bound by one read of the matrix): trans = 0: A is [rows][K] (one warp per row),
// trans = 1: A is [K][rows] (lanes along rows, K sliced over the grid, fixed-order reduction).
extern "C" int mpccb_gemv(mpccb_ctx* c, const double* A_dev, const double* x_dev, double* y_dev, long long rows,
                          long long K, int trans, double alpha, double beta) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws) return fail(c, "bind the workspace first");
  if (rows <= 0 || K <= 0 || rows > INT_MAX || K > INT_MAX) return fail(c, "gemv: bad sizes");
  if (!trans) {
    if (rows > c->P_count) return fail(c, "gemv: scratch too small");
    k_gemv_rows<<<cdiv(rows, 8), 256, 0, c->stream>>>(A_dev, x_dev, c->P, (int)rows, (int)K, K);
    LAUNCH_OK(c, "k_gemv_rows");
    return reduce_splits(c, y_dev, c->P, 1, rows, rows, alpha, beta);
  }
  const int nbx = cdiv(rows, 256);
  int slices = (int)std::min<i64>(std::max<i64>(1, (4 * NSM) / nbx), std::max<i64>(1, K / 64));
  slices = (int)std::min<i64>(slices, c->P_count / rows);
  if (slices < 1) return fail(c, "gemv: scratch too small");
  const i64 kchunk = (K + slices - 1) / slices;
  slices = (int)((K + kchunk - 1) / kchunk);
  k_gemv_cols_partial<<<dim3(nbx, slices), 256, 0, c->stream>>>(A_dev, x_dev, c->P, K, rows, kchunk);
  LAUNCH_OK(c, "k_gemv_cols_partial");
  return reduce_splits(c, y_dev, c->P, slices, rows, rows, alpha, beta);
}

// C[map(m, n)] = alpha * sum_k A[m][k] B[n][k] + beta * C[map(m, n)] with the two-level output map
//   map(m, n) = (m / mdiv) * m_hi + (m % mdiv) * m_lo + (n / ndiv) * n_hi + (n % ndiv) * n_lo
// so that a contraction lands in its (permuted) result tensor without a separate permute / accumulate pass
// (lib.einsum's transpose -> dot -> transpose, pyscf/lib/numpy_helper.py:101-252, minus the last transpose).
// The map must be injective on [0,M) x [0,N) (the engine derives it from a permutation of the output indices).
extern "C" int mpccb_gemm_nt_mapped(mpccb_ctx* c, const double* A_dev, const double* B_dev, double* C_dev,
                                    int M, int N, int K, int mdiv, long long m_hi, long long m_lo,
                                    int ndiv, long long n_hi, long long n_lo, double alpha, double beta) {
  CUDA_OK(c, enter_device(c));
  if (M <= 0 || N <= 0 || K <= 0 || mdiv <= 0 || ndiv <= 0) return fail(c, "gemm_nt_mapped: bad sizes");
  Output C = flat_output(C_dev, 0, 0);
  C.mdiv = mdiv; C.m_hi = m_hi; C.m_lo = m_lo;
  C.ndiv = ndiv; C.n_hi = n_hi; C.n_lo = n_lo;
  C.alpha = alpha; C.beta = beta;
  const bool aligned = (K % 2 == 0) && ((reinterpret_cast<uintptr_t>(A_dev) | reinterpret_cast<uintptr_t>(B_dev)) & 15) == 0;
  if (c->use_tma && aligned && K >= 16 && (double)M * N * K >= 5e8 && M >= 64 && N >= 64) {
    // short K (the W = Leff.Leff^T product of the DF-direct ladder, K = naux): the store epilogue is a visible share of
    // a tile's life, so 128 x 64 tiles with two CTAs per SM — one stores while the other multiplies
    const bool narrow = K <= c->mapped_narrow_k;
    const int TN = narrow ? TmaCfg<2>::TN : TmaCfg<4>::TN;
    CUtensorMap mA, mB;
    if (make_map_2d(c, &mA, A_dev, K, M, K, TMA_TM)) return 1;
    if (make_map_2d(c, &mB, B_dev, K, N, K, TN)) return 1;
    TmaGemmParams p;
    memset(&p, 0, sizeof(p));
    p.M = M; p.N = N;
    p.tm_eff = TMA_TM; p.tn_eff = TN;
    p.tiles_m = cdiv(M, TMA_TM); p.tiles_n = cdiv(N, TN);
    p.nk = cdiv(K, TMA_TK);
    p.a_mode = 0;
    p.a_bytes = TMA_TM * TMA_TK * 8; p.b_bytes = (uint32_t)(TN * TMA_TK * 8);
    p.ksplit = 1;
    p.out = C;
    T2Params t;
    memset(&t, 0, sizeof(t));
    return narrow ? launch_tma<2, 2>(c, "gemm_tma_kernel<mapped,narrow>", mA, mB, p, t)
                  : launch_tma<2, 4>(c, "gemm_tma_kernel<mapped>", mA, mB, p, t);
  }
  Operand A = flat_operand(A_dev, M, K, 1);
  Operand B = flat_operand(B_dev, N, K, 1);
  if (N <= 64) return launch_gemm<CfgN64KK>(c, "gemm_nt_mapped", A, B, C, M, N, K, 1, nullptr);
  return launch_gemm<CfgBigKK>(c, "gemm_nt_mapped", A, B, C, M, N, K, 1, nullptr);
}

// ===========================================================================
// tensor engine (rows f1/f2): permute/accumulate, denominators, active-space overwrite, dot
// ===========================================================================
// out = alpha * transpose(in, perm) + beta * out   (numpy semantics: out axis d = in axis perm[d])
extern "C" int mpccb_permute_axpby(mpccb_ctx* c, const double* in_dev, double* out_dev, int rank,
                                   const int* in_dims, const int* perm, double alpha, double beta) {
  CUDA_OK(c, enter_device(c));
  if (rank < 1 || rank > MAX_RANK) return fail(c, "permute: rank must be 1..%d", MAX_RANK);
  i64 istr[MAX_RANK], total = 1;
  int seen = 0;
  for (int d = rank - 1; d >= 0; --d) { istr[d] = total; total *= in_dims[d]; }
  for (int d = 0; d < rank; ++d) {
    if (perm[d] < 0 || perm[d] >= rank || (seen >> perm[d]) & 1) return fail(c, "permute: bad permutation");
    seen |= 1 << perm[d];
  }
  if (total == 0) return 0;
  int rpos = -1;                                 // output axis that is the input's unit-stride axis
  for (int d = 0; d < rank; ++d) if (perm[d] == rank - 1) rpos = d;
  const int nblk_max = NSM * 8;
  if (rpos == rank - 1 || in_dims[rank - 1] < 8 || in_dims[perm[rank - 1]] < 8) {
    PermuteDesc d;
    memset(&d, 0, sizeof(d));
    d.rank = rank; d.total = total;
    for (int k = 0; k < rank; ++k) { d.out_dim[k] = in_dims[perm[k]]; d.in_stride[k] = istr[perm[k]]; }
    const int nb = (int)std::min<i64>(cdiv(total, 256), nblk_max);
    k_permute_axpby<<<nb, 256, 0, c->stream>>>(in_dev, out_dev, d, alpha, beta);
    LAUNCH_OK(c, "k_permute_axpby");
    return 0;
  }
  PermuteTiled d;
  memset(&d, 0, sizeof(d));
  i64 ostr[MAX_RANK], acc = 1;
  for (int k = rank - 1; k >= 0; --k) { ostr[k] = acc; acc *= in_dims[perm[k]]; }
  d.R = in_dims[rank - 1]; d.C = in_dims[perm[rank - 1]];
  d.r_out_stride = ostr[rpos]; d.c_in_stride = istr[perm[rank - 1]];
  d.n_outer = 1;
  for (int k = 0; k < rank - 1; ++k) {
    if (k == rpos) continue;
    d.outer_dim[d.rank] = in_dims[perm[k]];
    d.outer_in_stride[d.rank] = istr[perm[k]];
    d.outer_out_stride[d.rank] = ostr[k];
    d.n_outer *= in_dims[perm[k]];
    d.rank++;
  }
  const i64 tiles = d.n_outer * cdiv(d.R, 32) * cdiv(d.C, 32);
  const int nb = (int)std::min<i64>(tiles, nblk_max);
  k_permute_tiled<<<nb, 256, 0, c->stream>>>(in_dev, out_dev, d, alpha, beta);
  LAUNCH_OK(c, "k_permute_tiled");
  return 0;
}

// rank 2: t1[i,a] /= eo[i]-ev[a];  rank 4: t2[i,j,a,b] /= eo[i]+eo[j]-ev[a]-ev[b]   (the context's eps_o / eps_v)
extern "C" int mpccb_div_denominators(mpccb_ctx* c, double* t_dev, int rank) {
  CUDA_OK(c, enter_device(c));
  if (!c->eo || !c->ev) return fail(c, "orbital energies not bound");
  if (rank != 2 && rank != 4) return fail(c, "div_denominators: rank 2 or 4");
  const i64 total = rank == 2 ? c->nov : c->nov * c->nov;
  const int nb = (int)std::min<i64>(cdiv(total, 256), NSM * 16);
  k_div_denominators<<<nb, 256, 0, c->stream>>>(t_dev, c->eo, c->ev, c->o, c->v, rank);
  LAUNCH_OK(c, "k_div_denominators");
  return 0;
}

// t[x0..] /= sum_d sign[d] * e_dev[d][x_d]  (rank <= 4; e_dev[d] has dims[d] entries)
extern "C" int mpccb_div_general(mpccb_ctx* c, double* t_dev, int rank, const int* dims,
                                 const double* const* e_dev, const double* signs) {
  CUDA_OK(c, enter_device(c));
  if (rank < 1 || rank > 4) return fail(c, "div_general: rank 1..4");
  DenomDesc d;
  memset(&d, 0, sizeof(d));
  d.rank = rank; d.total = 1;
  for (int k = 0; k < rank; ++k) { d.dim[k] = dims[k]; d.e[k] = e_dev[k]; d.sign[k] = signs[k]; d.total *= dims[k]; }
  if (d.total == 0) return 0;
  const int nb = (int)std::min<i64>(cdiv(d.total, 256), NSM * 16);
  k_div_general<<<nb, 256, 0, c->stream>>>(t_dev, d);
  LAUNCH_OK(c, "k_div_general");
  return 0;
}

// tnew[idx] = told[idx] for the active/inactive index classes in class_mask (see k_class_overwrite)
extern "C" int mpccb_class_overwrite(mpccb_ctx* c, double* tnew_dev, const double* told_dev, int rank,
                                     const int* act_h_dev, const int* act_p_dev, unsigned class_mask) {
  CUDA_OK(c, enter_device(c));
  if (rank != 2 && rank != 4) return fail(c, "class_overwrite: rank 2 or 4");
  const i64 total = rank == 2 ? c->nov : c->nov * c->nov;
  const int nb = (int)std::min<i64>(cdiv(total, 256), NSM * 16);
  k_class_overwrite<<<nb, 256, 0, c->stream>>>(tnew_dev, told_dev, act_h_dev, act_p_dev, class_mask, c->o, c->v, rank);
  LAUNCH_OK(c, "k_class_overwrite");
  return 0;
}

// Test hook: q[i] = fast_div(x[i], y[i]) — the branch-free quotient the fused t2 epilogue uses instead of the
// compiler's IEEE division sequence (gemm_tma.cuh). tests/test_gpu_parity.py sweeps it against x / y.
namespace {
__global__ void __launch_bounds__(256) k_test_fast_div(const double* __restrict__ x, const double* __restrict__ y,
                                                       double* __restrict__ q, i64 n) {
  const i64 stride = (i64)gridDim.x * blockDim.x * 8;
  for (i64 i0 = ((i64)blockIdx.x * blockDim.x + threadIdx.x) * 8; i0 < n; i0 += stride) {
    double xs[8], ys[8], qs[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) { xs[u] = i0 + u < n ? x[i0 + u] : 1.0; ys[u] = i0 + u < n ? y[i0 + u] : 1.0; }
    fast_div_batch<8>(xs, ys, qs);                    // the batched form the epilogue calls
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (i0 + u < n) q[i0 + u] = (u & 1) ? qs[u] : fast_div(xs[u], ys[u]);   // and the scalar form, alternating
  }
}
}  // namespace

extern "C" int mpccb_test_fast_div(mpccb_ctx* c, const double* x_dev, const double* y_dev, double* q_dev, long long n) {
  CUDA_OK(c, enter_device(c));
  if (n <= 0) return 0;
  const int nb = (int)std::min<i64>(cdiv(n, 2048), NSM * 8);
  k_test_fast_div<<<nb, 256, 0, c->stream>>>(x_dev, y_dev, q_dev, n);
  LAUNCH_OK(c, "k_test_fast_div");
  return 0;
}

// lib.take_2d / lib.takebak_2d on device buffers (pyscf/lib/numpy_helper.py:478-541): the gather / scatter-add behind
// ccsd.amplitudes_to_vector_s4 / _unpack_4fold (pyscf/cc/ccsd.py:705-757), which the unrestricted prototype's run_diis
// uses (pyscf/cc/umpcc_lowlevel.py:15-46). idx / idy: int32 device arrays of row / column numbers.
extern "C" int mpccb_take_2d(mpccb_ctx* c, const double* a_dev, long long lda, const int* idx_dev, int nx,
                             const int* idy_dev, int ny, double* out_dev) {
  CUDA_OK(c, enter_device(c));
  if (nx < 0 || ny < 0 || lda < 0) return fail(c, "take_2d: negative size");
  if (nx == 0 || ny == 0) return 0;
  const int nb = (int)std::min<i64>(cdiv((i64)nx * ny, 256), NSM * 8);
  k_take_2d<<<nb, 256, 0, c->stream>>>(a_dev, lda, idx_dev, nx, idy_dev, ny, out_dev);
  LAUNCH_OK(c, "k_take_2d");
  return 0;
}

extern "C" int mpccb_takebak_2d(mpccb_ctx* c, double* out_dev, long long ldo, const double* a_dev, const int* idx_dev,
                                int nx, const int* idy_dev, int ny, double alpha) {
  CUDA_OK(c, enter_device(c));
  if (nx < 0 || ny < 0 || ldo < 0) return fail(c, "takebak_2d: negative size");
  if (nx == 0 || ny == 0) return 0;
  const int nb = (int)std::min<i64>(cdiv((i64)nx * ny, 256), NSM * 8);
  k_takebak_2d<<<nb, 256, 0, c->stream>>>(out_dev, ldo, a_dev, idx_dev, nx, idy_dev, ny, alpha);
  LAUNCH_OK(c, "k_takebak_2d");
  return 0;
}

// out_dev[0] = sum_i x[i] y[i]  (deterministic two-pass reduction; needs the bound workspace)
extern "C" int mpccb_dot(mpccb_ctx* c, const double* x_dev, const double* y_dev, long long n, double* out_dev) {
  CUDA_OK(c, enter_device(c));
  if (!c->ws) return fail(c, "bind the workspace first");
  const int nb = (int)std::min<i64>(std::max<i64>(cdiv(n, 1024), 1), 592);
  k_dot_partial<<<nb, 256, 0, c->stream>>>(x_dev, y_dev, c->scal, n);
  LAUNCH_OK(c, "k_dot_partial");
  k_final_sums<<<1, 256, 0, c->stream>>>(c->scal, out_dev, nb, 1);
  LAUNCH_OK(c, "k_final_sums");
  return 0;
}

// ---------------------------------------------------------------------------
// mpccb_dot without any allocation (no counterpart in the reference; added for
// callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid it).
// Same two kernels and the same block count as mpccb_dot — hence the same bits — but the per-block partial sums go to a
// file-scope device array instead of the caller-allocated workspace, so no workspace needs to be bound.
//   * MPCCB_DOT_STATIC_PARTIALS doubles (8 KB) per GPU that loads this library, held until the process exits.
//   * `g_dot_static` is a HOST-SIDE STAND-IN in host code, not a device address: its device address is looked up once
//     per device, with that device current, by cudaGetSymbolAddress (dot_static_address) and that pointer is passed
//     to the kernels.
//   * Every call on a device shares the array. Nothing in it has to start at zero (each of the nb partials is written
//     before the second kernel reads it), so it is not cleared; it keeps the last call's partials. Calls that overlap
//     on the GPU (two streams, two contexts on one device) race on it: serialise them.
// ---------------------------------------------------------------------------
namespace {
constexpr int MPCCB_DOT_STATIC_PARTIALS = 1024;                   // mpccb_dot uses at most 592 blocks
constexpr long long MPCCB_DOT_STATIC_MAX_N = 1LL << 32;           // stated maximum length (34 GB per vector)
constexpr int MPCCB_MAX_DEVICES = 64;
static __device__ __align__(256) double g_dot_static[MPCCB_DOT_STATIC_PARTIALS];
double* g_dot_static_addr[MPCCB_MAX_DEVICES] = {};                // device address of g_dot_static, per device

// caller has made `device` current
int dot_static_address(mpccb_ctx* c, int device, double** out) {
  if (device < 0 || device >= MPCCB_MAX_DEVICES) return fail(c, "device %d outside the supported 0..%d", device, MPCCB_MAX_DEVICES - 1);
  if (!g_dot_static_addr[device]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_dot_static);
    if (e != cudaSuccess || !p) return fail(c, "cudaGetSymbolAddress(g_dot_static) failed: %s", cudaGetErrorString(e));
    g_dot_static_addr[device] = static_cast<double*>(p);
  }
  *out = g_dot_static_addr[device];
  return 0;
}
}  // namespace

extern "C" int mpccb_dot_static(mpccb_ctx* c, const double* x_dev, const double* y_dev, long long n, double* out_dev) {
  CUDA_OK(c, enter_device(c));
  if (!x_dev || !y_dev || !out_dev) return fail(c, "mpccb_dot_static: null pointer");
  if (n <= 0) return fail(c, "mpccb_dot_static: n = %lld must be positive", n);
  if (n > MPCCB_DOT_STATIC_MAX_N)
    return fail(c, "mpccb_dot_static: n = %lld exceeds the maximum of %lld elements of the allocation-free form "
                   "(use mpccb_dot with a bound workspace)", n, MPCCB_DOT_STATIC_MAX_N);
  const int nb = (int)std::min<i64>(std::max<i64>(cdiv(n, 1024), 1), 592);      // the rule of mpccb_dot
  static_assert(592 <= MPCCB_DOT_STATIC_PARTIALS, "static scratch smaller than the block count");
  double* partial = nullptr;
  if (dot_static_address(c, c->device, &partial)) return 1;
  k_dot_partial<<<nb, 256, 0, c->stream>>>(x_dev, y_dev, partial, n);
  LAUNCH_OK(c, "k_dot_partial(static)");
  k_final_sums<<<1, 256, 0, c->stream>>>(partial, out_dev, nb, 1);
  LAUNCH_OK(c, "k_final_sums(static)");
  return 0;
}
