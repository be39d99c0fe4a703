/* libmpcc_b200 — C ABI of the B200 (sm_100a) DF-MPCC low-level iteration.
 *
 * Drop-in boundary for the hot path of FabianFaulstich/pyscf_mpcc
 * (pyscf/mpcc/dfrmpcc_lowlevel.py::MPCC_LL). The reference has no native code
 * on this path besides lib.pack_tril/unpack_tril (ctypes -> NPdpack_tril_2d /
 * NPdunpack_tril_2d, pyscf/lib/numpy_helper.py:286-289,355-357); it follows the
 * same calling convention — plain pointers and integer sizes, the caller owns
 * every tensor — with two deliberate differences: every entry point returns an
 * int status (0 = ok; CUDA can fail) and an opaque context owns launch state.
 *
 * Pointer arguments named *_dev are device pointers (the Python host owns them
 * through torch tensors); *_host are host pointers. All arrays are C-contiguous
 * float64. Calls are asynchronous on the context's stream unless stated.
 *
 * Index names: o = nocc, v = nvir, L = naux, nov = o*v. A context works on the
 * virtual-index slab a in [a_begin, a_end) (whole range on one GPU).
 */
#ifndef MPCC_B200_H
#define MPCC_B200_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mpccb_ctx mpccb_ctx;

/* ---- lifetime -------------------------------------------------------------------- */
/* replaces the size bookkeeping of MPCC_LL.__init__ (dfrmpcc_lowlevel.py:23-31) */
int mpccb_create(mpccb_ctx** out, int nocc, int nvir, int naux, int a_begin, int a_end, int device);
void mpccb_destroy(mpccb_ctx* ctx);
const char* mpccb_last_error(const mpccb_ctx* ctx);
const char* mpccb_version(void);
int mpccb_set_stream(mpccb_ctx* ctx, void* cuda_stream);
int mpccb_synchronize(mpccb_ctx* ctx);

/* scratch: the caller allocates `mpccb_workspace_bytes` bytes of device memory and binds it */
size_t mpccb_workspace_bytes(const mpccb_ctx* ctx);
int mpccb_bind_workspace(mpccb_ctx* ctx, void* ws_dev, size_t bytes);

/* MO-basis three-index tensors and Fock blocks, the attributes MPCC_LL keeps after
 * compute_three_center_ints (dfrmpcc_lowlevel.py:46-69, :35-42):
 * Loo (L,o,o), Lov (L,o,v), Lvv (L,v,v), foo (o,o), fvv (v,v), eps_o (o), eps_v (v).
 * Builds the (L,v,o) transpose of Lov in the workspace (the reference's unused `Lvo`). */
int mpccb_bind_integrals(mpccb_ctx* ctx, const double* Loo_dev, const double* Lov_dev,
                         const double* Lvv_dev, const double* foo_dev, const double* fvv_dev,
                         const double* eps_o_dev, const double* eps_v_dev);
/* AO -> MO transform of one block of Cholesky rows on the device: MPCC_LL.compute_three_center_ints
 * (dfrmpcc_lowlevel.py:46-69: lib.unpack_tril + lib.einsum("Lab,ap,bq->Lpq") + slicing). packed_dev holds nb rows
 * of nao(nao+1)/2 doubles as with_df.loop() yields them; C_dev is mo_coeff (nao x nao row-major). Writes rows
 * [l0, l0+nb) of Loo (L,o,o), Lov (L,o,v), Lvv (L,v,v). scratch_dev: mpccb_ao2mo_scratch_bytes(ctx, nb) bytes. */
size_t mpccb_ao2mo_scratch_bytes(const mpccb_ctx* ctx, int nb);
int mpccb_ao2mo_block(mpccb_ctx* ctx, const double* packed_dev, int nb, int l0, const double* C_dev,
                      double* Loo_dev, double* Lov_dev, double* Lvv_dev, void* scratch_dev, size_t scratch_bytes);
/* t2 buffer in [a - a_begin, b, i, j] order: (a_end-a_begin)*v*o*o doubles */
int mpccb_bind_t2(mpccb_ctx* ctx, double* t2_abij_dev);

/* Several ranks, fused all-gather of the dressed integrals: every rank allocates its JK buffer
 * (nov * ldk doubles, ldk = naux rounded up to even, zero-filled) in peer-mapped symmetric memory and binds
 * (a) its own buffer, (b) either the device pointers of all ranks' buffers (npeers of them, own rank included)
 * or the NVSwitch multicast address of the allocation. mpccb_dress then stores the slab's JK rows directly into
 * every rank's buffer over NVLink (no separate all-gather); the caller only has to run a cross-rank barrier on
 * the stream before mpccb_dress_post. Optional: without it the caller all-gathers the "JK" buffer itself. */
int mpccb_bind_jk_exchange(mpccb_ctx* ctx, double* jk_local_dev, double* const* peer_jk_dev, int npeers,
                           double* jk_multicast_dev);

/* Subset form of the same exchange: with the mirror scheme (below) rank r's JK rows are needed by the nranks/2 + 1 ranks
 * r, r-1, .., r-nranks/2 (mod nranks) only — one NVLink store per needed peer. The symmetric allocation holds
 * nov*ldk (JK) + nov (w = Jvo.X, :128) + nov (Fvo = w - F2, :133) doubles; mpccb_dress forms the slab's rows of w and
 * of Fvo (F2 is then a GEMM over the slab's rows only) and broadcasts them (through w_multicast_dev when given, else
 * through the peer pointers); mpccb_dress_post has nothing left to do.
 * peer_jk_dev: the nranks buffers in rank order as mapped on this device. */
int mpccb_bind_jk_subset(mpccb_ctx* ctx, double* jk_local_dev, double* const* peer_jk_dev, int nranks, int rank,
                         double* w_multicast_dev);
/* the rule behind it, as a pure host function (no device): 1 if rank dst needs the JK rows of rank src's slab */
int mpccb_jk_subset_needs(int nranks, int src, int dst);

/* Xoo = Lov.t1^T (:120) and Joo = Xoo + Loo (:123) do not depend on the virtual slab; replicated they cost every rank
 * the same GEMM however many ranks there are. Sharded form: the ranks allocate [Xoo | Joo] (2 * naux * nocc^2 doubles)
 * in symmetric memory; rank r computes the rows of its slab of the AUXILIARY index, l in [naux r / P, naux (r+1) / P),
 * and stores its Xoo rows into every rank's buffer (through xoo_multicast_dev when given, else one store per peer
 * pointer; Joo is completed locally from the exchanged Xoo).
 * mpccb_dress then returns once the exchange and the slab's Xvo GEMM (:119) are issued; the caller runs a cross-rank
 * barrier on the stream and calls mpccb_dress_rest. xoo_local_dev == NULL goes back to the replicated form. */
int mpccb_bind_xoo_exchange(mpccb_ctx* ctx, double* xoo_local_dev, double* const* peer_xoo_dev, int nranks, int rank,
                            double* xoo_multicast_dev);

/* Mirror scheme of the eris -> t2 step: e[(a,i),(b,j)] is a Gram matrix, so t2[a,b,i,j] = t2[b,a,j,i]; only one
 * block of each pair is computed and its CTA also writes the mirror block (always inside one slab; option "t2_sym" 0
 * turns it off). With several ranks and peer-mapped t2 buffers (symmetric memory) this call extends it across slabs:
 * rank `rank` of `nranks` owns a in [slab_begin[rank], slab_begin[rank+1]); T_peers_dev[r] = rank r's t2 buffer as
 * mapped on this device (own buffer included; replaces mpccb_bind_t2). */
int mpccb_bind_t2_peers(mpccb_ctx* ctx, int nranks, int rank, const int* slab_begin, double* const* T_peers_dev);

/* ---- one residual evaluation: MPCC_LL.updated_amps (dfrmpcc_lowlevel.py:103-152) --- */
/* phase 1, :119-133 — T1 dressing and singles intermediates. Everything that carries a virtual index is
 * computed for the context's slab only (rows [a_begin*o, a_end*o) of JK = Jvo in [(a,i)][ldk] order,
 * ldk = naux rounded up to even). A whole-range context (one rank) also runs mpccb_dress_post itself. */
int mpccb_dress(mpccb_ctx* ctx, const double* t1_dev);
/* phase 1b — several ranks: after the caller has all-gathered the JK rows of all slabs (buffer "JK" of
 * mpccb_get_buffer; slabs are contiguous row ranges), form w = Jvo.X and Fov for every (b,j) (:128,:133) */
int mpccb_dress_post(mpccb_ctx* ctx);
/* second half of mpccb_dress (everything that reads Xoo / Joo for every auxiliary index); mpccb_dress runs it itself
 * unless the Xoo exchange is bound */
int mpccb_dress_rest(mpccb_ctx* ctx);
/* phase 2, :135-147 — eris/t2 (fused), Yvo, doubles->singles, energy, for the local slab.
 * Writes the slab's partial sums to red_dev[0..nov) = Omega part [a,i], red_dev[nov] = dE part;
 * with several ranks the caller all-reduces red_dev (nov+1 doubles) before phase 3. */
int mpccb_doubles(mpccb_ctx* ctx, double* red_dev);
/* the same in two halves: begin = eris/t2 (with the mirror scheme of mpccb_bind_t2_peers the kernel also stores
 * blocks into OTHER ranks' t2 buffers), end = Yvo, doubles->singles, energy. With peer-mapped t2 buffers the caller
 * runs a cross-rank barrier on the stream between the two. */
int mpccb_doubles_begin(mpccb_ctx* ctx, double* red_dev);
int mpccb_doubles_end(mpccb_ctx* ctx, double* red_dev);
/* phase 3, :128-131,:149-151 — assemble Omega, t1_new = -Omega^T/eia, scal_dev = {res^2, dE} */
int mpccb_finish(mpccb_ctx* ctx, const double* red_dev, double* t1_new_dev, double* scal_dev);
/* all three phases with HOST buffers and a final synchronize (single rank):
 * t1 (o,v) in, t1_new (o,v), res and dE out — the reference call with numpy arrays. */
int mpccb_updated_amps_host(mpccb_ctx* ctx, const double* t1_host, double* t1_new_host,
                            double* res_host, double* dE_host);

/* ---- amplitude <-> vector layouts (dfrmpcc_lowlevel.py:162-176; NPdpack_tril /
 *      NPdunpack_tril + NPdsymm_triu, pyscf/lib/np_helper/pack_tril.c:24-37,58-69,99-107) - */
/* (v,o,v,o) view of the bound t2 slab, as updated_amps returns it: out[(a-a_begin),i,b,j] */
int mpccb_t2_to_aibj(mpccb_ctx* ctx, double* out_dev);
/* lower-triangle rows [a_begin*o, a_end*o) of the t2 buffer read as an (nov,nov) matrix */
int mpccb_pack_tril_rows(mpccb_ctx* ctx, double* seg_dev);
/* Several ranks: switch the t2 part of the DIIS vector from row-owned (rank = owner of the first virtual index;
 * contiguous but (2P-1):1 unbalanced) to column-owned (rank = owner of the second virtual index, read from the
 * local slab through t2[A,B,I,J] = t2[B,A,J,I]; balanced). prefix_dev: na*nvir + 1 device int64, number of packed
 * elements before local block k = (a - a_begin)*nvir + b (sharding.column_owned_prefix). Affects
 * mpccb_pack_tril_rows and mpccb_diis_push; NULL switches back. */
int mpccb_bind_diis_columns(mpccb_ctx* ctx, const long long* prefix_dev);
/* column-owned t2 part -> its positions in a full reference-order packed vector (nov(nov+1)/2 doubles, which the
 * caller zero-fills and sums over the ranks): the host-facing gather of dfrmpcc_lowlevel.py:162-168's vector */
int mpccb_diis_scatter_tril(mpccb_ctx* ctx, const double* seg_dev, double* full_packed_dev);
/* packed (nov(nov+1)/2) -> symmetric (nov,nov) -> (o,v,o,v) -> transpose(0,2,1,3): out (o,o,v,v) */
int mpccb_unpack_oovv(mpccb_ctx* ctx, const double* packed_dev, double* out_dev);

/* ---- DIIS vector algebra: lib.diis.DIIS.push_vec / update / extrapolate
 *      (pyscf/lib/diis.py:148-185, 221-231, 268-275). The 7x7 system is solved by the
 *      Python host with the same SciPy/NumPy calls as diis.py:251-265. ------------------ */
/* x_slot = [t1_new | packed local rows of the bound t2]; unless first_call:
 * e_slot = x_slot - xprev and dots_dev[k] = e_slot . e_k for k < nd (e_list[slot] is ignored:
 * e_slot itself is used). count_t1 = 0 leaves the (replicated) t1 part out of the dots.
 * first_call != 0: only x_slot is written (it is the DIIS `xprev`).
 * xprev, the previous extrapolated vector (diis.py:268-275), is either the explicit vector xprev_dev (n_prev == 0)
 * or — never stored — the combination sum_{k<n_prev} coeff_prev_host[k] * x_prev_list_dev[k] of the stored trial
 * vectors in slot order, recomputed per element with the arithmetic of mpccb_diis_combine: one write and one read
 * of the whole vector less per iteration. x_prev_list_dev[slot] may be x_slot_dev (the slot being replaced). */
int mpccb_diis_push(mpccb_ctx* ctx, const double* t1_new_dev, const double* xprev_dev,
                    const double* const* x_prev_list_dev, const double* coeff_prev_host, int n_prev,
                    double* x_slot_dev, double* e_slot_dev, const double* const* e_list_dev,
                    int nd, int slot, int first_call, int count_t1, double* dots_dev);
/* the same push for a plain device vector of n doubles (the CCSD-type stage packs its own vector,
 * pyscf/cc/ccsd.py:719-726, and feeds it to lib.diis.DIIS.update, ccsd.py:1269-1276) */
int mpccb_vec_diis_push(mpccb_ctx* ctx, const double* x_new_dev, long long n, const double* xprev_dev,
                        const double* const* x_prev_list_dev, const double* coeff_prev_host, int n_prev,
                        double* x_slot_dev, double* e_slot_dev, const double* const* e_list_dev,
                        int nd, int slot, int first_call, double* dots_dev);
/* lib.pack_tril / lib.unpack_tril(filltriu=SYMMETRIC) of one (n,n) row-major matrix
 * (NPdpack_tril, NPdunpack_tril + NPdsymm_triu: pyscf/lib/np_helper/pack_tril.c:24-37,58-69,99-107) */
int mpccb_pack_tril(mpccb_ctx* ctx, const double* M_dev, long long n, double* packed_dev);
int mpccb_unpack_tril(mpccb_ctx* ctx, const double* packed_dev, long long n, double* full_dev);
/* xnew = sum_{k<nd} coeff[k] * x_list[k]  (n doubles each) */
int mpccb_diis_combine(mpccb_ctx* ctx, const double* const* x_list_dev, const double* coeff_host,
                       int nd, double* xnew_dev, long long n);

/* tiles the fused eris -> t2 kernel launches (mirror scheme) and the size of the full tile grid */
int mpccb_t2_tiles(const mpccb_ctx* ctx, long long* launched, long long* total);
/* the mirror-scheme rule as a pure host function (tests): bit 0 = the rank owning a computes block (a,b),
 * bit 1 = it also writes the mirror block (b,a), bits 8.. = rank owning b */
int mpccb_sym_query(int nranks, int rank, const int* slab_begin, int peers, int a, int b);

/* ---- introspection for tests and profiling -------------------------------------------- */
/* named scratch tensor (e.g. "JK", "Y", "Xoo", "Fvo"); see DESIGN.md for layouts */
int mpccb_get_buffer(mpccb_ctx* ctx, const char* name, double** ptr_dev, long long* count);
int mpccb_copy_to_host(mpccb_ctx* ctx, const double* src_dev, double* dst_host, long long count);
int mpccb_copy_to_device(mpccb_ctx* ctx, const double* src_host, double* dst_dev, long long count);
/* number of kernel launches issued by this context so far */
long long mpccb_launch_count(const mpccb_ctx* ctx);
/* per-stage device timing (CUDA events on the context's stream), enabled by option "profile":
 * stage names via mpccb_profile_stage_name(0..mpccb_profile_stages()-1); mpccb_profile_read
 * synchronises and returns summed milliseconds + counts per stage since the previous read */
int mpccb_profile_stages(void);
const char* mpccb_profile_stage_name(int stage);
int mpccb_profile_read(mpccb_ctx* ctx, double* ms_out, long long* count_out);
/* option by name: "tma" (0/1 use the TMA-fed kernels when shapes allow; default 1),
 * "profile" (0/1 record stage timings; default 0), "t2_wide_tiles" (1: 128x128 tiles / one CTA per SM for the
 * fused eris->t2 GEMM, 0: 128x64 tiles / two CTAs per SM; default: the better-filled shape; set before bind_workspace), "t2_unfused" (1: GEMM into the t2 buffer + in-place
 * t2 pass instead of the fused epilogue; default: only when whole (i,j) blocks fill the tile < 70 %, e.g. nocc = 80,
 * or nocc > 128), "yvo_split" (force the split-K factor of the Yvo GEMM; default 0 = chosen from the tile count so that
 * the last wave of CTAs is not mostly idle when the slab is small) */
int mpccb_set_option(mpccb_ctx* ctx, const char* name, long long value);
/* standalone C[m,n] = sum_k A[m,k] B[n,k] through the generic DMMA kernel (row-major, test hook) */
int mpccb_gemm_nt(mpccb_ctx* ctx, const double* A_dev, const double* B_dev, double* C_dev,
                  int M, int N, int K);
/* y = alpha * op(A) x + beta * y: trans 0: A [rows][K]; trans 1: A [K][rows]. The einsums of rmpccsd_slow.py /
 * rintermediates.py in which one operand is summed away entirely (e.g. 'kdac,kd->ac', rintermediates.py:77) */
int mpccb_gemv(mpccb_ctx* ctx, const double* A_dev, const double* x_dev, double* y_dev, long long rows, long long K,
               int trans, double alpha, double beta);
/* the same product stored through a two-level index map with axpby:
 *   C[(m/mdiv)*m_hi + (m%mdiv)*m_lo + (n/ndiv)*n_hi + (n%ndiv)*n_lo] = alpha * sum_k A[m][k] B[n][k] + beta * C[..]
 * — lib.einsum's final transpose (pyscf/lib/numpy_helper.py:246-252) and the "+=" of the residual terms
 * (rmpccsd_slow.py:93-158) folded into the GEMM's store epilogue */
int mpccb_gemm_nt_mapped(mpccb_ctx* ctx, const double* A_dev, const double* B_dev, double* C_dev, int M, int N, int K,
                         int mdiv, long long m_hi, long long m_lo, int ndiv, long long n_hi, long long n_lo,
                         double alpha, double beta);

/* ---- tensor engine for the CCSD-type residual terms: the reference evaluates
 *      pyscf/cc/rmpccsd_slow.py:73-180 and pyscf/cc/rintermediates.py:29-130 as a list of lib.einsum calls
 *      (-> lib.dot -> NPdgemm, pyscf/lib/numpy_helper.py:101-252, pyscf/lib/np_helper/npdot.c:32-139); here every
 *      einsum is permute -> mpccb_gemm_nt (DMMA) -> permute/accumulate on the device. ------------------------ */
/* out = alpha * transpose(in, perm) + beta * out; numpy semantics (out axis d = in axis perm[d]); rank <= 6 */
int mpccb_permute_axpby(mpccb_ctx* ctx, const double* in_dev, double* out_dev, int rank, const int* in_dims,
                        const int* perm, double alpha, double beta);
/* t1[i,a] /= eps_o[i]-eps_v[a] (rank 2) or t2[i,j,a,b] /= eps_o[i]+eps_o[j]-eps_v[a]-eps_v[b] (rank 4)
 * (rmpccsd_slow.py:164-169) */
int mpccb_div_denominators(mpccb_ctx* ctx, double* t_dev, int rank);
/* t[x0..x_{r-1}] /= sum_d signs[d] * e_dev[d][x_d], rank <= 4: denominators in any index order / spin combination
 * (pyscf/cc/umpcc_lowlevel.py:468-476 D_aa, D_ab, D_bb in [a,i,b,j] order; :350-354 the singles denominators) */
int mpccb_div_general(mpccb_ctx* ctx, double* t_dev, int rank, const int* dims, const double* const* e_dev,
                      const double* signs);
/* active-space overwrite tnew[idx] = told[idx] (rmpccsd_slow.py:171-178, index classes of get_index_tuples :39-71):
 * bit (2*act_h[i] + act_p[a]) of class_mask for singles, bit (8*act_h[i] + 4*act_h[j] + 2*act_p[a] + act_p[b]) for
 * doubles; act_h_dev / act_p_dev are 0/1 int flags per occupied / virtual orbital */
int mpccb_class_overwrite(mpccb_ctx* ctx, double* tnew_dev, const double* told_dev, int rank,
                          const int* act_h_dev, const int* act_p_dev, unsigned class_mask);
/* lib.take_2d / lib.takebak_2d (pyscf/lib/numpy_helper.py:478-541; NPdtake_2d / NPdtakebak_2d of libnp_helper) on
 * device buffers: out[r*ny + c] = a[idx[r]*lda + idy[c]]  and  out[idx[r]*ldo + idy[c]] += alpha * a[r*ny + c]
 * (idx / idy: int32 device arrays without duplicates). They carry ccsd.amplitudes_to_vector_s4 / _unpack_4fold
 * (pyscf/cc/ccsd.py:705-757) for the run_diis of the unrestricted prototype (pyscf/cc/umpcc_lowlevel.py:15-46,92-98) */
int mpccb_take_2d(mpccb_ctx* ctx, const double* a_dev, long long lda, const int* idx_dev, int nx,
                  const int* idy_dev, int ny, double* out_dev);
int mpccb_takebak_2d(mpccb_ctx* ctx, double* out_dev, long long ldo, const double* a_dev, const int* idx_dev, int nx,
                     const int* idy_dev, int ny, double alpha);
/* out_dev[0] = x . y (deterministic two-pass reduction) */
int mpccb_dot(mpccb_ctx* ctx, const double* x_dev, const double* y_dev, long long n, double* out_dev);
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it). mpccb_dot with its per-block partial sums in a
 * file-scope device array of the library (1024 doubles = 8 KB on every GPU that loads it, held until exit) instead of
 * the caller-allocated workspace: allocates nothing, needs no bound workspace, same kernels and block count as
 * mpccb_dot (identical bits). Maximum n = 2^32 elements; larger n is rejected with an error. All calls on a device
 * share the array (it keeps the last call's partials; nothing in it needs clearing): calls that overlap on the GPU
 * race — serialise them. */
int mpccb_dot_static(mpccb_ctx* ctx, const double* x_dev, const double* y_dev, long long n, double* out_dev);
/* test hook: q[i] = x[i] / y[i] with the branch-free quotient of the fused t2 epilogue (replaces the IEEE division of
 * dfrmpcc_lowlevel.py:137; within 1 ulp for normal denominators — tests/test_gpu_parity.py sweeps it) */
int mpccb_test_fast_div(mpccb_ctx* ctx, const double* x_dev, const double* y_dev, double* q_dev, long long n);

#ifdef __cplusplus
}
#endif
#endif /* MPCC_B200_H */
