"""Device tensor-contraction engine for the CCSD-type residual terms.

The reference writes those terms as a list of ``lib.einsum`` calls
(``pyscf/cc/rmpccsd_slow.py:73-180``, ``pyscf/cc/rintermediates.py:29-130``); ``lib.einsum`` itself
turns each pairwise contraction into transpose -> ``lib.dot`` (BLAS dgemm) -> transpose
(``pyscf/lib/numpy_helper.py:101-252``). This engine does the same on the GPU:

    permute A to [M.., K..]  ->  permute B to [N.., K..]  ->  C[m,n] = sum_k A[m,k] B[n,k]  (FP64 DMMA GEMM)
    ->  out = alpha * permute(C) + beta * out

with hand-written kernels behind the C ABI (``mpccb_permute_axpby``, ``mpccb_gemm_nt`` — TMA-fed for the
large products —, ``mpccb_div_denominators``, ``mpccb_class_overwrite``, ``mpccb_dot``). PyTorch only owns the
buffers. The numerical primitives live in a *backend* object; the product backend is :class:`CudaBackend`
and there is no CPU backend in this package (tests inject a NumPy stand-in to check the formula
transcription without a GPU).
"""
from __future__ import annotations

import ctypes

import numpy as np


class DTensor:
    """Shape + flat device buffer (a torch float64 CUDA tensor for the product backend). ``const`` marks tensors
    that do not change between calls (integral blocks): the engine keeps their permuted copies."""
    __slots__ = ("buf", "shape", "const")

    def __init__(self, buf, shape, const=False):
        self.buf = buf
        self.shape = tuple(int(s) for s in shape)
        self.const = bool(const)

    @property
    def size(self):
        n = 1
        for s in self.shape:
            n *= s
        return n

    @property
    def ndim(self):
        return len(self.shape)


class CudaBackend:
    """Primitives on the GPU through ``libmpcc_b200.so``. ``ctx`` is a ``_cabi.Context`` with its workspace,
    stream and orbital energies bound (``MPCC_LL`` does that)."""

    def __init__(self, ctx, device):
        import torch
        if not torch.cuda.is_available():
            raise RuntimeError("the tensor engine needs a CUDA device; there is no CPU path")
        self.torch = torch
        self.ctx = ctx
        self.device = device

    def empty(self, n):
        return self.torch.empty(int(n), dtype=self.torch.float64, device=self.device)

    def zeros(self, n):
        return self.torch.zeros(int(n), dtype=self.torch.float64, device=self.device)

    def upload(self, array):
        a = np.ascontiguousarray(array, dtype=np.float64).reshape(-1)
        return self.torch.from_numpy(a).to(self.device)

    def upload_int(self, array):
        a = np.ascontiguousarray(array, dtype=np.int32).reshape(-1)
        return self.torch.from_numpy(a).to(self.device)

    def download(self, buf):
        self.torch.cuda.current_stream(self.device).synchronize()
        return buf.cpu().numpy()

    def copy(self, buf):
        return buf.clone()

    def permute_axpby(self, src, dst, dims, perm, alpha, beta):
        r = len(dims)
        self.ctx.call("mpccb_permute_axpby", src.data_ptr(), dst.data_ptr(), r,
                      (ctypes.c_int * r)(*dims), (ctypes.c_int * r)(*perm), float(alpha), float(beta))

    def gemm_nt(self, a, b, c, m, n, k):
        self.ctx.call("mpccb_gemm_nt", a.data_ptr(), b.data_ptr(), c.data_ptr(), int(m), int(n), int(k))

    def gemv(self, a, x, y, rows, k, trans, alpha, beta):
        self.ctx.call("mpccb_gemv", a.data_ptr(), x.data_ptr(), y.data_ptr(), int(rows), int(k), int(bool(trans)),
                      float(alpha), float(beta))

    def gemm_nt_mapped(self, a, b, c, m, n, k, mmap, nmap, alpha, beta):
        """c[map(m, n)] = alpha * a.b^T + beta * c[map(m, n)]; mmap / nmap = (div, hi, lo) two-level index maps."""
        self.ctx.call("mpccb_gemm_nt_mapped", a.data_ptr(), b.data_ptr(), c.data_ptr(), int(m), int(n), int(k),
                      int(mmap[0]), int(mmap[1]), int(mmap[2]), int(nmap[0]), int(nmap[1]), int(nmap[2]),
                      float(alpha), float(beta))

    def div_denominators(self, buf, rank):
        self.ctx.call("mpccb_div_denominators", buf.data_ptr(), int(rank))

    def div_general(self, buf, dims, evecs, signs):
        r = len(dims)
        self.ctx.call("mpccb_div_general", buf.data_ptr(), r, (ctypes.c_int * r)(*dims),
                      (ctypes.c_void_p * r)(*[e.data_ptr() for e in evecs]), (ctypes.c_double * r)(*signs))

    def class_overwrite(self, new, old, rank, act_h, act_p, mask):
        self.ctx.call("mpccb_class_overwrite", new.data_ptr(), old.data_ptr(), int(rank), act_h.data_ptr(),
                      act_p.data_ptr(), int(mask))

    def dot(self, x, y, n):
        out = self.empty(1)
        self.ctx.call("mpccb_dot", x.data_ptr(), y.data_ptr(), int(n), out.data_ptr())
        return float(self.download(out)[0])

    def take_2d(self, a, lda, idx, idy, out):
        """out[r, c] = a[idx[r], idy[c]] (lib.take_2d); idx / idy are int32 device arrays."""
        self.ctx.call("mpccb_take_2d", a.data_ptr(), int(lda), idx.data_ptr(), int(idx.numel()), idy.data_ptr(),
                      int(idy.numel()), out.data_ptr())

    def takebak_2d(self, out, ldo, a, idx, idy, alpha=1.0):
        """out[idx[r], idy[c]] += alpha * a[r, c] (lib.takebak_2d)."""
        self.ctx.call("mpccb_takebak_2d", out.data_ptr(), int(ldo), a.data_ptr(), idx.data_ptr(), int(idx.numel()),
                      idy.data_ptr(), int(idy.numel()), float(alpha))

    # -- vector layouts and DIIS of the CCSD-type stage (ccsd.py:719-735, 1269-1276) ------------------
    def pack_tril(self, mat, n, out):
        self.ctx.call("mpccb_pack_tril", mat.data_ptr(), int(n), out.data_ptr())

    def unpack_tril(self, packed, n, out):
        self.ctx.call("mpccb_unpack_tril", packed.data_ptr(), int(n), out.data_ptr())

    @staticmethod
    def _ptr_list(bufs):
        arr = (ctypes.c_void_p * 6)()
        for k, b in enumerate(bufs):
            arr[k] = b.data_ptr() if b is not None else None
        return arr

    def vec_diis_push(self, x, n, xprev, prev_x, prev_c, x_slot, e_slot, e_list, nd, slot, first):
        """One ``push_vec`` + B-matrix row (diis.py:148-185,221-231) on device vectors; returns the nd dots."""
        dots = self.empty(8)
        coeff = (ctypes.c_double * 6)(*([float(c) for c in prev_c] + [0.0] * (6 - len(prev_c))))
        self.ctx.call("mpccb_vec_diis_push", x.data_ptr(), int(n), xprev.data_ptr() if xprev is not None else None,
                      self._ptr_list(prev_x), coeff, len(prev_x), x_slot.data_ptr(),
                      e_slot.data_ptr() if e_slot is not None else None, self._ptr_list(e_list), int(nd), int(slot),
                      int(bool(first)), dots.data_ptr())
        return None if first else self.download(dots)[:nd]

    def combine(self, xs, coeffs, out, n):
        coeff = (ctypes.c_double * 6)(*([float(c) for c in coeffs] + [0.0] * (6 - len(coeffs))))
        self.ctx.call("mpccb_diis_combine", self._ptr_list(xs), coeff, len(xs), out.data_ptr(), int(n))

    def all_reduce(self, buf, group):
        """Sum over the ranks of ``group`` (NCCL); used by the a-slab sharded particle-particle ladder."""
        self.torch.cuda.current_stream(self.device).synchronize()
        self.torch.distributed.all_reduce(buf, group=group)

    def tick(self):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        return e

    def tock(self, start):
        e = self.torch.cuda.Event(enable_timing=True)
        e.record()
        e.synchronize()
        return start.elapsed_time(e)

    def rank_and_size(self, group):
        d = self.torch.distributed
        return d.get_rank(group), d.get_world_size(group)


def _prod(xs):
    n = 1
    for x in xs:
        n *= x
    return n


def two_level_map(indices, ext, stride):
    """(div, hi, lo) with  offset(r) = (r // div) * hi + (r % div) * lo  for the mixed-radix index r over `indices`
    (last fastest; extents `ext`, target strides `stride`), or None when no split into two linear groups exists.
    A group is linear when its strides nest: stride[x_k] == stride[x_{k+1}] * ext[x_{k+1}]."""
    idx = [ch for ch in indices if ext[ch] > 1]
    if not idx:
        return (1 << 30, 0, 0)

    def linear(group):
        return all(stride[a] == stride[b] * ext[b] for a, b in zip(group[:-1], group[1:]))

    for cut in range(len(idx) + 1):
        head, tail = idx[:cut], idx[cut:]
        if linear(head) and linear(tail):
            div = _prod(ext[ch] for ch in tail) if tail else 1
            hi = stride[head[-1]] if head else 0
            lo = stride[tail[-1]] if tail else 0
            if not head:
                div = 1 << 30
            return (div, hi, lo)
    return None


class Engine:
    """Pairwise einsum / transpose / axpy on :class:`DTensor` objects. ``flops`` accumulates 2*M*N*K of
    every GEMM issued (the algorithmic count used for throughput reports)."""

    def __init__(self, backend):
        self.be = backend
        self.flops = 0.0
        self.gemms = 0
        self.profile = None          # {spec: [ms, flops, calls]} when switched on (tools/cc_bench.py); synchronises per call
        self.fuse_output = True      # store products through index maps instead of permute passes (tests switch it off)
        self.fused_outputs = 0
        self.plan_layouts = True     # choose operand roles / index orders that avoid permuted copies (tests switch it off)
        # permuted copies of const tensors (integral blocks), made once: {(id(tensor), perm): (tensor, buffer)}
        self._perm_cache = {}
        self.perm_cache_bytes = 0
        self.perm_cache_limit = 16 << 30

    # -- construction ------------------------------------------------------------------------------
    def tensor(self, array):
        a = np.asarray(array, dtype=np.float64)
        return DTensor(self.be.upload(a), a.shape)

    def zeros(self, shape):
        return DTensor(self.be.zeros(_prod(shape)), shape)

    def numpy(self, t):
        return self.be.download(t.buf).reshape(t.shape)

    def copy(self, t):
        return DTensor(self.be.copy(t.buf), t.shape)

    # -- layout ------------------------------------------------------------------------------------
    def transpose(self, t, perm, alpha=1.0):
        """alpha * t.transpose(perm) as a new contiguous tensor."""
        perm = tuple(perm)
        shape = tuple(t.shape[p] for p in perm)
        out = DTensor(self.be.empty(t.size), shape)
        if t.size:
            self.be.permute_axpby(t.buf, out.buf, t.shape, perm, alpha, 0.0)
        return out

    def add(self, out, x, alpha=1.0, perm=None):
        """out += alpha * x.transpose(perm)   (perm None = identity)."""
        perm = tuple(range(x.ndim)) if perm is None else tuple(perm)
        if tuple(x.shape[p] for p in perm) != out.shape:
            raise ValueError(f"add: shape {x.shape} permuted by {perm} does not match {out.shape}")
        if out.size:
            self.be.permute_axpby(x.buf, out.buf, x.shape, perm, alpha, 1.0)
        return out

    def scale(self, t, alpha):
        return self.transpose(t, range(t.ndim), alpha)

    # -- contraction -------------------------------------------------------------------------------
    def einsum(self, spec, a, b, alpha=1.0, out=None, beta=0.0):
        """``out = alpha * einsum(spec, a, b) + beta * out`` for two operands without batch indices."""
        if self.profile is None:
            return self._einsum(spec, a, b, alpha, out, beta)
        f0 = self.flops
        t0 = self.be.tick()
        res = self._einsum(spec, a, b, alpha, out, beta)
        ms = self.be.tock(t0)
        rec = self.profile.setdefault(spec, [0.0, 0.0, 0])
        rec[0] += ms
        rec[1] += self.flops - f0
        rec[2] += 1
        return res

    def _einsum(self, spec, a, b, alpha=1.0, out=None, beta=0.0):
        lhs, io = spec.replace(" ", "").split("->")
        ia, ib = lhs.split(",")
        if len(ia) != a.ndim or len(ib) != b.ndim:
            raise ValueError(f"einsum {spec}: operand ranks {a.ndim},{b.ndim}")
        if len(set(ia)) != len(ia) or len(set(ib)) != len(ib) or len(set(io)) != len(io):
            raise NotImplementedError(f"einsum {spec}: repeated index inside one operand")
        ext = {}
        for idx, t in ((ia, a), (ib, b)):
            for ch, n in zip(idx, t.shape):
                if ext.setdefault(ch, n) != n:
                    raise ValueError(f"einsum {spec}: extent of '{ch}' differs ({ext[ch]} vs {n})")
        ksum = [ch for ch in ia if ch in ib and ch not in io]
        if any(ch in ia and ch in ib for ch in io):
            raise NotImplementedError(f"einsum {spec}: batch index (present in both operands and the output)")
        if any((ch not in ib and ch not in io) for ch in ia) or any((ch not in ia and ch not in io) for ch in ib):
            raise NotImplementedError(f"einsum {spec}: index summed inside a single operand")
        if self.plan_layouts and self.fuse_output:
            a, b, ia, ib, mi, ni, ksum = self._plan(a, b, ia, ib, io, ksum, ext)
        else:
            if io and io[-1] in ia and io[-1] not in ib and self.fuse_output:
                # the unit-stride index of the result should be a COLUMN index of the product (lanes of the store
                # epilogues run along n): C = A.B^T -> C^T = B.A^T, same flops
                a, b, ia, ib = b, a, ib, ia
            mi = [ch for ch in io if ch in ia]
            ni = [ch for ch in io if ch in ib]
        if len(mi) + len(ni) != len(io):
            raise ValueError(f"einsum {spec}: output index not found in the operands")
        M, N, K = _prod(ext[c] for c in mi), _prod(ext[c] for c in ni), _prod(ext[c] for c in ksum)
        oshape = tuple(ext[c] for c in io)
        if out is not None and out.shape != oshape:
            raise ValueError(f"einsum {spec}: out has shape {out.shape}, expected {oshape}")
        if out is None:
            if beta != 0.0:
                raise ValueError("beta without out")
            out = DTensor(self.be.empty(_prod(oshape)), oshape)
        if M * N == 0:
            return out
        if K == 0:
            raise NotImplementedError("empty contraction index")
        if (not mi or not ni) and hasattr(self.be, "gemv"):
            self.flops += 2.0 * M * N * K
            self.gemms += 1
            if not ni:
                return self._gemv(a, ia, b, ib, mi, ksum, ext, io, alpha, out, beta)
            return self._gemv(b, ib, a, ia, ni, ksum, ext, io, alpha, out, beta)
        am = self._to_matrix(a, ia, mi + ksum)
        bm = self._to_matrix(b, ib, ni + ksum)
        cm_idx = mi + ni
        perm = tuple(cm_idx.index(ch) for ch in io)
        direct = perm == tuple(range(len(io))) and alpha == 1.0 and beta == 0.0
        self.flops += 2.0 * M * N * K
        self.gemms += 1
        if direct:
            self.be.gemm_nt(am, bm, out.buf, M, N, K)
            return out
        # the permutation of the result (lib.einsum's last transpose) and "out = alpha * .. + beta * out" go into the
        # GEMM's store epilogue when rows and columns of the product each map onto the result through a two-level
        # index map; otherwise the product is permuted / accumulated in a second pass
        ostride, acc = {}, 1
        for ch in reversed(io):
            ostride[ch] = acc
            acc *= ext[ch]
        # results of less than one wave of tiles with a long contraction index ("->ia", "->ki", "->ac", "->klij":
        # K = o*v*v or v*v) stay on the split-K path of mpccb_gemm_nt (K spread over the machine, fixed-order
        # reduction); permuting the small result costs nothing
        tiles = -(-M // 128) * -(-N // 128)
        fuse = self.fuse_output and not (tiles < 148 and K >= 2048)
        mmap = two_level_map(mi, ext, ostride) if fuse else None
        nmap = two_level_map(ni, ext, ostride) if fuse else None
        if mmap is not None and nmap is not None:
            self.be.gemm_nt_mapped(am, bm, out.buf, M, N, K, mmap, nmap, alpha, beta)
            self.fused_outputs += 1
            return out
        cbuf = self.be.empty(M * N)
        self.be.gemm_nt(am, bm, cbuf, M, N, K)
        self.be.permute_axpby(cbuf, out.buf, tuple(ext[c] for c in cm_idx), perm, alpha, beta)
        return out

    def _gemv(self, mat, imat, vec, ivec, free, ksum, ext, io, alpha, out, beta):
        """One operand (``vec``) is summed away entirely: a matrix-vector product bound by one read of ``mat``. The
        matrix is used in place when its contracted indices are its leading or its trailing ones."""
        nk = len(ksum)
        kset = set(ksum)
        if set(imat[:nk]) == kset:
            korder, forder, trans, mbuf = list(imat[:nk]), list(imat[nk:]), True, mat.buf
        elif set(imat[nk:]) == kset and len(imat) > nk:
            korder, forder, trans, mbuf = list(imat[len(imat) - nk:]), list(imat[:len(imat) - nk]), False, mat.buf
        else:
            korder, forder, trans = list(ksum), list(free), False
            mbuf = self._to_matrix(mat, imat, forder + korder)
        xbuf = self._to_matrix(vec, ivec, korder)
        rows, K = _prod(ext[c] for c in forder), _prod(ext[c] for c in korder)
        if forder == list(io):
            self.be.gemv(mbuf, xbuf, out.buf, rows, K, trans, alpha, beta)
            return out
        tmp = self.be.empty(rows)
        self.be.gemv(mbuf, xbuf, tmp, rows, K, trans, 1.0, 0.0)
        self.be.permute_axpby(tmp, out.buf, tuple(ext[c] for c in forder), tuple(forder.index(ch) for ch in io), alpha, beta)
        return out

    def _to_matrix(self, t, idx, order):
        perm = tuple(idx.index(ch) for ch in order)
        if perm == tuple(range(len(idx))):
            return t.buf
        if t.const:
            hit = self._perm_cache.get((id(t), perm))
            if hit is not None:
                return hit[1]
        tmp = self.be.empty(t.size)
        self.be.permute_axpby(t.buf, tmp, t.shape, perm, 1.0, 0.0)
        if t.const and self.perm_cache_bytes + 8 * t.size <= self.perm_cache_limit:
            self._perm_cache[(id(t), perm)] = (t, tmp)        # holds `t`, so its id stays unique
            self.perm_cache_bytes += 8 * t.size
        return tmp

    def _plan(self, a, b, ia, ib, io, ksum, ext):
        """Operand roles and index orders of C[m,n] = sum_k A[m,k] B[n,k] that move the fewest bytes outside the GEMM.
        Any order of the row indices, of the column indices and of the contraction indices gives the same numbers up
        to the order of the K summation; what differs is which operands need a permuted copy first and whether the
        result can be stored in place. Candidates: either operand as A; K in the order of either operand; rows /
        columns in the order of the result or of their operand. Cost = bytes of the permuted copies (const tensors:
        made once, cached) + bytes of a separate result pass when the store cannot be fused, or is fused but scattered
        (neither the last row index nor the last column index is the result's unit-stride index)."""
        ostride, acc = {}, 1
        for ch in reversed(io):
            ostride[ch] = acc
            acc *= ext[ch]
        csize = 8.0 * acc

        def operand_cost(t, idx, order):
            if list(idx) == list(order):
                return 0.0
            if t.const:
                if (id(t), tuple(idx.index(ch) for ch in order)) in self._perm_cache:
                    return 0.0
                return 0.25 * 8.0 * t.size          # made once, amortised over the iterations
            return 2.0 * 8.0 * t.size

        best = None
        for swap in (False, True):
            xa, xb, xia, xib = (b, a, ib, ia) if swap else (a, b, ia, ib)
            fa = [ch for ch in io if ch in xia]
            fb = [ch for ch in io if ch in xib]
            m_orders = [fa] + ([[ch for ch in xia if ch in fa]] if [ch for ch in xia if ch in fa] != fa else [])
            n_orders = [fb] + ([[ch for ch in xib if ch in fb]] if [ch for ch in xib if ch in fb] != fb else [])
            k_orders = [[ch for ch in xia if ch in ksum]]
            kb = [ch for ch in xib if ch in ksum]
            if kb != k_orders[0]:
                k_orders.append(kb)
            for ko in k_orders:
                for mo in m_orders:
                    for no in n_orders:
                        cost = operand_cost(xa, xia, mo + ko) + operand_cost(xb, xib, no + ko)
                        if mo + no != list(io):
                            M, N, K = _prod(ext[c] for c in mo), _prod(ext[c] for c in no), _prod(ext[c] for c in ko)
                            tiles = -(-M // 128) * -(-N // 128)
                            mmap = two_level_map(mo, ext, ostride)
                            nmap = two_level_map(no, ext, ostride)
                            fused = mmap is not None and nmap is not None and not (tiles < 148 and K >= 2048)
                            unit = io[-1]
                            coalesced = (no and no[-1] == unit) or (mo and mo[-1] == unit)
                            if not fused:
                                cost += 3.0 * csize
                            elif not coalesced:
                                cost += 4.0 * csize
                        # mild preferences on ties: the result's unit-stride index among the columns, no role swap
                        cost += 1e-3 * csize * ((0 if (no and no[-1] == io[-1]) else 1) + (1 if swap else 0))
                        if best is None or cost < best[0]:
                            best = (cost, xa, xb, xia, xib, mo, no, ko)
        return best[1:]

    # -- reductions / elementwise ------------------------------------------------------------------
    def dot(self, x, y):
        if x.shape != y.shape:
            raise ValueError("dot: shapes differ")
        return self.be.dot(x.buf, y.buf, x.size) if x.size else 0.0

    def div_denominators(self, t):
        self.be.div_denominators(t.buf, t.ndim)
        return t

    def div_general(self, t, evecs, signs):
        """t[x0,..] /= sum_d signs[d] * evecs[d][x_d]; evecs are 1-D device tensors."""
        if t.size:
            self.be.div_general(t.buf, t.shape, [e.buf for e in evecs], signs)
        return t

    def class_overwrite(self, tnew, told, act_h, act_p, mask):
        if mask:
            self.be.class_overwrite(tnew.buf, told.buf, tnew.ndim, act_h, act_p, mask)
        return tnew

    # -- amplitude <-> vector (ccsd.py:719-735) ------------------------------------------------------
    def amplitudes_to_vector(self, t1, t2):
        """[t1 | pack_tril(t2.transpose(0,2,1,3).reshape(nov, nov))] as a flat device buffer."""
        o, v = t1.shape
        nov = o * v
        vec = self.be.empty(nov + nov * (nov + 1) // 2)
        self.be.permute_axpby(t1.buf, vec, t1.shape, (0, 1), 1.0, 0.0)
        m = self.transpose(t2, (0, 2, 1, 3))
        self.be.pack_tril(m.buf, nov, vec[nov:])
        return vec

    def vector_to_amplitudes(self, vec, o, v):
        """t1 (o,v) copy; t2 = unpack_tril(SYMMETRIC).reshape(o,v,o,v).transpose(0,2,1,3), contiguous."""
        nov = o * v
        t1 = DTensor(self.be.copy(vec[:nov]), (o, v))
        full = DTensor(self.be.empty(nov * nov), (o, v, o, v))
        self.be.unpack_tril(vec[nov:], nov, full.buf)
        return t1, self.transpose(full, (0, 2, 1, 3))


def make_cuda_engine(nocc, nvir, naux, device=None):
    """Stand-alone engine (no MPCC_LL around it): a native context sized for (nocc, nvir, naux) — the sizes
    only dimension its split-K / reduction scratch — with workspace and stream bound. Returns (engine, keepalive)."""
    import torch
    from . import _cabi
    if not torch.cuda.is_available():
        raise RuntimeError("the tensor engine needs a CUDA device; there is no CPU path")
    device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    if device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    torch.cuda.set_device(device)
    ctx = _cabi.Context(int(nocc), int(nvir), int(naux), 0, int(nvir), device.index)
    ctx.call("mpccb_set_stream", torch.cuda.current_stream(device).cuda_stream)
    nbytes = ctx.workspace_bytes()
    ws = torch.empty(nbytes + 256, dtype=torch.uint8, device=device)
    ctx.call("mpccb_bind_workspace", (ws.data_ptr() + 255) // 256 * 256, nbytes)
    return Engine(CudaBackend(ctx, device)), (ctx, ws)
