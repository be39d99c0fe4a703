"""Device-resident Pulay DIIS with the semantics of ``pyscf.lib.diis.DIIS``.

The reference low-level solver creates ``lib.diis.DIIS()`` and calls
``update(vec)`` without an error vector (``pyscf/mpcc/dfrmpcc_lowlevel.py:80,155-160``).
This class keeps the same bookkeeping on the host (``pyscf/lib/diis.py:148-185``
slot ring, ``:221-231`` B matrix, ``:251-265`` eigh guard + ``numpy.linalg.solve``)
while the vectors (``x0..x5``, ``e0..e5``, ``xprev``) stay in HBM and are touched
only by the kernels behind ``mpccb_diis_push`` / ``mpccb_diis_combine``:

* push: one pass that packs the new amplitudes into slot ``s``, forms
  ``e_s = x_s - xprev`` and accumulates all ``e_s . e_k`` (deterministic reduction);
* the (nd+1) x (nd+1) system is solved on the host with the reference's calls;
* combine: ``xprev = sum_k c_k x_k`` in slot order (``diis.py:268-275``) — only its first
  nov elements (the t1 the next iteration starts from) are formed right away. The rest of
  ``xprev`` is never stored: the next push recomputes it per element from the stored
  trial vectors and the coefficients (one write and one read of the vector less per
  iteration), and a host-facing access (``.t2`` after ``run_diis``) materialises it on demand.

Unlike the reference there is no HDF5 spill (``diis.py:123-139``): 13 vectors of
4.3 GB fit in 180 GB of HBM at the target shape.
"""
from __future__ import annotations

import ctypes

import numpy as np
import scipy.linalg
import torch

from . import _cabi


_SYEVR_WORK = {}


def _eigh(h: np.ndarray):
    """``scipy.linalg.eigh(h)`` without its Python wrapper: the same LAPACK routine (dsyevr, lower triangle, all
    eigenpairs, workspace sizes from dsyevr_lwork) called directly — bit-identical eigenpairs (checked in
    tests/test_host_logic.py) in half the time; this solve sits on the critical path of every iteration."""
    n = h.shape[0]
    if h.dtype != np.float64:
        return scipy.linalg.eigh(h)
    work = _SYEVR_WORK.get(n)
    if work is None:
        r = scipy.linalg.lapack.dsyevr_lwork(n, lower=1)
        work = _SYEVR_WORK[n] = (int(r[0]), int(r[1]))
    w, v, _, _, info = scipy.linalg.lapack.dsyevr(h, lwork=work[0], liwork=work[1], lower=1, compute_v=1)
    if info != 0:
        return scipy.linalg.eigh(h)
    return w, v


def solve_diis_system(h: np.ndarray) -> np.ndarray:
    """c with ``h c = (1, 0, ..)``; pseudo-inverse if |eigenvalue| < 1e-14 (diis.py:251-265)."""
    g = np.zeros(h.shape[0], h.dtype)
    g[0] = 1
    w, v = _eigh(h)
    if np.any(abs(w) < 1e-14):
        idx = abs(w) > 1e-14
        return np.dot(v[:, idx] * (1. / w[idx]), np.dot(v[:, idx].T.conj(), g))
    return np.linalg.solve(h, g)


class DIIS:
    """``update(amps)`` where ``amps`` is an :class:`AmplitudeVector` bound to a solver."""

    def __init__(self, dev=None, filename=None, incore=True):
        self.space = 6
        self.min_space = 1
        self.verbose = getattr(dev, "verbose", 4)
        self.stdout = getattr(dev, "stdout", None)
        self.filename = filename
        self._bookkeep = []
        self._head = 0
        self._H = None
        self._xprev_buf = None      # torch tensor (device): the first vector; later the on-demand materialisation
        self._prev = None           # (coefficients, slots): xprev as a combination of the stored trial vectors
        self._prev_stored = False   # does _xprev_buf hold the current xprev?
        self._t1_next = None
        self._x = {}
        self._e = {}
        self.last_coeff = None
        self._pool = []
        self.generation = 0         # bumped whenever the extrapolation buffer (xprev) is overwritten in place

    def get_num_vec(self):
        return len(self._bookkeep)

    def get_vec(self, idx):
        return self._x[idx]

    def get_err_vec(self, idx):
        return self._e[idx]

    # -- device plumbing ----------------------------------------------------
    def _alloc(self, like_len, device):
        if self._pool:
            return self._pool.pop()
        return torch.empty(like_len, dtype=torch.float64, device=device)

    def reserve(self, solver):
        """Allocate the whole history (xprev + ``space`` x/e pairs) up front, so that no cudaMalloc —
        slow once NCCL has mapped peer memory — lands inside the iteration loop. Optional: without it
        the vectors are allocated on first use, as the reference's ``push_vec`` does (diis.py:148-185)."""
        need = 1 + 2 * self.space - (0 if self._xprev_buf is None else 1) - 2 * len(self._x) - len(self._pool)
        for _ in range(max(need, 0)):
            self._pool.append(torch.empty(solver._vec_len_local, dtype=torch.float64, device=solver.device))
        return self

    def update(self, amps):
        """Push the amplitudes the solver just produced and return the extrapolated
        :class:`AmplitudeVector` (``diis.py:202-243``)."""
        ll = amps.solver
        ctx = ll._ctx
        self.synced = False          # did this call wait for the solver's stream (after the push)?
        n_local = ll._vec_len_local
        dev = ll.device
        if len(self._bookkeep) >= self.space:
            self._bookkeep = self._bookkeep[1 - self.space:]

        if self._xprev_buf is None:
            # first trial vector only becomes xprev (diis.py:160-167); returned unchanged
            self._xprev_buf = self._alloc(n_local, dev)
            self._solver = ll
            ctx.call("mpccb_diis_push", amps.t1_dev.data_ptr(), None, None, None, 0, self._xprev_buf.data_ptr(), None,
                     None, 0, 0, 1, int(ll._count_t1), None)
            self._prev_stored = True
            self.generation += 1
            return AmplitudeVector(ll, vec=self._xprev_buf, t1_dev=self._xprev_buf[:ll.nov], owner=self,
                                   serial=self.generation)

        if self._head >= self.space:
            self._head = 0
        slot = self._head
        self._bookkeep.append(slot)
        if slot not in self._x:
            self._x[slot] = self._alloc(n_local, dev)
            self._e[slot] = self._alloc(n_local, dev)
        self._head += 1
        nd = self.get_num_vec()

        e_list = (ctypes.c_void_p * 6)()
        for k in range(6):
            e_list[k] = self._e[k].data_ptr() if (k < nd and k in self._e) else None
        xp_list = (ctypes.c_void_p * 6)()
        xp_coeff = (ctypes.c_double * 6)()
        n_prev = 0
        # the column-owned push kernel (several ranks) walks scattered block fragments and is bound by its address
        # arithmetic, not by HBM: six more vector streams cost it more than the combine pass saves (measured at two
        # ranks: 8.3 vs 5.1 ms per update), so only the row-owned form recomputes xprev on the fly
        implicit = ll._diis_implicit
        if self._prev is not None and implicit:
            coeffs, slots = self._prev
            n_prev = len(slots)
            for k, sl in enumerate(slots):
                xp_list[k] = self._x[sl].data_ptr()
                xp_coeff[k] = coeffs[k]
        dots = ll._dots_dev
        ctx.call("mpccb_diis_push", amps.t1_dev.data_ptr(), self._xprev_buf.data_ptr(), xp_list, xp_coeff, n_prev,
                 self._x[slot].data_ptr(), self._e[slot].data_ptr(), e_list, nd, slot, 0,
                 int(ll._count_t1), dots.data_ptr())
        if ll.world_size > 1:
            torch.distributed.all_reduce(dots, group=ll.group)
        ll._dots_host.copy_(dots, non_blocking=True)
        ll._sync()
        self.synced = True
        d = ll._dots_host.numpy()

        if nd < self.min_space:
            return AmplitudeVector(ll, vec=self._x[slot], t1_dev=self._x[slot][:ll.nov])

        if self._H is None:
            self._H = np.zeros((self.space + 1, self.space + 1))
            self._H[0, 1:] = self._H[1:, 0] = 1
        for i in range(nd):
            self._H[self._head, i + 1] = d[i]
            self._H[i + 1, self._head] = d[i]

        c = solve_diis_system(self._H[:nd + 1, :nd + 1])
        self.last_coeff = c
        # xprev = sum_k c[k+1] x_k (diis.py:268-275): kept as a recipe; only the t1 part is formed now
        self._prev = ([float(c[k + 1]) for k in range(nd)], list(range(nd)))
        self._prev_stored = False
        self.generation += 1
        if not implicit:
            # explicit xprev, formed at once. (Forming its t2 part on a side stream next to the following T1 dressing
            # was tried: a streaming kernel that fills every SM does not share them with the dressing's GEMMs, whose
            # CTAs take the whole register file — the two simply ran one after the other.)
            xp = self._xprev
            return AmplitudeVector(ll, vec=xp, t1_dev=xp[:ll.nov], owner=self, serial=self.generation)
        if self._t1_next is None:
            self._t1_next = torch.empty(ll.nov, dtype=torch.float64, device=dev)
        self._combine_into(self._t1_next, ll.nov)
        return AmplitudeVector(ll, vec=None, t1_dev=self._t1_next, owner=self, serial=self.generation, lazy=True)

    def _combine_into(self, out, n, offset=0):
        """out[offset : offset + n] = sum_k c_k x_k[offset : offset + n]"""
        coeffs, slots = self._prev
        x_list = (ctypes.c_void_p * 6)()
        coeff = (ctypes.c_double * 6)()
        for k, sl in enumerate(slots):
            x_list[k] = self._x[sl].data_ptr() + 8 * offset
            coeff[k] = coeffs[k]
        self._solver._ctx.call("mpccb_diis_combine", x_list, coeff, len(slots), out.data_ptr() + 8 * offset, n)

    @property
    def _xprev(self):
        """The current extrapolated vector as a device tensor (materialised on demand; host-facing paths only)."""
        if self._xprev_buf is None:
            return None
        if not self._prev_stored:
            self._combine_into(self._xprev_buf, self._xprev_buf.numel())
            self._prev_stored = True
        return self._xprev_buf


class AmplitudeVector:
    """What ``amplitudes_to_vector`` returns on the GPU path: either a handle on
    the solver's freshly built amplitudes (``vec is None``: t1 on device + the
    bound t2 buffer) or a packed device vector ``[t1 | tril rows]``.
    ``numpy()`` materialises the reference's flat vector on the host."""

    def __init__(self, solver, vec=None, t1_dev=None, owner=None, serial=0, lazy=False):
        self.solver = solver
        self.vec = vec
        self.t1_dev = t1_dev
        self.lazy = lazy            # the extrapolated vector of `owner`, not materialised yet
        # vectors returned by DIIS.update alias its extrapolation buffer: `owner.generation` moves on when the next
        # update overwrites it (the reference returns a fresh array, diis.py:268-275)
        self.owner = owner
        self.serial = serial

    def _check_fresh(self):
        if self.owner is not None and self.serial != self.owner.generation:
            raise RuntimeError("this amplitude vector is stale: a later DIIS update overwrote the extrapolated vector")

    def packed_device(self):
        ll = self.solver
        if self.vec is not None:
            self._check_fresh()
            return self.vec
        if self.lazy:
            self._check_fresh()
            self.vec = self.owner._xprev
            return self.vec
        out = torch.empty(ll._vec_len_local, dtype=torch.float64, device=ll.device)
        out[:ll.nov].copy_(self.t1_dev.reshape(-1))
        ll._ctx.call("mpccb_pack_tril_rows", out[ll.nov:].data_ptr())
        return out

    def numpy(self):
        """The reference's flat host vector. With several ranks this is a collective call
        (the packed t2 segments are all-gathered; the t1 part is replicated)."""
        ll = self.solver
        v = self.packed_device()
        if ll.world_size > 1 or ll._diis_columns:
            v = torch.cat([v[:ll.nov], ll._gather_packed(v[ll.nov:])])
        ll._sync()
        return v.cpu().numpy()

    def __array__(self, dtype=None, copy=None):
        return self.numpy()


class VectorDIIS:
    """``lib.diis.DIIS`` (``pyscf/lib/diis.py:100-275``) for plain device vectors, as the CCSD-type stage uses it
    (``ccsd.kernel`` builds ``lib.diis.DIIS(mycc, ...)`` with ``space = mycc.diis_space`` and ``run_diis`` feeds it
    ``amplitudes_to_vector(t1, t2)``, ``pyscf/cc/ccsd.py:71-76,1269-1276``). Same ring / B-matrix / solve as
    :class:`DIIS`; the vectors are flat buffers of a backend (``tensor_engine.CudaBackend`` in the product, a NumPy
    stand-in in the CPU tests) which provides ``vec_diis_push`` and ``combine``."""

    def __init__(self, backend, space=6, min_space=1):
        self.be = backend
        self.space = space
        self.min_space = min_space
        self._bookkeep = []
        self._head = 0
        self._H = None
        self._xprev = None
        self._prev = None
        self._x = {}
        self._e = {}
        self.last_coeff = None

    def get_num_vec(self):
        return len(self._bookkeep)

    def update(self, x, n):
        """Push the device vector ``x`` (n doubles) and return a NEW device vector with the extrapolation."""
        be = self.be
        if len(self._bookkeep) >= self.space:
            self._bookkeep = self._bookkeep[1 - self.space:]
        if self._xprev is None:
            self._xprev = be.empty(n)
            be.vec_diis_push(x, n, None, [], [], self._xprev, None, [], 0, 0, True)
            return be.copy(self._xprev)
        if self._head >= self.space:
            self._head = 0
        slot = self._head
        self._bookkeep.append(slot)
        if slot not in self._x:
            self._x[slot] = be.empty(n)
            self._e[slot] = be.empty(n)
        self._head += 1
        nd = self.get_num_vec()
        prev_x, prev_c = [], []
        if self._prev is not None:
            prev_c, slots = self._prev
            prev_x = [self._x[s] for s in slots]
        d = be.vec_diis_push(x, n, self._xprev, prev_x, prev_c, self._x[slot], self._e[slot],
                             [self._e[k] for k in range(nd)], nd, slot, False)
        if nd < self.min_space:
            return be.copy(self._x[slot])
        if self._H is None:
            self._H = np.zeros((self.space + 1, self.space + 1))
            self._H[0, 1:] = self._H[1:, 0] = 1
        for i in range(nd):
            self._H[self._head, i + 1] = d[i]
            self._H[i + 1, self._head] = d[i]
        c = solve_diis_system(self._H[:nd + 1, :nd + 1])
        self.last_coeff = c
        self._prev = ([float(c[k + 1]) for k in range(nd)], list(range(nd)))
        out = be.empty(n)
        be.combine([self._x[k] for k in range(nd)], self._prev[0], out, n)
        return out
