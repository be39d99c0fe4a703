"""TEST INFRASTRUCTURE — CPU restatement (NumPy) of the DF-MPCC low-level iteration.

This file is the *oracle* for the CUDA path. It is imported only by ``tests/``,
``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference`` legs
of ``bench.py``; nothing under ``pyscf_mpcc_b200/`` may import it.

Parity pin: every function below is checked against the reference's own code,
executed unmodified through ``oracle/ref_shim.py`` in the build container
(``tests/test_oracle_vs_reference.py``), and against the committed fixtures in
``tests/golden/`` that the same reference run produced
(``tests/golden/generate_golden.py``). The reference has no MPCC tests of its
own, so those reference-generated vectors are the pin.

What is restated (all citations relative to ``/root/reference``):

* ``LowLevelOracle.__init__``          — ``pyscf/mpcc/dfrmpcc_lowlevel.py:7-43``
* ``LowLevelOracle.mo_three_center``   — ``dfrmpcc_lowlevel.py:46-69``
* ``LowLevelOracle.updated_amps``      — ``dfrmpcc_lowlevel.py:103-152``
* ``LowLevelOracle.kernel``            — ``dfrmpcc_lowlevel.py:72-100``
* ``amps_to_vector / vector_to_amps``  — ``dfrmpcc_lowlevel.py:162-176``
* ``tril_pack / tril_unpack_symmetric``— ``pyscf/lib/numpy_helper.py:256-358``,
  ``pyscf/lib/np_helper/pack_tril.c:24-37,58-69,99-107``
* ``PulayDIIS``                        — ``pyscf/lib/diis.py:100-121,148-185,202-275``

Two arithmetic modes:

* ``mode="einsum"`` performs each contraction with a plain ``np.einsum`` (no
  ``optimize``), i.e. the same single-threaded sum-of-products loops the
  reference runs. This is the timed "reference CPU path" (kind ``"port"``).
* ``mode="blas"`` evaluates the same tensors through GEMM-shaped ``@`` products
  (multi-threaded OpenBLAS): the "BLAS-fair" line of SURVEY.md §8d, and the
  only mode that finishes in seconds at the large shapes.

Both agree to rounding (tests pin 1e-12); neither is used by the product.
"""
from __future__ import annotations

import numpy as np
import scipy.linalg

__all__ = [
    "tril_pack", "tril_unpack_symmetric", "amps_to_vector", "vector_to_amps",
    "PulayDIIS", "LowLevelOracle", "flops_per_updated_amps", "doubles_sample",
]


# --------------------------------------------------------------------------
# triangular packing helpers
# --------------------------------------------------------------------------
def tril_pack(mat: np.ndarray) -> np.ndarray:
    """Lower triangle, row by row: out[r(r+1)/2 + c] = mat[r, c] for c <= r.

    ``NPdpack_tril`` (pack_tril.c:99-107). ``mat`` need not be symmetric —
    only the lower triangle is looked at.
    """
    n = mat.shape[0]
    out = np.empty(n * (n + 1) // 2, dtype=mat.dtype)
    pos = 0
    for r in range(n):
        out[pos:pos + r + 1] = mat[r, :r + 1]
        pos += r + 1
    return out


def tril_unpack_symmetric(packed: np.ndarray) -> np.ndarray:
    """Inverse of :func:`tril_pack` with the upper triangle mirrored.

    ``NPdunpack_tril`` + ``NPdsymm_triu`` (pack_tril.c:58-69, 24-37).
    """
    n = int(round((np.sqrt(8.0 * packed.size + 1.0) - 1.0) / 2.0))
    assert n * (n + 1) // 2 == packed.size
    full = np.empty((n, n), dtype=packed.dtype)
    pos = 0
    for r in range(n):
        row = packed[pos:pos + r + 1]
        full[r, :r + 1] = row
        full[:r + 1, r] = row
        pos += r + 1
    return full


def amps_to_vector(t1: np.ndarray, t2: np.ndarray, nocc: int, nvir: int) -> np.ndarray:
    """[t1 | tril(M)] with M = t2.transpose(0,2,1,3) viewed as (nov, nov).

    dfrmpcc_lowlevel.py:162-168. The reference applies this to whatever 4-index
    array it is handed; ``updated_amps`` hands it the (v,o,v,o) tensor, so M is
    the [a,b,i,j]-ordered buffer read as an (nov x nov) row-major matrix — not
    symmetric; only its lower triangle survives.
    """
    nov = nocc * nvir
    vec = np.empty(nov + nov * (nov + 1) // 2, dtype=t1.dtype)
    vec[:nov] = t1.reshape(-1)
    m = np.ascontiguousarray(t2.transpose(0, 2, 1, 3)).reshape(nov, nov)
    vec[nov:] = tril_pack(m)
    return vec


def vector_to_amps(vec: np.ndarray, nocc: int, nvir: int):
    """dfrmpcc_lowlevel.py:170-176: t1 (o,v) copy; t2 symmetric-unpacked,
    viewed (o,v,o,v), transposed to (o,o,v,v), made C-contiguous."""
    nov = nocc * nvir
    t1 = np.array(vec[:nov]).reshape(nocc, nvir)
    full = tril_unpack_symmetric(vec[nov:])
    t2 = np.ascontiguousarray(full.reshape(nocc, nvir, nocc, nvir).transpose(0, 2, 1, 3))
    return t1, t2


# --------------------------------------------------------------------------
# DIIS
# --------------------------------------------------------------------------
class PulayDIIS:
    """Pulay DIIS as ``pyscf.lib.diis.DIIS`` runs it when ``update(x)`` is
    called without an error vector (diis.py:202-275).

    * ring of ``space`` = 6 slots, ``min_space`` = 1 (diis.py:109-110);
    * the first vector only becomes ``xprev`` and is returned unchanged
      (diis.py:160-167, 216-218);
    * afterwards slot ``s`` stores x and e = x - xprev (diis.py:169-185), and
      B[s+1, k+1] = e_s . e_k for the ``nd`` filled slots (diis.py:221-231);
    * row/column 0 of B is 1 with B[0,0] = 0; the system B c = (1,0,..)ᵀ is
      solved with ``numpy.linalg.solve`` unless ``scipy.linalg.eigh`` finds an
      eigenvalue below 1e-14 in magnitude, in which case the pseudo-inverse on
      the remaining eigenpairs is used (diis.py:251-265);
    * x_new = Σ c[k+1] x_k in slot order, and becomes the next ``xprev``
      (diis.py:268-275, 234-242).
    """

    def __init__(self, space: int = 6, min_space: int = 1):
        self.space = space
        self.min_space = min_space
        self.xs = {}
        self.es = {}
        self.order = []
        self.head = 0
        self.B = None
        self.xprev = None
        self.last_coeff = None

    def update(self, x: np.ndarray) -> np.ndarray:
        x = x.reshape(-1)
        if len(self.order) >= self.space:
            self.order = self.order[1 - self.space:]
        if self.xprev is None:
            self.xprev = x
            return x
        if self.head >= self.space:
            self.head = 0
        slot = self.head
        self.order.append(slot)
        self.xs[slot] = x
        self.es[slot] = x - self.xprev
        self.head += 1
        nd = len(self.order)
        if nd < self.min_space:
            return x
        if self.B is None:
            self.B = np.zeros((self.space + 1, self.space + 1))
            self.B[0, 1:] = 1.0
            self.B[1:, 0] = 1.0
        newest = self.es[slot]
        for k in range(nd):
            d = float(np.dot(newest, self.es[k]))
            self.B[self.head, k + 1] = d
            self.B[k + 1, self.head] = d
        c = self.solve(self.B[:nd + 1, :nd + 1])
        self.last_coeff = c
        # xnew = sum_k xs[k] * c[k+1] in slot order (diis.py:268-275), formed block by block so that no vector-sized
        # temporary exists (same products, same order of additions: bit-identical to the whole-array form)
        xnew = np.zeros_like(x)
        step = 1 << 22
        buf = np.empty(min(step, x.size))
        for lo in range(0, x.size, step):
            hi = min(lo + step, x.size)
            for k in range(nd):
                np.multiply(self.xs[k][lo:hi], c[k + 1], out=buf[:hi - lo])
                xnew[lo:hi] += buf[:hi - lo]
        self.xprev = xnew
        return xnew

    @staticmethod
    def solve(h: np.ndarray) -> np.ndarray:
        g = np.zeros(h.shape[0])
        g[0] = 1.0
        w, v = scipy.linalg.eigh(h)
        if np.any(np.abs(w) < 1e-14):
            keep = np.abs(w) > 1e-14
            return np.dot(v[:, keep] * (1.0 / w[keep]), np.dot(v[:, keep].T, g))
        return np.linalg.solve(h, g)


# --------------------------------------------------------------------------
# the low-level solver
# --------------------------------------------------------------------------
def flops_per_updated_amps(o: int, v: int, L: int) -> float:
    """Algorithmic FP64 flop count of one ``updated_amps``."""
    nov = o * v
    return (4.0 * L * nov * nov + 2.0 * nov * nov + 6.0 * L * v * v * o
            + 14.0 * L * o * o * v + 8.0 * L * o * v)


def doubles_sample(Jvo, Lov, eo, ev, a0, a1, mode="einsum"):
    """The two O(L·nov²) contractions and the t2 build (dfrmpcc_lowlevel.py:135-138)
    restricted to rows a0 <= a < a1 — the bounded CPU sample timed by bench.py.
    Returns (t2_slab [a,i,b,j], Yvo_slab [L,a,i], flops)."""
    L, v, o = Jvo.shape
    na = a1 - a0
    D = (eo[None, :, None, None] + eo[None, None, None, :]
         - ev[a0:a1, None, None, None] - ev[None, None, :, None])
    if mode == "einsum":
        eris = np.einsum("Lai,Ljb->aijb", Jvo[:, a0:a1], Jvo)
        t2 = (2 * eris - np.transpose(eris, (0, 3, 2, 1))) / D
        Y = np.einsum("aibj,Ljb->Lai", t2, Lov)
    else:
        Jm = Jvo.reshape(L, v * o)
        eris = (Jm[:, a0 * o:a1 * o].T @ Jm).reshape(na, o, v, o)
        t2 = (2 * eris - eris.transpose(0, 3, 2, 1)) / D
        Lvo = np.ascontiguousarray(Lov.transpose(0, 2, 1)).reshape(L, v * o)
        Y = (Lvo @ t2.reshape(na * o, v * o).T).reshape(L, na, o)
    return t2, Y, 4.0 * L * (na * o) * (o * v)


class LowLevelOracle:
    """State + methods of the reference's ``MPCC_LL`` restated on plain arrays."""

    def __init__(self, mf=None, *, blocks=None, mode: str = "blas", verbose: bool = False):
        """Either ``mf`` (duck-typed mean field, dfrmpcc_lowlevel.py:7-43) or
        ``blocks = (Loo, Lov, Lvv, foo, fvv, mo_energy)`` already in the MO basis."""
        assert mode in ("blas", "einsum")
        self.mode = mode
        self.verbose = verbose
        self.diis = True
        self.con_tol = 1e-6
        self.max_its = 50
        if mf is not None:
            self.nao = int(mf.mol.nao)
            self.nocc = int(mf.mol.nelec[0])
            self.nvir = self.nao - self.nocc
            self.naux = int(mf.with_df.get_naoaux())
            self.mo_energy = np.asarray(mf.mo_energy, dtype=float)
            self.mo_coeff = np.asarray(mf.mo_coeff, dtype=float)
            self.Loo, self.Lov, self.Lvv = self.mo_three_center(mf.with_df)
            fmo = self.mo_coeff.T @ np.asarray(mf.get_fock()) @ self.mo_coeff
            self.foo = fmo[:self.nocc, :self.nocc]
            self.fvv = fmo[self.nocc:, self.nocc:]
        else:
            self.Loo, self.Lov, self.Lvv, self.foo, self.fvv, self.mo_energy = blocks
            self.naux, self.nocc, self.nvir = self.Lov.shape
            self.nao = self.nocc + self.nvir
        eo = self.mo_energy[:self.nocc]
        ev = self.mo_energy[self.nocc:]
        self.eia = ev[None, :] - eo[:, None]              # [i,a] = e_a - e_i  (:40-42)
        self.t1 = None
        self.t2 = None
        self.history = []

    # -- :46-69 ----------------------------------------------------------
    def mo_three_center(self, with_df):
        nao, o = self.nao, self.nocc
        C = self.mo_coeff
        rows, cols = np.tril_indices(nao)
        chunks = []
        for packed in with_df.loop():
            blk = np.zeros((packed.shape[0], nao, nao))
            blk[:, rows, cols] = packed
            blk[:, cols, rows] = packed
            half = blk.reshape(-1, nao) @ C                      # (L*mu, q)
            half = half.reshape(-1, nao, nao).transpose(0, 2, 1)  # (L, q, mu)
            full = (half.reshape(-1, nao) @ C).reshape(-1, nao, nao).transpose(0, 2, 1)  # (L, p, q)
            chunks.append(full)
        Lpq = np.concatenate(chunks, axis=0)
        return (np.ascontiguousarray(Lpq[:, :o, :o]),
                np.ascontiguousarray(Lpq[:, :o, o:]),
                np.ascontiguousarray(Lpq[:, o:, o:]))

    def denominators(self):
        """D[a,i,b,j] = -eia[i,a] - eia[j,b] (:43) — materialised only by tests."""
        e = self.eia
        return -(e.T[:, :, None, None] + e.T[None, None, :, :])

    # -- :103-152 --------------------------------------------------------
    def updated_amps(self, t1=None, a_slab=None):
        """One residual evaluation from ``t1`` (defaults to ``self.t1``).

        Returns (res, dE, t1_new, t2) with t2 shaped (v,o,v,o) = [a,i,b,j].
        ``a_slab=(a0,a1)`` restricts the two O(L·nov²) contractions (eris, Yvo)
        and the t2 build to rows a0 <= a < a1 (used only to time a bounded
        sample of a large workload; outputs are then partial and meaningless
        as amplitudes).
        """
        if t1 is None:
            t1 = self.t1
        if self.mode == "einsum":
            return self._updated_amps_einsum(t1, a_slab)
        return self._updated_amps_blas(t1, a_slab)

    def _updated_amps_blas(self, t1, a_slab):
        o, v, L = self.nocc, self.nvir, self.naux
        Loo, Lov, Lvv = self.Loo, self.Lov, self.Lvv
        # T1 dressing (:119-124)
        Xvo = (Lvv.reshape(L * v, v) @ t1.T).reshape(L, v, o)
        Xoo = (Lov.reshape(L * o, v) @ t1.T).reshape(L, o, o)
        X = Lov.reshape(L, o * v) @ t1.reshape(o * v)
        Joo = Xoo + Loo
        Joo_t1 = (Joo.reshape(L * o, o) @ t1).reshape(L, o, v)          # [L,i,a]
        Jvo = Xvo + Lov.transpose(0, 2, 1) - Joo_t1.transpose(0, 2, 1)
        # singles pieces (:126-131)
        Xoo_t1 = (Xoo.reshape(L * o, o) @ t1).reshape(L, o, v)          # [L,j,a]
        W = Xoo_t1.transpose(0, 2, 1) - Xvo                             # [L,a,j]
        Om = np.einsum("Laj,Lji->ai", W, Joo, optimize=True)
        JvoX = np.tensordot(X, Jvo, axes=(0, 0))                        # [a,i]
        Om += JvoX
        Om += (t1 @ self.fvv).T
        Om -= (self.foo @ t1).T
        # Fov (:133)
        Fov = JvoX.T - np.einsum("Lij,Lib->jb", Xoo, Lov, optimize=True)
        # doubles (:135-139)
        Jm = Jvo.reshape(L, v * o)
        a0, a1 = (0, v) if a_slab is None else a_slab
        eo = self.mo_energy[:o]
        ev = self.mo_energy[o:]
        rows = Jm[:, a0 * o:a1 * o]
        eris = (rows.T @ Jm).reshape(a1 - a0, o, v, o)                  # [a,i,b,j]
        denom = (eo[None, :, None, None] + eo[None, None, None, :]
                 - ev[a0:a1, None, None, None] - ev[None, None, :, None])
        t2 = (2.0 * eris - eris.transpose(0, 3, 2, 1)) / denom
        del eris
        Lvo_flat = Lov.transpose(0, 2, 1).reshape(L, v * o)             # [L,(b,j)]
        Y = np.zeros((L, v, o))
        Y[:, a0:a1, :] = (Lvo_flat @ t2.reshape((a1 - a0) * o, v * o).T).reshape(L, a1 - a0, o)
        Om[a0:a1, :] += (t2.reshape((a1 - a0) * o, v * o)
                         @ Fov.T.reshape(v * o)).reshape(a1 - a0, o)
        # doubles -> singles (:141-143); '+' sign kept as written
        Jvv = np.einsum("Ljb,ja->Lba", Lov, t1, optimize=True) + Lvv
        Om += np.einsum("Lba,Lbi->ai", Jvv, Y, optimize=True)
        Om -= np.einsum("Lji,Laj->ai", Joo, Y, optimize=True)
        # energy (:146-147)
        e1 = Xoo_t1.transpose(0, 2, 1) + X[:, None, None] * t1.T[None, :, :] + Jvo
        dE = float(np.vdot(e1, Y))
        # update (:149-151)
        upd = Om.T / self.eia
        res = float(np.linalg.norm(t1 + upd))
        return res, dE, -upd, t2

    def _updated_amps_einsum(self, t1, a_slab):
        """Same tensors, each formed by one un-optimised ``np.einsum`` exactly
        as the reference spells them (:119-147)."""
        o, v, L = self.nocc, self.nvir, self.naux
        Loo, Lov, Lvv = self.Loo, self.Lov, self.Lvv
        es = np.einsum
        Xvo = es("Lab,ib->Lai", Lvv, t1)
        Xoo = es("Lia,ja->Lij", Lov, t1)
        X = es("Lia,ia->L", Lov, t1)
        Joo = Xoo + Loo
        Jvo = Xvo + np.transpose(Lov, (0, 2, 1)) - es("Lij,ja->Lai", Joo, t1)
        Om = -1 * es("Laj,Lji->ai", Xvo, Joo)
        Om += es("Ljk,ka,Lji->ai", Xoo, t1, Joo)
        Om += es("Lai,L->ai", Jvo, X)
        Om += es("ib,ba->ai", t1, self.fvv)
        Om -= es("ka,ik->ai", t1, self.foo)
        Fov = es("Lbj,L->jb", Jvo, X) - es("Lij,Lib->jb", Xoo, Lov)
        a0, a1 = (0, v) if a_slab is None else a_slab
        eris = es("Lai,Ljb->aijb", Jvo[:, a0:a1], Jvo)                 # really [a,i,b,j]
        e = self.eia
        D = -(e.T[a0:a1, :, None, None] + e.T[None, None, :, :])
        t2 = (2 * eris - np.transpose(eris, (0, 3, 2, 1))) / D
        Y = np.zeros((L, v, o))
        Y[:, a0:a1] = es("aibj,Ljb->Lai", t2, Lov)
        Om[a0:a1] += es("aijb,bj->ai", t2, Fov)
        Jvv = es("Ljb,ja->Lba", Lov, t1) + Lvv
        Om += es("Lba,Lbi->ai", Jvv, Y)
        Om -= es("Lji,Laj->ai", Joo, Y)
        e1 = es("Lij,ja->Lai", Xoo, t1) + es("L,ia->Lai", X, t1) + Jvo
        dE = float(es("Lai,Lai", e1, Y))
        res = float(np.linalg.norm(t1 + Om.T / self.eia))
        return res, dE, -Om.T / self.eia, t2

    def updated_amps_packed(self, t1=None, slab=32):
        """``updated_amps`` (:103-152) followed by ``amplitudes_to_vector`` (:162-168) for shapes whose t2 (8.6 GB at
        (64,512,1536)) should not be held next to ``eris``: the doubles part runs over slabs of ``slab`` values of
        the first virtual index — ``eris``, ``t2``, ``Yvo`` rows and the rows of the (nov x nov) matrix the DIIS
        vector packs are all slab-local. GEMM-shaped arithmetic as ``mode="blas"``; agrees with
        ``_updated_amps_blas`` to rounding (``tests/test_oracle_golden.py::test_packed_form_equals_plain``).

        Returns (res, dE, t1_new, vec, t2_sumsq) with ``vec = [t1_new | tril(M)]`` the reference's DIIS vector.
        """
        if t1 is None:
            t1 = self.t1
        o, v, L = self.nocc, self.nvir, self.naux
        nov = o * v
        Loo, Lov, Lvv = self.Loo, self.Lov, self.Lvv
        Xvo = (Lvv.reshape(L * v, v) @ t1.T).reshape(L, v, o)
        Xoo = (Lov.reshape(L * o, v) @ t1.T).reshape(L, o, o)
        X = Lov.reshape(L, o * v) @ t1.reshape(o * v)
        Joo = Xoo + Loo
        Joo_t1 = (Joo.reshape(L * o, o) @ t1).reshape(L, o, v)
        Jvo = Xvo + Lov.transpose(0, 2, 1) - Joo_t1.transpose(0, 2, 1)
        del Joo_t1
        Xoo_t1 = (Xoo.reshape(L * o, o) @ t1).reshape(L, o, v)
        W = Xoo_t1.transpose(0, 2, 1) - Xvo
        Om = np.einsum("Laj,Lji->ai", W, Joo, optimize=True)
        del W
        JvoX = np.tensordot(X, Jvo, axes=(0, 0))
        Om += JvoX
        Om += (t1 @ self.fvv).T
        Om -= (self.foo @ t1).T
        Fov = JvoX.T - np.einsum("Lij,Lib->jb", Xoo, Lov, optimize=True)
        Jm = np.ascontiguousarray(Jvo.reshape(L, nov))
        eo = self.mo_energy[:o]
        ev = self.mo_energy[o:]
        Lvo_flat = np.ascontiguousarray(Lov.transpose(0, 2, 1)).reshape(L, nov)
        Fvo_flat = np.ascontiguousarray(Fov.T).reshape(nov)
        Y = np.zeros((L, v, o))
        vec = np.empty(nov + nov * (nov + 1) // 2)
        t2_sumsq = 0.0
        for a0 in range(0, v, slab):
            a1 = min(a0 + slab, v)
            na = a1 - a0
            eris = (Jm[:, a0 * o:a1 * o].T @ Jm).reshape(na, o, v, o)               # [a,i,b,j] (:135)
            denom = (eo[None, :, None, None] + eo[None, None, None, :]
                     - ev[a0:a1, None, None, None] - ev[None, None, :, None])
            t2 = (2.0 * eris - eris.transpose(0, 3, 2, 1)) / denom                   # (:137)
            del eris, denom
            t2m = t2.reshape(na * o, nov)
            Y[:, a0:a1, :] = (Lvo_flat @ t2m.T).reshape(L, na, o)                     # (:138)
            Om[a0:a1, :] += (t2m @ Fvo_flat).reshape(na, o)                           # (:139)
            t2_sumsq += float(np.vdot(t2m, t2m))
            # rows a0*o .. a1*o of M = t2.transpose(0,2,1,3) read as (nov, nov)  (:166)
            rows = np.ascontiguousarray(t2.transpose(0, 2, 1, 3)).reshape(na * o, nov)
            del t2, t2m
            for q in range(na * o):
                r = a0 * o + q
                p0 = nov + r * (r + 1) // 2
                vec[p0:p0 + r + 1] = rows[q, :r + 1]
            del rows
        Om += np.einsum("Lba,Lbi->ai", Lvv, Y, optimize=True)                         # Jvv = Lov.t1 + Lvv (:141-142)
        Z = np.einsum("Ljb,Lbi->ji", Lov, Y, optimize=True)
        Om += t1.T @ Z
        Om -= np.einsum("Lji,Laj->ai", Joo, Y, optimize=True)                         # (:143)
        e1 = Xoo_t1.transpose(0, 2, 1) + X[:, None, None] * t1.T[None, :, :] + Jvo    # (:146)
        dE = float(np.vdot(e1, Y))
        upd = Om.T / self.eia
        res = float(np.linalg.norm(t1 + upd))
        vec[:nov] = (-upd).reshape(-1)
        return res, dE, -upd, vec, t2_sumsq

    # -- :72-100, :155-160 -------------------------------------------------
    def run_diis(self, t1, t2, adiis):
        vec = amps_to_vector(t1, t2, self.nocc, self.nvir)
        return vector_to_amps(adiis.update(vec), self.nocc, self.nvir)

    def kernel(self):
        o, v = self.nocc, self.nvir
        self.t1 = np.zeros((o, v))
        self.t2 = np.zeros((o, o, v, v))
        self.history = []
        err, count = np.inf, 0
        adiis = PulayDIIS()
        dE = None
        while err > self.con_tol and count < self.max_its:
            res, dE, t1_new, t2_new = self.updated_amps()
            if self.diis:
                t1_next, t2_next = self.run_diis(t1_new, t2_new, adiis)
            else:
                t1_next, t2_next = t1_new, t2_new
            self.t1, self.t2 = t1_next, t2_next
            count += 1
            err = res
            self.history.append((count, dE, res))
            if self.verbose:
                print(f"It {count} Energy progress {dE:.6e} residual progress {res:.6e}")
        self.iterations = count
        return dE
