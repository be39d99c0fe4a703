"""CCSD-type doubles/singles residual of the multi-level solver on the GPU (SURVEY.md §8f rows f1 + f2).

Counterpart of ``pyscf/cc/rmpccsd_slow.py`` — the "high-level" MPCC stage the DF low-level solver is meant
to be embedded in (``pyscf/mpcc/mpcc.py:22-23``): the full RCCSD amplitude equations (Hirata et al., JCP 120,
2581 (2004), Eqs. 35-36) followed by the active-space overwrite that keeps the inactive ("environment")
amplitude classes at their input values.

    update_amps(cc, t1, t2, eris, act_hole, act_particle, idx_singles, idx_doubles) -> (t1new, t2new)   (:73-180)
    energy(cc, t1, t2, eris)                                                                            (:183-194)
    get_index_tuples(act_hole, act_particle, nocc, nvir)                                                (:39-71)
    init_amps(eris)                                                                                     (:206-219)

The four-index integrals come from the three-index DF tensors of the low-level solver
(``(pq|rs) = sum_L Lpq Lrs``): :class:`DFERIs` builds the oooo ... ovvv blocks as DMMA GEMMs on the device; the
particle-particle ladder ``sum_cd W_abcd tau_ijcd`` (:138, ``cc_Wvvvv`` rintermediates.py:92-99) is evaluated
**DF-direct in slabs of the first virtual index** so that v^4 integrals never exist in memory (the same blocking
idea as ``pyscf/cc/dfccsd.py:54-117``). The hole-hole ladder (``cc_Woooo``), the ring/exchange terms
(``cc_Wvoov``, ``cc_Wvovo``) and all singles terms follow the reference einsum by einsum through
:class:`pyscf_mpcc_b200.tensor_engine.Engine` (permute -> FP64 tensor-core GEMM -> permute/accumulate).

Amplitude layouts are the reference's: t1 (o,v), t2 (o,o,v,v); integrals in chemists' notation.
"""
from __future__ import annotations

import numpy as np

from .tensor_engine import DTensor, Engine


def get_index_tuples(act_hole, act_particle, nocc, nvir):
    """Index blocks of the singles (4) and doubles (16) amplitude classes, in the reference's order
    (rmpccsd_slow.py:39-71). Host-side helper; the device path uses :func:`class_masks`."""
    inact_particle = np.delete(np.arange(nvir), act_particle)
    inact_hole = np.delete(np.arange(nocc), act_hole)
    ah, ih, ap, ip = act_hole, inact_hole, act_particle, inact_particle
    singles = [np.ix_(ah, ip), np.ix_(ih, ap), np.ix_(ih, ip), np.ix_(ah, ap)]
    doubles = []
    for pa, pb in ((ip, ap), (ap, ip), (ip, ip), (ap, ap)):
        for hi, hj in ((ah, ah), (ih, ah), (ah, ih), (ih, ih)):
            doubles.append(np.ix_(hi, hj, pa, pb))
    # the reference lists the all-active block last and (inact,inact,act,act) before it (:65-68)
    doubles[12:16] = [doubles[13], doubles[14], doubles[15], doubles[12]]
    return singles, doubles


# class bit used by mpccb_class_overwrite: singles 2*act_h[i] + act_p[a]; doubles 8*act_h[i] + 4*act_h[j] + 2*act_p[a] + act_p[b]
_SINGLES_BITS = [0b10, 0b01, 0b00, 0b11]          # order of get_index_tuples()[0]


def _doubles_bits():
    bits = []
    for pa, pb in ((0, 1), (1, 0), (0, 0), (1, 1)):
        for hi, hj in ((1, 1), (0, 1), (1, 0), (0, 0)):
            bits.append(8 * hi + 4 * hj + 2 * pa + pb)
    bits[12:16] = [bits[13], bits[14], bits[15], bits[12]]
    return bits


_DOUBLES_BITS = _doubles_bits()


def class_masks(idx_singles, idx_doubles):
    """Bit masks over the active/inactive classes for the reference's class numbers (list positions)."""
    ms = 0
    for s in idx_singles:
        ms |= 1 << _SINGLES_BITS[s]
    md = 0
    for d in idx_doubles:
        md |= 1 << _DOUBLES_BITS[d]
    return ms, md


class DFERIs:
    """Four-index MO integrals from the DF tensors, on the device (chemists' notation, reference attribute
    names of ``ccsd._ChemistsERIs``): fock (host), oooo, ovoo, ooov, oovv, ovov, ovvo, ovvv; ``vvvv`` stays
    implicit — ``Lvv`` is kept for the DF-direct ladder."""

    def __init__(self, eng: Engine, Loo, Lov, Lvv, fock):
        self.eng = eng
        self.fock = np.asarray(fock, dtype=np.float64)
        L, o, v = Lov.shape
        self.nocc, self.nvir, self.naux = o, v, L
        E = eng.einsum
        self.Loo, self.Lov, self.Lvv = Loo, Lov, Lvv
        self.oooo = E("Lij,Lkl->ijkl", Loo, Loo)
        self.ovoo = E("Lia,Ljk->iajk", Lov, Loo)
        self.ooov = eng.transpose(self.ovoo, (2, 3, 0, 1))
        self.oovv = E("Lij,Lab->ijab", Loo, Lvv)
        self.ovov = E("Lia,Ljb->iajb", Lov, Lov)
        self.ovvo = eng.transpose(self.ovov, (0, 1, 3, 2))
        self.ovvv = E("Lia,Lbc->iabc", Lov, Lvv)
        self.vvvv = None
        # the integrals do not change during a solve: the engine keeps the permuted copies it needs of them
        for t in (Loo, Lov, Lvv, self.oooo, self.ovoo, self.ooov, self.oovv, self.ovov, self.ovvo, self.ovvv):
            t.const = True

    @classmethod
    def from_numpy(cls, eng, Loo, Lov, Lvv, fock):
        return cls(eng, eng.tensor(Loo), eng.tensor(Lov), eng.tensor(Lvv), fock)

    def get_ovvv(self):
        return self.ovvv


def _sym_add(eng, t2new, tmp, alpha):
    """t2new += alpha * (tmp + tmp.transpose(1,0,3,2))   — the P(ia,jb) symmetrisation of :142-155."""
    eng.add(t2new, tmp, alpha)
    eng.add(t2new, tmp, alpha, (1, 0, 3, 2))


def _rows(t, lo, hi):
    """Zero-copy view of t[lo:hi] (leading axis) of a contiguous tensor."""
    inner = t.size // t.shape[0] if t.shape[0] else 0
    return DTensor(t.buf[lo * inner:hi * inner], (hi - lo,) + t.shape[1:])


def _ladder_slabs(v, na_max, rank, world, o2=0, halve=True):
    """Slabs [a0, a1) of the first virtual index for this rank. One rank: pieces of at most na_max, sized so that the
    product against tau — (v - a0) * na rows, o2 columns, the larger of the slab's two GEMMs — fills whole waves of
    128 x 128 tiles on 148 SMs (equal pieces of v/6 left 12 waves where 9.1 were needed at (32,192,576)). Several
    ranks: 2*world (or a multiple) equal pieces, piece s and piece n-1-s to the same rank — the work of a slab is
    proportional to v - a0 (only the blocks b >= a0 are formed), so the two halves of a folded pair add up to the same
    for every rank."""
    if world == 1:
        if not o2:
            return [(a0, min(v, a0 + na_max)) for a0 in range(0, v, na_max)]
        ncol = -(-o2 // 128)
        out, a0 = [], 0
        while a0 < v:
            nb = (v - a0) if halve else v
            best = None
            for na in range(max(1, na_max // 2), min(na_max, v - a0) + 1):
                if v - (a0 + na) < max(1, na_max // 2) and a0 + na != v:
                    continue                      # would leave a sliver for the last slab
                tiles = -(-(na * nb) // 128) * ncol
                eff = round(tiles / (-(-tiles // 148) * 148.0), 6)
                key = (eff, na if tiles <= 148 else -na)      # ties: fewer launches below one wave, fewer flops above
                if best is None or key > best[0]:
                    best = (key, na)
            na = best[1] if best else min(na_max, v - a0)
            out.append((a0, a0 + na))
            a0 += na
        return out
    n = 2 * world
    while -(-v // n) > na_max:
        n += 2 * world
    bounds = [v * s // n for s in range(n + 1)]
    mine = []
    for s in range(n):
        owner = s % (2 * world)
        owner = owner if owner < world else 2 * world - 1 - owner
        if owner == rank and bounds[s + 1] > bounds[s]:
            mine.append((bounds[s], bounds[s + 1]))
    return mine


def _pair_symmetric(eng, t):
    """Is t[i,j,a,b] == t[j,i,b,a] (to rounding)? True for amplitudes that went through the DIIS vector (symmetric
    unpack of a lower triangle, ccsd.py:594-603) or came out of the symmetrised equations; not guaranteed for an
    arbitrary input of update_amps, e.g. after a class overwrite with a class list that is not closed under the swap."""
    d = eng.copy(t)
    eng.add(d, t, -1.0, (1, 0, 3, 2))
    return eng.dot(d, d) <= 1e-24 * max(eng.dot(t, t), 1e-300)


def _pp_ladder(eng, eris, t1, tau, t2new, slab_bytes=4 << 30, group=None, ring=None, halve=True):
    """t2new[i,j,a,b] += sum_cd W[a,b,c,d] tau[i,j,c,d]  (:138), W = cc_Wvvvv (rintermediates.py:92-99):
        W[a,b,c,d] = (ac|bd) - sum_k (kd|ac) t1[k,b] - sum_k (kc|bd) t1[k,a]
    DF-direct, in slabs of the first virtual index a, so that no v^4 array is ever stored (the blocking idea of
    ``pyscf/cc/dfccsd.py:54-117``). Both t1 terms fold into the DF tensor: with Lt[L,a,c] = sum_k t1[k,a] Lov[L,k,c]
    and Leff = Lvv - Lt,
        W[a,b,c,d] = sum_L Leff[L,a,c] Leff[L,b,d]  -  sum_L Lt[L,a,c] Lt[L,b,d],
    and the second piece is of rank o^2:  sum_L Lt[L,a,c] Lt[L,b,d] = sum_kl t1[k,a] t1[l,b] (kc|ld), so its ladder
    contribution is  sum_kl t1[k,a] t1[l,b] X[k,l,i,j],  X = sum_cd (kc|ld) tau[i,j,c,d]  — O(o^4 v^2) flops instead
    of a pass over every slab of W. Per slab: one GEMM (ac|bd)-shaped product stored straight in [a,b,c,d] order by
    the index-mapped epilogue, one GEMM against tau. Several ranks: slabs are dealt in folded pairs (_ladder_slabs),
    ``ring`` (a callable R -> None, see _ring_terms_into) adds this rank's share of the ring / exchange terms to the
    same [a,b,i,j] accumulator, and one all-reduce completes both."""
    E = eng.einsum
    be = eng.be
    v, L = eris.nvir, eris.naux
    o2 = tau.shape[0] * tau.shape[1]
    Lt = E("ka,Lkc->Lac", t1, eris.Lov)
    Leff = eng.copy(eris.Lvv)
    eng.add(Leff, Lt, -1.0)                                       # (Lvv - t1^T Lov)[L,a,c]
    LA = eng.transpose(Leff, (1, 2, 0))                           # [a, c, L]   rows of a slab are contiguous
    del Leff, Lt
    R = eng.zeros((v, v) + tau.shape[:2])                         # [a, b, i, j]
    # W[a,b,c,d] = W[b,a,d,c] and tau[i,j,c,d] = tau[j,i,d,c], hence R[a,b,i,j] = R[b,a,j,i]: per slab [a0,a1) only
    # the blocks b >= a0 are formed (the triangle pyscf's _add_vvvv keeps, ccsd.py / dfccsd.py:54-117); the square
    # diagonal slab block holds both members of a pair and enters with weight 1/2, and the final
    # t2new += R^T(2,3,0,1) + R^T(3,2,1,0) restores everything. About (1 + 1/nslabs)/2 of the ladder flops.
    na = max(1, min(v, int(slab_bytes // (8 * v * v * v)), -(-v // 4)))
    # several ranks (north_star: "the ladder ... shards naturally over ab"): every rank fills its blocks of R, one
    # all-reduce (sum with zeros) completes it.
    # halve=False (tau not pair-symmetric): every block b of a slab is formed and R enters once, untransposed; the ring
    # terms, whose symmetrisation is part of the equations, then get an accumulator of their own.
    rank, world = (0, 1) if group is None else be.rank_and_size(group)
    BIG = 1 << 30
    for a0, a1 in _ladder_slabs(v, na, rank, world, o2=o2, halve=halve):
        if not halve:
            n_a = a1 - a0
            W = be.empty(v * n_a * v * v)
            be.gemm_nt_mapped(LA.buf[a0 * v * L:], LA.buf, W, n_a * v, v * v, L, (v, v * v, v), (v, n_a * v * v, 1),
                              1.0, 0.0)
            be.gemm_nt_mapped(W, tau.buf, R.buf[a0 * v * o2:], v * n_a, o2, v * v, (n_a, o2, v * o2), (BIG, 0, 1),
                              1.0, 0.0)
            eng.flops += 2.0 * n_a * v * v * v * (L + o2)
            eng.gemms += 2
            del W
            continue
        n_a, n_b = a1 - a0, v - a0
        # W[b', a, c, d] = sum_L Leff[a0+a, c, L] Leff[a0+b', d, L]: product rows (a,c), columns (b',d); stored b'-major
        # by the index-mapped epilogue so that the diagonal block (b' < n_a) is a contiguous row range
        # (the weight 1/2 of the diagonal block rides on this product: columns b' < n_a with alpha = 0.5)
        W = be.empty(n_b * n_a * v * v)
        rows_a = LA.buf[a0 * v * L:]
        be.gemm_nt_mapped(rows_a, rows_a, W, n_a * v, n_a * v, L, (v, v * v, v), (v, n_a * v * v, 1), 0.5, 0.0)
        if n_b > n_a:
            be.gemm_nt_mapped(rows_a, LA.buf[a1 * v * L:], W[n_a * n_a * v * v:], n_a * v, (n_b - n_a) * v, L,
                              (v, v * v, v), (v, n_a * v * v, 1), 1.0, 0.0)
        eng.flops += 2.0 * n_a * v * n_b * v * L
        # R[a0+a, a0+b', i, j] = sum_cd W[(b',a),(c,d)] tau[(i,j),(c,d)]: one product over all rows (b', a)
        base = (a0 * v + a0) * o2
        be.gemm_nt_mapped(W, tau.buf, R.buf[base:], n_b * n_a, o2, v * v, (n_a, o2, v * o2), (BIG, 0, 1), 1.0, 0.0)
        eng.flops += 2.0 * n_b * n_a * o2 * v * v
        eng.gemms += 3
        del W
    if ring is not None and halve:
        ring(R)
    if world > 1:
        be.all_reduce(R.buf, group)
    eng.add(t2new, R, 1.0, (2, 3, 0, 1))
    if halve:
        eng.add(t2new, R, 1.0, (3, 2, 1, 0))
    del R
    if ring is not None and not halve:
        R2 = eng.zeros((v, v) + tau.shape[:2])
        ring(R2)
        be.all_reduce(R2.buf, group)
        eng.add(t2new, R2, 1.0, (2, 3, 0, 1))
        eng.add(t2new, R2, 1.0, (3, 2, 1, 0))
        del R2
    X = E("kcld,ijcd->klij", eris.ovov, tau)                      # rank-o^2 correction: - t1 t1 X
    Y = E("ka,klij->alij", t1, X)
    E("lb,alij->ijab", t1, Y, alpha=-1.0, out=t2new, beta=1.0)
    return t2new


def _intermediates(eng, t1, t2, tau, eris, f, cc2=False, ring_gemms=True):
    """Foo, Fvv, Fov (rintermediates.py:29-59), Loo, Lvv (:63-79), Woooo, Wvoov, Wvovo (:83-126).
    `f` holds the fock blocks as device tensors (fov, foo, fvv)."""
    E = eng.einsum
    ovov, ovoo, ovvv = eris.ovov, eris.ovoo, eris.ovvv
    # 2 (kc|ld) - (kd|lc): every "2*X - X'" pair of the kappa intermediates contracts this combination
    lov = eng.scale(ovov, 2.0)
    eng.add(lov, ovov, -1.0, (0, 3, 2, 1))
    dFov = E("kcld,ld->kc", lov, t1)                              # Fov - fov                      (:55-57)
    Fov = eng.copy(dFov)
    eng.add(Fov, f["fov"])
    Foo = E("kcld,ilcd->ki", lov, t2)                             # (:33-34)
    E("kc,ic->ki", dFov, t1, out=Foo, beta=1.0)                   # (:35-36)
    eng.add(Foo, f["foo"])
    Fvv = E("kcld,klad->ac", lov, t2, alpha=-1.0)                 # (:44-45)
    E("kc,ka->ac", dFov, t1, alpha=-1.0, out=Fvv, beta=1.0)       # (:46-47)
    eng.add(Fvv, f["fvv"])
    if cc2:                      # the CC2 doubles equation (:115-133) uses none of the kappa / W intermediates
        return dict(Foo=Foo, Fvv=Fvv, Fov=Fov)
    Loo = eng.copy(Foo)                                           # (:63-70)
    E("kc,ic->ki", f["fov"], t1, out=Loo, beta=1.0)
    E("lcki,lc->ki", ovoo, t1, alpha=2.0, out=Loo, beta=1.0)
    E("kcli,lc->ki", ovoo, t1, alpha=-1.0, out=Loo, beta=1.0)
    Lvv = eng.copy(Fvv)                                           # (:72-79)
    E("kc,ka->ac", f["fov"], t1, alpha=-1.0, out=Lvv, beta=1.0)
    E("kdac,kd->ac", ovvv, t1, alpha=2.0, out=Lvv, beta=1.0)
    E("kcad,kd->ac", ovvv, t1, alpha=-1.0, out=Lvv, beta=1.0)
    # hole-hole ladder intermediate (:83-91); the t2 and t1*t1 terms contract the same integrals -> tau
    Woooo = eng.transpose(eris.oooo, (0, 2, 1, 3))
    E("lcki,jc->klij", ovoo, t1, out=Woooo, beta=1.0)
    E("kclj,ic->klij", ovoo, t1, out=Woooo, beta=1.0)
    E("kcld,ijcd->klij", ovov, tau, out=Woooo, beta=1.0)
    # ring / exchange intermediates (:101-126)
    Wvoov = eng.transpose(eris.ovvo, (2, 0, 3, 1))
    E("kcad,id->akic", ovvv, t1, out=Wvoov, beta=1.0)
    E("kcli,la->akic", ovoo, t1, alpha=-1.0, out=Wvoov, beta=1.0)
    if ring_gemms:
        E("ldkc,ilda->akic", ovov, t2, alpha=-0.5, out=Wvoov, beta=1.0)
        # (:113) -0.5 (lc|kd) t2[i,l,a,d]  and  (:117) + (ld|kc) t2[i,l,a,d]  contract the same amplitudes: one GEMM
        # with (ld|kc) - 0.5 (lc|kd) = 0.5 * lov[l,d,k,c]
        E("ldkc,ilad->akic", lov, t2, alpha=0.5, out=Wvoov, beta=1.0)
    y = E("ldkc,id->lkci", ovov, t1)
    E("lkci,la->akic", y, t1, alpha=-1.0, out=Wvoov, beta=1.0)
    Wvovo = eng.transpose(eris.oovv, (2, 0, 3, 1))
    E("kdac,id->akci", ovvv, t1, out=Wvovo, beta=1.0)
    E("lcki,la->akci", ovoo, t1, alpha=-1.0, out=Wvovo, beta=1.0)
    if ring_gemms:
        E("lckd,ilda->akci", ovov, t2, alpha=-0.5, out=Wvovo, beta=1.0)
    y = E("lckd,id->lcki", ovov, t1)
    E("lcki,la->akci", y, t1, alpha=-1.0, out=Wvovo, beta=1.0)
    return dict(Foo=Foo, Fvv=Fvv, Fov=Fov, Loo=Loo, Lvv=Lvv, Woooo=Woooo, Wvoov=Wvoov, Wvovo=Wvovo, lov=lov)


def _ring_terms_into(eng, eris, t2, im, R, rank, world):
    """This rank's share of the ring / exchange terms (:143-149), accumulated into R[a,b,i,j] for its slab of the
    virtual index a (the ROW index of Wvoov / Wvovo): the three o^3 v^3 products that complete the intermediates for
    the slab (``im`` was built with ``ring_gemms=False``) and the three that contract them with t2. With
    tmp[i,j,a,b] -> R[a,b,i,j] the symmetrisation tmp + tmp^T(1,0,3,2) of :142-155 is the R^T(2,3,0,1) + R^T(3,2,1,0)
    of the ladder; the third term enters in its transposed form so that its row index is the slab index as well:
        - sum_kc Wvovo[b,k,c,i] t2[k,j,a,c]  (+ transpose)  ==  - sum_kc Wvovo[a,k,c,j] t2[k,i,b,c]  (+ transpose)."""
    E = eng.einsum
    v = eris.nvir
    a0, a1 = v * rank // world, v * (rank + 1) // world
    if a1 <= a0:
        return
    # t2 with the slab index leading, so that the slab is a row range: t2[i,l,d,a] -> [a,i,l,d] and t2[i,l,a,d] -> [a,i,l,d]
    tA = _rows(eng.transpose(t2, (3, 0, 1, 2)), a0, a1)
    tB = _rows(eng.transpose(t2, (2, 0, 1, 3)), a0, a1)
    Wvoov, Wvovo = _rows(im["Wvoov"], a0, a1), _rows(im["Wvovo"], a0, a1)
    E("aild,ldkc->akic", tA, eris.ovov, alpha=-0.5, out=Wvoov, beta=1.0)          # (:113)
    E("aild,ldkc->akic", tB, im["lov"], alpha=0.5, out=Wvoov, beta=1.0)           # (:113) + (:117)
    E("aild,lckd->akci", tA, eris.ovov, alpha=-0.5, out=Wvovo, beta=1.0)          # (:124)
    del tA, tB
    ring = eng.scale(Wvoov, 2.0)
    eng.add(ring, Wvovo, -1.0, (0, 1, 3, 2))
    Rs = _rows(R, a0, a1)
    E("akic,kjcb->abij", ring, t2, out=Rs, beta=1.0)                              # ring (:143-145)
    del ring
    E("akic,kjbc->abij", Wvoov, t2, alpha=-1.0, out=Rs, beta=1.0)                 # exchange (:146-147)
    E("akcj,kibc->abij", Wvovo, t2, alpha=-1.0, out=Rs, beta=1.0)                 # exchange (:148-149), transposed form


def _fock_blocks(eng, eris, remove_diag=False):
    o = eris.nocc
    fock = eris.fock
    foo, fvv, fov = fock[:o, :o].copy(), fock[o:, o:].copy(), fock[:o, o:].copy()
    f = {"fov": eng.tensor(fov), "foo": eng.tensor(foo), "fvv": eng.tensor(fvv)}
    f["foo_diag"] = eng.tensor(np.diag(np.diag(foo)))
    f["fvv_diag"] = eng.tensor(np.diag(np.diag(fvv)))
    return f


def make_tau(eng, t1, t2):
    """tau[i,j,a,b] = t2[i,j,a,b] + t1[i,a] t1[j,b]   (:136)."""
    tau = eng.copy(t2)
    eng.einsum("ia,jb->ijab", t1, t1, out=tau, beta=1.0)
    return tau


def update_amps(cc, t1, t2, eris, act_hole=(), act_particle=(), idx_singles=(), idx_doubles=()):
    """One amplitude update with active-space overwrite — ``rmpccsd_slow.update_amps`` (:73-180): the full RCCSD
    doubles equation (:134-149) or, with ``cc.cc2``, the CC2 one (:115-133). ``t1``/``t2`` are :class:`DTensor`
    (o,v)/(o,o,v,v) or numpy arrays (uploaded; the result then comes back as numpy). ``eris`` is a :class:`DFERIs`.
    ``cc.group`` (optional torch.distributed group): the particle-particle ladder is sharded over its ranks."""
    cc2 = bool(getattr(cc, "cc2", False))
    group = getattr(cc, "group", None)
    eng = eris.eng
    as_numpy = not isinstance(t1, DTensor)
    if as_numpy:
        t1, t2 = eng.tensor(t1), eng.tensor(t2)
    E = eng.einsum
    o, v = eris.nocc, eris.nvir
    f = _fock_blocks(eng, eris)
    ovvv, ooov, ovvo, oovv, ovov = eris.ovvv, eris.ooov, eris.ovvo, eris.oovv, eris.ovov
    tau = make_tau(eng, t1, t2)
    rank, world = (0, 1) if group is None else eng.be.rank_and_size(group)
    shard_ring = world > 1 and not cc2
    im = _intermediates(eng, t1, t2, tau, eris, f, cc2=cc2, ring_gemms=not shard_ring)
    Foo, Fvv, Fov = im["Foo"], im["Fvv"], im["Fov"]
    # move energy terms to the other side (:90-91, :135)
    shifted = ((Foo, "foo_diag"), (Fvv, "fvv_diag"))
    if not cc2:
        shifted += ((im["Loo"], "foo_diag"), (im["Lvv"], "fvv_diag"))
    for x, d in shifted:
        eng.add(x, f[d], -1.0)

    # ---- T1 equation (:93-111) ----
    t1new = eng.copy(f["fov"])
    z = E("kc,ic->ki", f["fov"], t1)
    E("ki,ka->ia", z, t1, alpha=-2.0, out=t1new, beta=1.0)
    E("ac,ic->ia", Fvv, t1, out=t1new, beta=1.0)
    E("ki,ka->ia", Foo, t1, alpha=-1.0, out=t1new, beta=1.0)
    # The "2 x - x'" pairs of :99-100, :105-106, :109-110 each contract the same integrals with t2 in two index
    # orders: one product with 2 t2 - t2' instead of two with t2 (no symmetry of t2 is assumed: the class overwrite
    # of a previous update may have broken t2[ijab] = t2[jiba]).
    theta_o = eng.scale(t2, 2.0)                                  # 2 t2[k,i,c,a] - t2[i,k,c,a]
    eng.add(theta_o, t2, -1.0, (1, 0, 2, 3))
    E("kc,kica->ia", Fov, theta_o, out=t1new, beta=1.0)
    z = E("kc,ic->ki", Fov, t1)
    E("ki,ka->ia", z, t1, out=t1new, beta=1.0)
    E("kcai,kc->ia", ovvo, t1, alpha=2.0, out=t1new, beta=1.0)
    E("kiac,kc->ia", oovv, t1, alpha=-1.0, out=t1new, beta=1.0)
    theta_v = eng.scale(t2, 2.0)                                  # 2 t2[i,k,c,d] - t2[i,k,d,c]
    eng.add(theta_v, t2, -1.0, (0, 1, 3, 2))
    E("kdac,ikcd->ia", ovvv, theta_v, out=t1new, beta=1.0)       # 2 (kd|ac) t2[ikcd] - (kc|ad) t2[ikcd], c <-> d in the second
    del theta_v
    w = E("kdac,kd->ac", ovvv, t1)
    E("ac,ic->ia", w, t1, alpha=2.0, out=t1new, beta=1.0)
    w = E("kcad,kd->ca", ovvv, t1)
    E("ca,ic->ia", w, t1, alpha=-1.0, out=t1new, beta=1.0)
    E("kilc,klac->ia", ooov, theta_o, alpha=-1.0, out=t1new, beta=1.0)  # -2 (ki|lc) t2[klac] + (li|kc) t2[klac], k <-> l
    del theta_o
    u = E("kilc,lc->ki", ooov, t1)
    E("ki,ka->ia", u, t1, alpha=-2.0, out=t1new, beta=1.0)
    u = E("likc,lc->ik", ooov, t1)
    E("ik,ka->ia", u, t1, out=t1new, beta=1.0)

    # ---- T2 equation (:114-162) ----
    t2new = eng.transpose(ovov, (0, 2, 1, 3))
    if cc2:
        # CC2 (:115-133): both ladders act on t1 (x) t1 only — Woooo2 / Wvvvv of :116-126 are cc_Woooo / cc_Wvvvv
        # with t2 = 0, and 'klij,ka,lb->ijab' / 'abcd,ic,jd->ijab' contract them with tau1[i,j,a,b] = t1[i,a] t1[j,b]
        tau1 = E("ia,jb->ijab", t1, t1)
        Woooo2 = eng.transpose(eris.oooo, (0, 2, 1, 3))                   # (:116)
        E("lcki,jc->klij", eris.ovoo, t1, out=Woooo2, beta=1.0)           # (:117)  ooov[k,i,l,c] = ovoo[l,c,k,i]
        E("kclj,ic->klij", eris.ovoo, t1, out=Woooo2, beta=1.0)           # (:118)
        E("kcld,ijcd->klij", ovov, tau1, out=Woooo2, beta=1.0)            # (:119)
        E("klij,klab->ijab", Woooo2, tau1, out=t2new, beta=1.0)           # (:120)
        _pp_ladder(eng, eris, t1, tau1, t2new, group=group)               # (:121-125)
        Lvv2 = E("kc,ka->ac", f["fov"], t1, alpha=-1.0)                   # (:126-127)
        eng.add(Lvv2, f["fvv"])
        eng.add(Lvv2, f["fvv_diag"], -1.0)
        _sym_add(eng, t2new, E("ac,ijcb->ijab", Lvv2, t2), 1.0)           # (:128-129)
        Loo2 = E("kc,ic->ki", f["fov"], t1)                               # (:130-131)
        eng.add(Loo2, f["foo"])
        eng.add(Loo2, f["foo_diag"], -1.0)
        _sym_add(eng, t2new, E("ki,kjab->ijab", Loo2, t2), -1.0)          # (:132-133)
    else:
        E("klij,klab->ijab", im["Woooo"], tau, out=t2new, beta=1.0)          # hole-hole ladder (:137)
        # particle-particle ladder (:138); several ranks: the ring / exchange terms ride in the same accumulator
        _pp_ladder(eng, eris, t1, tau, t2new, group=group, halve=_pair_symmetric(eng, t2),
                   ring=(lambda R: _ring_terms_into(eng, eris, t2, im, R, rank, world)) if shard_ring else None)
        _sym_add(eng, t2new, E("ac,ijcb->ijab", im["Lvv"], t2), 1.0)          # (:139-140)
        _sym_add(eng, t2new, E("ki,kjab->ijab", im["Loo"], t2), -1.0)         # (:141-142)
    if not cc2 and not shard_ring:
        ring = eng.scale(im["Wvoov"], 2.0)                                    # ring (:143-145): 2 Wvoov[a,k,i,c] -
        eng.add(ring, im["Wvovo"], -1.0, (0, 1, 3, 2))                        # Wvovo[a,k,c,i] share t2[k,j,c,b]: one GEMM
        _sym_add(eng, t2new, E("akic,kjcb->ijab", ring, t2), 1.0)
        del ring
        _sym_add(eng, t2new, E("akic,kjbc->ijab", im["Wvoov"], t2), -1.0)     # exchange (:146-149)
        _sym_add(eng, t2new, E("bkci,kjac->ijab", im["Wvovo"], t2), -1.0)
    tmp2 = eng.transpose(ovvv, (1, 3, 0, 2))                              # (:151-154)
    E("kibc,ka->abic", oovv, t1, alpha=-1.0, out=tmp2, beta=1.0)
    _sym_add(eng, t2new, E("abic,jc->ijab", tmp2, t1), 1.0)
    tmp2 = eng.transpose(ooov, (3, 1, 2, 0))                              # (:155-158)
    E("kcai,jc->akij", ovvo, t1, out=tmp2, beta=1.0)
    _sym_add(eng, t2new, E("akij,kb->ijab", tmp2, t1), -1.0)

    eng.div_denominators(t1new)                                           # (:160-164)
    eng.div_denominators(t2new)

    # ---- active-space overwrite (:166-178) ----
    ms, md = class_masks(idx_singles, idx_doubles)
    if ms or md:
        ah = np.zeros(o, dtype=np.int32); ah[np.asarray(act_hole, dtype=int)] = 1
        ap = np.zeros(v, dtype=np.int32); ap[np.asarray(act_particle, dtype=int)] = 1
        ah_d, ap_d = eng.be.upload_int(ah), eng.be.upload_int(ap)
        eng.class_overwrite(t1new, t1, ah_d, ap_d, ms)
        eng.class_overwrite(t2new, t2, ah_d, ap_d, md)
    if as_numpy:
        return eng.numpy(t1new), eng.numpy(t2new)
    return t1new, t2new


def energy(cc, t1, t2, eris):
    """``rmpccsd_slow.energy`` (:183-194): 2 f_ia t1_ia + sum tau_ijab (2 (ia|jb) - (ib|ja))."""
    eng = eris.eng
    if not isinstance(t1, DTensor):
        t1, t2 = eng.tensor(t1), eng.tensor(t2)
    o = eris.nocc
    e = 2.0 * eng.dot(eng.tensor(eris.fock[:o, o:]), t1)
    tau = make_tau(eng, t1, t2)
    e += 2.0 * eng.dot(tau, eng.transpose(eris.ovov, (0, 2, 1, 3)))
    e -= eng.dot(tau, eng.transpose(eris.ovov, (0, 2, 3, 1)))
    return e


def init_amps(cc, eris):
    """MP2 start (:206-219): t1 = f_ia / e_ia, t2 = (ia|jb) / e_ijab; returns (emp2, t1, t2) on the device."""
    eng = eris.eng
    o = eris.nocc
    t1 = eng.div_denominators(eng.tensor(eris.fock[:o, o:]))
    t2 = eng.div_denominators(eng.transpose(eris.ovov, (0, 2, 1, 3)))
    emp2 = 2.0 * eng.dot(t2, eng.transpose(eris.ovov, (0, 2, 1, 3))) - eng.dot(t2, eng.transpose(eris.ovov, (0, 2, 3, 1)))
    return emp2, t1, t2


# --------------------------------------------------------------------------------------------------------------
# driver loop: pyscf/cc/ccsd.py:42-150 (``kernel``) as the fork runs it for ``RMPCCSD.ccsd`` (rmpccsd_slow.py:223-258)
# --------------------------------------------------------------------------------------------------------------
def ccsd_kernel(mycc, eris, t1=None, t2=None, max_cycle=50, tol=1e-7, tolnormt=1e-5, act_hole=(), act_particle=(),
                idx_s=(), idx_d=(), log=None):
    """The amplitude iteration with every vector resident on the device (``ccsd.kernel``, ccsd.py:42-150, the
    ``act_particle is not None and not oo_mp2 and not pert_triples`` path):

        eccsd = energy(t1, t2);  each cycle:  t1new, t2new = update_amps(...)
        normt = |vec(t1new, t2new) - vec(t1, t2)|                    (:99-101)
        t1, t2 = run_diis(t1new, t2new, istep, normt, eccsd - eold, adiis)   (:121, ccsd.py:1269-1276)
        eold, eccsd = eccsd, energy(t1, t2);  stop when |dE| < tol and normt < tolnormt   (:130-137)

    ``iterative_damping`` < 1 mixes the old amplitudes in as :107-119. Returns (converged, e_corr, t1, t2) with
    device tensors. Only scalars (energies, norms, DIIS dots) cross to the host."""
    from .diis import VectorDIIS
    eng = eris.eng
    o, v = eris.nocc, eris.nvir
    if t1 is None or t2 is None:
        guess = mycc.init_amps(eris) if callable(getattr(mycc, "init_amps", None)) else init_amps(mycc, eris)
        g1, g2 = guess[1:]                                      # get_init_guess = init_amps()[1:] (ccsd.py)
        t1 = g1 if t1 is None else t1
        t2 = g2 if t2 is None else t2
    if not isinstance(t1, DTensor):
        t1 = eng.tensor(t1)
    if not isinstance(t2, DTensor):
        t2 = eng.tensor(t2)
    eold = 0.0
    eccsd = energy(mycc, t1, t2, eris)
    diis = getattr(mycc, "diis", True)
    adiis = None
    if isinstance(diis, VectorDIIS):
        adiis = diis
    elif diis:
        adiis = VectorDIIS(eng.be, space=getattr(mycc, "diis_space", 6))
    damping = float(getattr(mycc, "iterative_damping", 1.0))
    start_cycle = getattr(mycc, "diis_start_cycle", 0)
    start_de = getattr(mycc, "diis_start_energy_diff", 1e9)
    n = o * v + (o * v) * (o * v + 1) // 2
    converged = False
    mycc.cycles = 0
    history = []
    vec_old = eng.amplitudes_to_vector(t1, t2)
    for istep in range(max_cycle):
        t1new, t2new = update_amps(mycc, t1, t2, eris, act_hole, act_particle, idx_s, idx_d)
        vec_new = eng.amplitudes_to_vector(t1new, t2new)
        diff = DTensor(eng.be.copy(vec_new), (n,))
        eng.add(diff, DTensor(vec_old, (n,)), -1.0)
        normt = float(np.sqrt(eng.dot(diff, diff)))
        del diff
        if damping < 1.0:                                        # (:107-119)
            mixed = eng.scale(t1new, damping)
            eng.add(mixed, t1, 1.0 - damping)
            t1new = mixed
            mixed = eng.scale(t2new, damping)
            eng.add(mixed, t2, 1.0 - damping)
            t2new = mixed
            vec_new = eng.amplitudes_to_vector(t1new, t2new)
        t1, t2 = t1new, t2new
        if adiis is not None and istep >= start_cycle and abs(eccsd - eold) < start_de:   # run_diis (ccsd.py:1269-1276)
            vec_new = adiis.update(vec_new, n)
            t1, t2 = eng.vector_to_amplitudes(vec_new, o, v)
        vec_old = vec_new
        eold, eccsd = eccsd, energy(mycc, t1, t2, eris)
        mycc.cycles = istep + 1
        history.append((istep + 1, eccsd, eccsd - eold, normt))
        if log is not None:
            log(f"cycle = {istep + 1}  E_corr(RMPCCSD) = {eccsd:.15g}  dE = {eccsd - eold:.9g}  norm(t1,t2) = {normt:.6g}")
        if abs(eccsd - eold) < tol and normt < tolnormt:
            converged = True
            break
    mycc.history = history
    return converged, eccsd, t1, t2


class RMPCCSD:
    """Device counterpart of ``pyscf.cc.rmpccsd_slow.RMPCCSD`` (:197-271): the high-level solver object whose
    ``kernel(act_hole, act_particle, idx_s, idx_d, ...)`` runs MBPT2, CC2 or CCSD with the active-space overwrite.
    Built on a low-level solver (``MPCC_LL``, whose DF tensors provide the integrals) instead of ``ccsd.CCSD``'s
    AO integrals; option names and defaults are ``ccsd.CCSD``'s (ccsd.py:970-984)."""

    max_cycle = 50
    conv_tol = 1e-7
    conv_tol_normt = 1e-5
    iterative_damping = 1.0
    diis = True
    diis_space = 6
    diis_start_cycle = 0
    diis_start_energy_diff = 1e9

    def __init__(self, lowlevel=None, eris=None, fock=None, group=None, verbose=False, stdout=None):
        if eris is None:
            if lowlevel is None:
                raise ValueError("RMPCCSD needs a low-level solver (MPCC_LL) or ready-made DFERIs")
            eris = lowlevel.high_level_eris(fock=fock)
        self.lowlevel = lowlevel
        self.eris = eris
        self.nocc, self.nvir = eris.nocc, eris.nvir
        self.nmo = self.nocc + self.nvir
        self.group = group
        self.cc2 = False
        self.max_space = 20                       # (:201) kept for attribute parity; EOM is not part of this path
        self.verbose = verbose
        self.stdout = stdout
        self.converged = False
        self.e_corr = None
        self.emp2 = None
        self.t1 = self.t2 = None
        self.cycles = 0

    def _log(self, msg):
        if self.verbose:
            import sys
            print(msg, file=self.stdout or sys.stdout)

    def ao2mo(self, mo_coeff=None):
        return self.eris

    def init_amps(self, eris=None):
        self.emp2, t1, t2 = init_amps(self, eris or self.eris)
        self._log(f"Init t2, MP2 energy = {self.emp2:.15g}")
        return self.emp2, t1, t2

    def get_init_guess(self, eris=None):
        return self.init_amps(eris)[1:]

    def energy(self, t1=None, t2=None, eris=None):
        return energy(self, self.t1 if t1 is None else t1, self.t2 if t2 is None else t2, eris or self.eris)

    def update_amps(self, t1, t2, eris, act_hole=(), act_particle=(), idx_s=(), idx_d=()):
        return update_amps(self, t1, t2, eris, act_hole, act_particle, idx_s, idx_d)

    def amplitudes_to_vector(self, t1, t2, out=None):
        eng = self.eris.eng
        t1 = t1 if isinstance(t1, DTensor) else eng.tensor(t1)
        t2 = t2 if isinstance(t2, DTensor) else eng.tensor(t2)
        return eng.amplitudes_to_vector(t1, t2)

    def vector_to_amplitudes(self, vec, nmo=None, nocc=None):
        nocc = self.nocc if nocc is None else nocc
        nmo = self.nmo if nmo is None else nmo
        return self.eris.eng.vector_to_amplitudes(vec, nocc, nmo - nocc)

    def vector_size(self, nmo=None, nocc=None):
        nocc = self.nocc if nocc is None else nocc
        nov = nocc * ((self.nmo if nmo is None else nmo) - nocc)
        return nov + nov * (nov + 1) // 2

    def kernel(self, act_hole, act_particle, idx_s, idx_d, t1=None, t2=None, eris=None, mbpt2=False, cc2=False):
        return self.ccsd(act_hole, act_particle, idx_s, idx_d, t1, t2, eris, mbpt2, cc2)

    def ccsd(self, act_hole, act_particle, idx_s, idx_d, t1=None, t2=None, eris=None, mbpt2=False, cc2=False):
        """(:223-258) MBPT2 (one shot), CC2 or CCSD ground state with the amplitude classes ``idx_s`` / ``idx_d``
        frozen at their input values. Returns (e_corr, t1, t2); amplitudes stay on the device (``eng.numpy``)."""
        if mbpt2 and cc2:
            raise RuntimeError("MBPT2 and CC2 are mutually exclusive approximations to the CCSD ground state.")
        if eris is None:
            eris = self.eris
        self.eris = eris
        if mbpt2:
            self.e_corr, self.t1, self.t2 = self.init_amps(eris)
            return self.e_corr, self.t1, self.t2
        self.cc2 = bool(cc2)
        cctyp = "CC2" if cc2 else "CCSD"
        self.converged, self.e_corr, self.t1, self.t2 = ccsd_kernel(
            self, eris, t1, t2, max_cycle=self.max_cycle, tol=self.conv_tol, tolnormt=self.conv_tol_normt,
            act_hole=act_hole, act_particle=act_particle, idx_s=idx_s, idx_d=idx_d, log=self._log)
        self._log(f"{cctyp} converged" if self.converged else f"{cctyp} not converged")
        self._log(f"E_corr = {self.e_corr:.16g}")
        return self.e_corr, self.t1, self.t2
