"""Unrestricted DF low-level residual on the GPU.

Counterpart of ``updated_amp`` in ``pyscf/cc/umpcc_lowlevel.py:174-357`` — the alpha/beta version of the
restricted iteration (T1 dressing per spin, same-spin doubles antisymmetrised, opposite-spin doubles plain,
``Table I of Mester, Nagy, Kallay, JCP 146, 194102``). The reference class around it cannot be instantiated
(``mpccLL.__init__`` reads an undefined global ``mol``, :393) and the factory rejects UHF
(``pyscf/mpcc/__init__.py:4-5``), so the drop-in unit is the function: it reads ``t1, Loo, Lov, Lvv, LOO, LOV, LVV,
foo, fOO, fvv, fVV`` (and the orbital energies behind ``fock``/``D``) from a state object and returns
``(res, dE, t1, [t2aa, t2ab, t2bb])``. Details kept as written there: ``Jvv`` carries a MINUS sign (:314; the
restricted file has +), ``Fov`` uses the bare ``Lov`` (:239), dE = (dEa + dEb)/2 (:346), res is the SUM of the
two spin norms (:348-349), t2 blocks are [a,i,b,j] ordered.

Every contraction runs through :class:`pyscf_mpcc_b200.tensor_engine.Engine` (permute -> FP64 DMMA GEMM ->
permute/accumulate); denominators through ``mpccb_div_general``.

The rest of the prototype's module-level functions is here as well:
``amplitudes_to_vector`` / ``vector_to_amplitudes`` (:15-46, on top of ``ccsd.amplitudes_to_vector_s4`` /
``vector_to_amplitudes_s4``, ``pyscf/cc/ccsd.py:705-757`` -> device gather / scatter ``mpccb_take_2d`` /
``mpccb_takebak_2d``), ``run_diis`` (:92-98, DIIS on the packed device vector), ``kernel`` (:48-89) and
``compute_CC2_energy`` (:100-127 == ``pyscf/cc/rmpcc_lowlevel.py:82-109``).
"""
from __future__ import annotations

import numpy as np

from .tensor_engine import DTensor, Engine


class UMPCCState:
    """What ``updated_amp`` reads from ``myll`` (:179-196), as device tensors. Lower case = alpha, upper = beta."""

    def __init__(self, eng: Engine, alpha, beta):
        """alpha / beta: dict(Loo=, Lov=, Lvv=, foo=, fvv=, eo=, ev=) of numpy arrays (MO basis of that spin)."""
        self.eng = eng
        T = eng.tensor
        for up, blk in ((False, alpha), (True, beta)):
            for key in ("Loo", "Lov", "Lvv"):
                name = key[0] + (key[1:].upper() if up else key[1:])
                setattr(self, name, T(blk[key]))
            setattr(self, "fOO" if up else "foo", T(blk["foo"]))
            setattr(self, "fVV" if up else "fvv", T(blk["fvv"]))
        self.eps = [(T(alpha["eo"]), T(alpha["ev"])), (T(beta["eo"]), T(beta["ev"]))]
        self.nocca, self.nvira = alpha["Lov"].shape[1:]
        self.noccb, self.nvirb = beta["Lov"].shape[1:]
        self.naux = alpha["Lov"].shape[0]
        self.t1 = [eng.zeros((self.nocca, self.nvira)), eng.zeros((self.noccb, self.nvirb))]
        self.t2 = None
        self.diis = True                                   # mpccLL.__init__ (:379)
        self.nmo = self.nocca + self.nvira                 # (:391)
        self.nocc = (self.nocca, self.noccb)               # what ump2's get_nocc gives; read by run_diis (:95)

    def set_t1(self, t1a, t1b):
        self.t1 = [self.eng.tensor(t1a), self.eng.tensor(t1b)]


def _dress(eng, Loo, Lov, Lvv, t1):
    """Steps 1-2 for one spin (:198-214)."""
    E = eng.einsum
    Xvo = E("Lab,ib->Lai", Lvv, t1)
    Xoo = E("Lia,ja->Lij", Lov, t1)
    Joo = eng.copy(Xoo)
    eng.add(Joo, Loo)
    Om = E("Laj,Lji->ai", Xvo, Joo, alpha=-1.0)
    Jvo = eng.copy(Xvo)
    eng.add(Jvo, Lov, 1.0, (0, 2, 1))
    E("Lij,ja->Lai", Joo, t1, alpha=-1.0, out=Jvo, beta=1.0)
    Xs = E("Lia,ia->L", Lov, t1)
    return dict(Xvo=Xvo, Xoo=Xoo, Joo=Joo, Om=Om, Jvo=Jvo, Xs=Xs)


def _singles_from_x(eng, d, Lov, t1, X, fvv, foo):
    """Step 4 + Fock terms + Fov of step 5 for one spin (:222-239)."""
    E = eng.einsum
    tmp = E("Ljk,ka->Lja", d["Xoo"], t1)
    E("Lja,Lji->ai", tmp, d["Joo"], out=d["Om"], beta=1.0)
    E("Lai,L->ai", d["Jvo"], X, out=d["Om"], beta=1.0)
    E("ib,ba->ai", t1, fvv, out=d["Om"], beta=1.0)
    E("ka,ik->ai", t1, foo, alpha=-1.0, out=d["Om"], beta=1.0)
    Fov = E("Ljb,L->jb", Lov, X)
    E("Lij,Lib->jb", d["Xoo"], Lov, alpha=-1.0, out=Fov, beta=1.0)
    d["XT"] = tmp
    return Fov


def updated_amp(myll: UMPCCState):
    """One unrestricted residual evaluation from ``myll.t1``; updates ``myll.t1`` in place like the reference
    (:351-352) and returns (res, dE, t1, [t2aa, t2ab, t2bb]) with device tensors."""
    eng = myll.eng
    E = eng.einsum
    t1a, t1b = myll.t1
    (eoa, eva), (eob, evb) = myll.eps
    a = _dress(eng, myll.Loo, myll.Lov, myll.Lvv, t1a)
    b = _dress(eng, myll.LOO, myll.LOV, myll.LVV, t1b)
    X = eng.copy(a["Xs"])                                            # step 3 (:217-219)
    eng.add(X, b["Xs"])
    Fov = _singles_from_x(eng, a, myll.Lov, t1a, X, myll.fvv, myll.foo)
    FOV = _singles_from_x(eng, b, myll.LOV, t1b, X, myll.fVV, myll.fOO)
    # steps 6-7 (:242-249): dressed integrals and doubles, [a,i,b,j]
    t2aa = E("Lai,Lbj->aibj", a["Jvo"], a["Jvo"])
    eng.add(t2aa, eng.copy(t2aa), -1.0, (0, 3, 2, 1))
    eng.div_general(t2aa, (eva, eoa, eva, eoa), (-1.0, 1.0, -1.0, 1.0))
    t2bb = E("Lai,Lbj->aibj", b["Jvo"], b["Jvo"])
    eng.add(t2bb, eng.copy(t2bb), -1.0, (0, 3, 2, 1))
    eng.div_general(t2bb, (evb, eob, evb, eob), (-1.0, 1.0, -1.0, 1.0))
    t2ab = E("Lai,Lbj->aibj", a["Jvo"], b["Jvo"])
    eng.div_general(t2ab, (eva, eoa, evb, eob), (-1.0, 1.0, -1.0, 1.0))
    # doubles -> Y (:299-303)
    Yvo = E("aibj,Ljb->Lai", t2aa, myll.Lov)
    E("aiBJ,LJB->Lai", t2ab, myll.LOV, out=Yvo, beta=1.0)
    YVO = E("aibj,Ljb->Lai", t2bb, myll.LOV)
    E("aiBJ,Lia->LBJ", t2ab, myll.Lov, out=YVO, beta=1.0)
    # doubles -> singles through Fov (:305-309)
    E("aibj,jb->ai", t2aa, Fov, out=a["Om"], beta=1.0)
    E("aiBJ,JB->ai", t2ab, FOV, out=a["Om"], beta=1.0)
    E("aibj,jb->ai", t2bb, FOV, out=b["Om"], beta=1.0)
    E("aiBJ,ia->BJ", t2ab, Fov, out=b["Om"], beta=1.0)
    dE = 0.0
    for d, Lov, Lvv, t1, Y in ((a, myll.Lov, myll.Lvv, t1a, Yvo), (b, myll.LOV, myll.LVV, t1b, YVO)):
        Jvv = eng.copy(Lvv)                                          # step 8 (:312-320), MINUS sign as written
        E("Ljb,ja->Lba", Lov, t1, alpha=-1.0, out=Jvv, beta=1.0)
        E("Lba,Lbi->ai", Jvv, Y, out=d["Om"], beta=1.0)
        E("Lji,Laj->ai", d["Joo"], Y, alpha=-1.0, out=d["Om"], beta=1.0)
        e1 = eng.copy(d["Jvo"])                                      # step 9 (:324-328)
        E("Lij,ja->Lai", d["Xoo"], t1, out=e1, beta=1.0)
        E("L,ia->Lai", X, t1, out=e1, beta=1.0)
        dE += eng.dot(e1, Y)
    dE *= 0.5                                                        # (:346)
    res = 0.0
    new_t1 = []
    for d, t1, (eo, ev) in ((a, t1a, myll.eps[0]), (b, t1b, myll.eps[1])):
        tn = eng.transpose(d["Om"], (1, 0), alpha=-1.0)              # -Omega^T / (e_a - e_i)  (:351-352)
        eng.div_general(tn, (eo, ev), (-1.0, 1.0))
        diff = eng.copy(t1)
        eng.add(diff, tn, -1.0)
        res += float(np.sqrt(eng.dot(diff, diff)))                   # (:348-349)
        new_t1.append(tn)
    myll.t1 = new_t1
    myll.t2 = [t2aa, t2ab, t2bb]
    return res, dE, myll.t1, myll.t2


# ---------------------------------------------------------------------------------------------------------------
# vector layout of the prototype's DIIS (:15-46)
# ---------------------------------------------------------------------------------------------------------------
def _tril_rows(eng, n):
    """Device int32 arrays of p*n+q and q*n+p over the strict lower triangle p > q (numpy.tril_indices(n, k=-1) order),
    cached on the engine."""
    cache = eng.__dict__.setdefault("_tril_index_cache", {})
    if n not in cache:
        p, q = np.tril_indices(n, k=-1)
        cache[n] = (eng.be.upload_int(p * n + q), eng.be.upload_int(q * n + p), len(p))
    return cache[n]


def _s4_size(nocc, nvir):
    return nocc * nvir + nocc * (nocc - 1) // 2 * nvir * (nvir - 1) // 2


def _pack_s4(eng, t1, t2, out):
    """ccsd.amplitudes_to_vector_s4 (ccsd.py:738-748) into the flat device buffer ``out``: t1, then the nocc^2 x nvir^2
    VIEW of t2 at rows i>j, columns a>b. Like the reference's ``t2.reshape(nocc**2, nvir**2)`` this reinterprets whatever
    4-index layout t2 has (the prototype passes its [a,i,b,j] blocks, :245-249)."""
    nocc, nvir = t1.shape
    nov = nocc * nvir
    eng.be.permute_axpby(t1.buf, out, t1.shape, (0, 1), 1.0, 0.0)
    if nocc > 1 and nvir > 1:
        orow, _, _ = _tril_rows(eng, nocc)
        vcol, _, _ = _tril_rows(eng, nvir)
        eng.be.take_2d(t2.buf, nvir * nvir, orow, vcol, out[nov:])


def _unpack_s4(eng, vec, nocc, nvir):
    """ccsd.vector_to_amplitudes_s4 -> _unpack_4fold (ccsd.py:705-717, 750-757): t1 copy and the antisymmetric t2."""
    nov = nocc * nvir
    t1 = DTensor(eng.be.copy(vec[:nov]), (nocc, nvir))
    t2 = eng.zeros((nocc, nocc, nvir, nvir))
    if nocc > 1 and nvir > 1:
        o_pq, o_qp, _ = _tril_rows(eng, nocc)
        v_pq, v_qp, _ = _tril_rows(eng, nvir)
        tril, ld = vec[nov:], nvir * nvir
        eng.be.takebak_2d(t2.buf, ld, tril, o_pq, v_pq, 1.0)
        eng.be.takebak_2d(t2.buf, ld, tril, o_qp, v_qp, 1.0)
        eng.be.takebak_2d(t2.buf, ld, tril, o_pq, v_qp, -1.0)
        eng.be.takebak_2d(t2.buf, ld, tril, o_qp, v_pq, -1.0)
    return t1, t2


def amplitudes_to_vector(t1, t2, out=None, *, eng):
    """[s4(t1a, t2aa) | s4(t1b, t2bb) | t2ab.ravel()] as one flat device buffer (:15-25). ``t1`` = [t1a, t1b],
    ``t2`` = [t2aa, t2ab, t2bb] device tensors; ``eng`` is the engine that owns them."""
    (oa, va), (ob, vb) = t1[0].shape, t1[1].shape
    sizea, sizeb, sizeab = _s4_size(oa, va), _s4_size(ob, vb), oa * ob * va * vb
    if t2[1].size != sizeab:
        raise ValueError("t2ab does not have nocca*noccb*nvira*nvirb elements")
    vector = eng.be.empty(sizea + sizeb + sizeab) if out is None else out
    _pack_s4(eng, t1[0], t2[0], vector[:sizea])
    _pack_s4(eng, t1[1], t2[2], vector[sizea:sizea + sizeb])
    eng.be.permute_axpby(t2[1].buf, vector[sizea + sizeb:], (sizeab,), (0,), 1.0, 0.0)
    return vector


def vector_to_amplitudes(vector, nmo, nocc, *, eng):
    """Inverse layout (:27-46): ([t1a, t1b], [t2aa (o,o,v,v) antisymmetric, t2ab reshaped (oa,ob,va,vb), t2bb])."""
    nocca, noccb = nocc
    nvira, nvirb = nmo - nocca, nmo - noccb
    sizea, sizeb = _s4_size(nocca, nvira), _s4_size(noccb, nvirb)
    n_all, n_v = nocca + noccb, nvira + nvirb
    if len(vector) == n_all * n_v + n_all * (n_all - 1) // 2 * n_v * (n_v - 1) // 2:
        raise RuntimeError("Input vector is GCCSD vecotr")                       # (:36-38), message as written
    if len(vector) != sizea + sizeb + nocca * noccb * nvira * nvirb:
        raise ValueError(f"vector has {len(vector)} elements, expected {sizea + sizeb + nocca * noccb * nvira * nvirb}")
    t1a, t2aa = _unpack_s4(eng, vector[:sizea], nocca, nvira)
    t1b, t2bb = _unpack_s4(eng, vector[sizea:sizea + sizeb], noccb, nvirb)
    t2ab = DTensor(eng.be.copy(vector[sizea + sizeb:]), (nocca, noccb, nvira, nvirb))
    return [t1a, t1b], [t2aa, t2ab, t2bb]


def run_diis(myll, t1, t2, adiis):
    """(:92-98) ``adiis`` is a :class:`pyscf_mpcc_b200.diis.VectorDIIS` on the engine's backend."""
    eng = myll.eng
    vec = amplitudes_to_vector(t1, t2, eng=eng)
    return vector_to_amplitudes(adiis.update(vec, len(vec)), myll.nmo, myll.nocc, eng=eng)


def kernel(myll, tol=1e-6, maxcount=100, verbose=True, stdout=None):
    """The iteration of the prototype's ``kernel`` (:48-89); tol / maxcount are its literals (:53-54). What precedes
    the loop there cannot be reproduced — ``compute_three_center_ints`` needs the PySCF mean field (the state already
    holds the MO three-index tensors) and ``initialize_t1`` (:130-171) runs MP2 / DF-MP2 / UHF on an undefined global
    ``mol``; what it leaves behind is ``myll.t1 = 0`` (:171), the state's initial value. The ``breakpoint()`` after
    the loop (:88) is dropped. Returns DE; ``myll.history`` holds (count, DE, res) per iteration."""
    import sys
    from .diis import VectorDIIS
    err, count, DE = np.inf, 0, None
    myll.history = []
    adiis = VectorDIIS(myll.eng.be)
    while err > tol and count < maxcount:
        res, DE, t1_new, t2_new = updated_amp(myll)
        if myll.diis:
            t1, t2 = run_diis(myll, t1_new, t2_new, adiis)
        else:
            t1, t2 = t1_new, t2_new
        myll.t1 = t1
        myll.t2 = t2
        count += 1
        err = res
        myll.history.append((count, DE, res))
        if verbose:
            print(f"It {count} Energy progress {DE:.8f} residual progress {res:.6e}", file=stdout or sys.stdout)
    myll.iterations = count
    return DE


def compute_CC2_energy(myll, t1, t2):
    """(:100-127; the same text as ``pyscf/cc/rmpcc_lowlevel.py:82-109``) one-spin energy expression
    sum_{L,a,i} e1[L,a,i] Yvo[L,a,i] from ``myll.Loo / Lov / Lvv``, t1 (o,v) and t2 in [a,i,b,j] order. The quantities
    the reference forms and never uses (Omega_vo, Fov, eris, Jvv) are not computed."""
    eng = myll.eng
    E = eng.einsum
    Loo, Lov, Lvv = myll.Loo, myll.Lov, myll.Lvv
    Xvo = E("Lab,ib->Lai", Lvv, t1)
    Xoo = E("Lia,ja->Lij", Lov, t1)
    Joo = eng.copy(Xoo)
    eng.add(Joo, Loo)
    Jvo = Xvo
    eng.add(Jvo, Lov, 1.0, (0, 2, 1))
    E("Lij,ja->Lai", Joo, t1, alpha=-1.0, out=Jvo, beta=1.0)
    X = E("Lia,ia->L", Lov, t1)
    Yvo = E("aibj,Ljb->Lai", t2, Lov)
    e1 = Jvo
    E("Lij,ja->Lai", Xoo, t1, out=e1, beta=1.0)
    E("L,ia->Lai", X, t1, out=e1, beta=1.0)
    return eng.dot(e1, Yvo)
