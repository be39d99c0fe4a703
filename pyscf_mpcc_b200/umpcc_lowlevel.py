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
