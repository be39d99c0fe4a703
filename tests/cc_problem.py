"""Shared builders for the CCSD-residual tests: a seeded DF problem, the reference-style eris namespace
(numpy arrays in chemists' notation, attribute names of ``ccsd._ChemistsERIs``) and random amplitudes."""
import types

import numpy as np


def make_problem(o, v, L, seed=0, scale=0.25, fock_offdiag=0.02):
    rng = np.random.default_rng(seed)
    n = o + v
    Lpq = rng.standard_normal((L, n, n)) * scale / np.sqrt(L)
    Lpq = Lpq + Lpq.transpose(0, 2, 1)
    eo = np.sort(rng.uniform(-2.0, -0.5, o))
    ev = np.sort(rng.uniform(0.3, 3.0, v))
    off = rng.standard_normal((n, n)) * fock_offdiag
    fock = np.diag(np.concatenate([eo, ev])) + off + off.T
    np.fill_diagonal(fock, np.concatenate([eo, ev]))
    Loo = np.ascontiguousarray(Lpq[:, :o, :o])
    Lov = np.ascontiguousarray(Lpq[:, :o, o:])
    Lvv = np.ascontiguousarray(Lpq[:, o:, o:])
    t1 = 0.05 * rng.standard_normal((o, v))
    t2 = 0.05 * rng.standard_normal((o, o, v, v))
    t2 = t2 + t2.transpose(1, 0, 3, 2)
    return dict(o=o, v=v, L=L, Loo=Loo, Lov=Lov, Lvv=Lvv, fock=fock, eo=eo, ev=ev, t1=t1, t2=t2)


def reference_eris(p):
    """What the reference's update_amps reads from ``eris`` (rmpccsd_slow.py:78-158, rintermediates.py)."""
    Loo, Lov, Lvv = p["Loo"], p["Lov"], p["Lvv"]
    Lvo = Lov.transpose(0, 2, 1)
    e = types.SimpleNamespace()
    e.fock = p["fock"]
    e.oooo = np.einsum("Lij,Lkl->ijkl", Loo, Loo, optimize=True)
    e.ovoo = np.einsum("Lia,Ljk->iajk", Lov, Loo, optimize=True)
    e.ooov = np.einsum("Lij,Lka->ijka", Loo, Lov, optimize=True)
    e.oovv = np.einsum("Lij,Lab->ijab", Loo, Lvv, optimize=True)
    e.ovov = np.einsum("Lia,Ljb->iajb", Lov, Lov, optimize=True)
    e.ovvo = np.einsum("Lia,Lbj->iabj", Lov, Lvo, optimize=True)
    e.ovvv = np.einsum("Lia,Lbc->iabc", Lov, Lvv, optimize=True)
    e.vvvv = np.einsum("Lab,Lcd->abcd", Lvv, Lvv, optimize=True)
    e.get_ovvv = lambda *a: e.ovvv
    return e
