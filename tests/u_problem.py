"""Seeded unrestricted DF problem for row f4 and the namespace the reference's updated_amp reads (umpcc_lowlevel.py:179-196)."""
import types

import numpy as np


def make_u_problem(oa, ob, n, L, seed=0, scale=0.2):
    rng = np.random.default_rng(seed)
    out = {}
    for tag, o in (("a", oa), ("b", ob)):
        v = n - o
        Lpq = rng.standard_normal((L, n, n)) * scale / np.sqrt(L)
        Lpq = Lpq + Lpq.transpose(0, 2, 1)
        eo = np.sort(rng.uniform(-2.0, -0.5, o))
        ev = np.sort(rng.uniform(0.3, 3.0, v))
        off = rng.standard_normal((n, n)) * 0.02
        f = off + off.T
        foo = f[:o, :o].copy(); np.fill_diagonal(foo, eo)
        fvv = f[o:, o:].copy(); np.fill_diagonal(fvv, ev)
        out[tag] = dict(Loo=np.ascontiguousarray(Lpq[:, :o, :o]), Lov=np.ascontiguousarray(Lpq[:, :o, o:]),
                        Lvv=np.ascontiguousarray(Lpq[:, o:, o:]), foo=foo, fvv=fvv, eo=eo, ev=ev,
                        t1=0.05 * rng.standard_normal((o, v)))
    return out


def reference_state(p):
    a, b = p["a"], p["b"]
    m = types.SimpleNamespace()
    m.t1 = [a["t1"].copy(), b["t1"].copy()]
    m.Loo, m.Lov, m.Lvv = a["Loo"], a["Lov"], a["Lvv"]
    m.LOO, m.LOV, m.LVV = b["Loo"], b["Lov"], b["Lvv"]
    m.foo, m.fvv, m.fOO, m.fVV = a["foo"], a["fvv"], b["foo"], b["fvv"]
    m.fock = [a["ev"][None, :] - a["eo"][:, None], b["ev"][None, :] - b["eo"][:, None]]
    eia_a = a["eo"][:, None] - a["ev"][None, :]
    eia_b = b["eo"][:, None] - b["ev"][None, :]
    ds = lambda x, y: (x[:, :, None, None] + y[None, None, :, :]).transpose(1, 0, 3, 2)     # 'ia+jb->aibj'
    m.D = [ds(eia_a, eia_a), ds(eia_a, eia_b), ds(eia_b, eia_b)]
    return m
