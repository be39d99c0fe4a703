"""Fixtures for the CC2 branch and the driver loop of the high-level stage, from the REFERENCE's own code.

Run in the build container only (needs /root/reference):
    python tests/golden/generate_golden_cc2.py            (--asym: only the non-symmetric t2 case at the end of this file)
* ``rmpccsd_cc2_*.npz``    — ``rmpccsd_slow.update_amps`` with ``cc.cc2 = True`` (:115-133).
* ``rmpccsd_kernel_*.npz`` — the iteration ``RMPCCSD.ccsd`` runs (:223-258 -> ``ccsd.kernel``, ccsd.py:42-150): the loop
  is ``oracle/cc_oracle.ccsd_kernel`` (restated; ccsd.py itself is not importable here) around the reference's
  unmodified ``update_amps`` / ``energy`` (rmpccsd_slow.py) and ``lib.diis.DIIS`` (pyscf/lib/diis.py); start = the
  MP2 amplitudes of ``init_amps`` (:206-219). Recorded: cycles, converged, per-cycle E_corr / dE / norm, final t1,
  samples and norm of the final t2.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import cc_oracle, ref_shim                  # noqa: E402
from cc_problem import make_problem, reference_eris     # noqa: E402

CC2_CASES = [  # o, v, L, seed, act_hole, act_particle, idx_singles, idx_doubles
    (3, 5, 11, 1, [0, 2], [1, 3], [0, 2], [1, 5, 10, 12]),
    (10, 40, 96, 3, [6, 7, 8, 9], list(range(12)), [0, 1, 2], list(range(15))),
    (16, 33, 70, 4, [], [], [], []),
]
KERNEL_CASES = [  # o, v, L, seed, cc2, act_hole, act_particle, idx_s, idx_d
    (4, 8, 20, 5, False, [0, 2], [1, 3], [0], [1, 5]),
    (4, 8, 20, 5, True, [0, 2], [1, 3], [0], [1, 5]),
    (6, 14, 40, 2, False, [1, 4, 5], [0, 3, 7, 8, 13], [1], [0, 3, 6, 9, 11, 14]),
    (8, 24, 60, 7, False, [], [], [], []),
]
STRIDE = 101


def main():
    slow, _ = ref_shim.load_reference_cc()
    _, diis = ref_shim.load_reference()
    for (o, v, L, seed, ah, ap, ids, idd) in CC2_CASES:
        p = make_problem(o, v, L, seed=seed)
        er = reference_eris(p)
        cc = types.SimpleNamespace(cc2=True)
        t1n, t2n = slow.update_amps(cc, p["t1"], p["t2"], er, ah, ap, ids, idd)
        path = os.path.join(HERE, f"rmpccsd_cc2_{o}_{v}_{L}.npz")
        np.savez_compressed(path, shape=np.array([o, v, L]), seed=seed, act_hole=np.array(ah, dtype=int),
                            act_particle=np.array(ap, dtype=int), idx_singles=np.array(ids, dtype=int),
                            idx_doubles=np.array(idd, dtype=int), t1new=t1n, stride=STRIDE,
                            t2new_samples=t2n.reshape(-1)[::STRIDE].copy(), t2new_norm=np.linalg.norm(t2n))
        print(path, "|t2new| =", np.linalg.norm(t2n))
    for (o, v, L, seed, cc2, ah, ap, ids, idd) in KERNEL_CASES:
        p = make_problem(o, v, L, seed=seed)
        er = reference_eris(p)
        cc = types.SimpleNamespace(cc2=cc2)
        emp2, t1, t2 = cc_oracle.init_amps(p["fock"], er.ovov, o)
        conv, e, t1, t2, hist = cc_oracle.ccsd_kernel(
            lambda a, b: slow.update_amps(cc, a, b, er, ah, ap, ids, idd), lambda a, b: slow.energy(cc, a, b, er),
            lambda: diis.DIIS(), t1, t2, o + v, o)
        path = os.path.join(HERE, f"rmpccsd_kernel_{'cc2' if cc2 else 'ccsd'}_{o}_{v}_{L}.npz")
        np.savez_compressed(path, shape=np.array([o, v, L]), seed=seed, cc2=cc2, act_hole=np.array(ah, dtype=int),
                            act_particle=np.array(ap, dtype=int), idx_singles=np.array(ids, dtype=int),
                            idx_doubles=np.array(idd, dtype=int), converged=conv, e_corr=e, emp2=emp2,
                            history=np.array(hist), t1=t1, stride=STRIDE, t2_samples=t2.reshape(-1)[::STRIDE].copy(),
                            t2_norm=np.linalg.norm(t2))
        print(path, "cycles =", len(hist), "converged =", conv, "E_corr =", e)


if __name__ == "__main__" and "--asym" not in sys.argv:
    main()


def asym_case():
    """``rmpccsd_asym_4_9_21.npz`` — update_amps of the reference on a t2 that is NOT pair-symmetric
    (t2[ijab] != t2[jiba]): make_problem's t2 plus seeded noise, both branches."""
    slow, _ = ref_shim.load_reference_cc()
    o, v, L, seed, noise_seed, eps = 4, 9, 21, 11, 5, 0.05
    p = make_problem(o, v, L, seed=seed)
    t2 = p["t2"] + eps * np.random.default_rng(noise_seed).standard_normal(p["t2"].shape)
    er = reference_eris(p)
    out = dict(shape=np.array([o, v, L]), seed=seed, noise_seed=noise_seed, eps=eps, act_hole=np.array([0, 2]),
               act_particle=np.array([1, 3]), idx_singles=np.array([0]), idx_doubles=np.array([1, 5]))
    for cc2 in (False, True):
        t1n, t2n = slow.update_amps(types.SimpleNamespace(cc2=cc2), p["t1"], t2, er, [0, 2], [1, 3], [0], [1, 5])
        tag = "cc2" if cc2 else "ccsd"
        out[f"t1new_{tag}"], out[f"t2new_{tag}"] = t1n, t2n
    path = os.path.join(HERE, f"rmpccsd_asym_{o}_{v}_{L}.npz")
    np.savez_compressed(path, **out)
    print(path)


if __name__ == "__main__" and "--asym" in sys.argv:
    asym_case()
