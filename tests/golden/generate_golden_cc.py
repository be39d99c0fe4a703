"""Regenerate tests/golden/rmpccsd_*.npz by running the REFERENCE's own CCSD-residual code.

Run in the build container only (needs /root/reference):
    python tests/golden/generate_golden_cc.py
Loads ``pyscf/cc/rmpccsd_slow.py`` and ``pyscf/cc/rintermediates.py`` unmodified through ``oracle/ref_shim.py`` and
records ``update_amps`` (t1new whole, t2new as norm + strided samples) and ``energy`` for seeded problems of
``tests/cc_problem.py`` (DF-built integrals, non-diagonal Fock matrix, random amplitudes, active-space classes).
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim                      # noqa: E402
from cc_problem import make_problem, reference_eris   # noqa: E402

CASES = [  # o, v, L, seed, act_hole, act_particle, idx_singles, idx_doubles
    (3, 5, 11, 1, [0, 2], [1, 3], [0, 2], [1, 5, 10, 12]),
    (6, 14, 40, 2, [1, 4, 5], [0, 3, 7, 8, 13], [1], [0, 3, 6, 9, 11, 14]),
    (10, 40, 96, 3, [6, 7, 8, 9], list(range(12)), [0, 1, 2], list(range(15))),
    (16, 33, 70, 4, [], [], [], []),
]
STRIDE = 101


def main():
    slow, _ = ref_shim.load_reference_cc()
    cc = types.SimpleNamespace(cc2=False)
    for (o, v, L, seed, ah, ap, ids, idd) in CASES:
        p = make_problem(o, v, L, seed=seed)
        er = reference_eris(p)
        t1n, t2n = slow.update_amps(cc, p["t1"], p["t2"], er, ah, ap, ids, idd)
        e = slow.energy(cc, p["t1"], p["t2"], er)
        path = os.path.join(HERE, f"rmpccsd_{o}_{v}_{L}.npz")
        np.savez_compressed(path, shape=np.array([o, v, L]), seed=seed, act_hole=np.array(ah, dtype=int),
                            act_particle=np.array(ap, dtype=int), idx_singles=np.array(ids, dtype=int),
                            idx_doubles=np.array(idd, dtype=int), t1new=t1n, stride=STRIDE,
                            t2new_samples=t2n.reshape(-1)[::STRIDE].copy(), t2new_norm=np.linalg.norm(t2n), energy=e)
        print(path, "E =", e, "|t2new| =", np.linalg.norm(t2n))


if __name__ == "__main__":
    main()
