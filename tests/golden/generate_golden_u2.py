"""Fixtures for the rest of the unrestricted prototype (row f4), from the REFERENCE's own functions.

Run in the build container only (needs /root/reference):
    python tests/golden/generate_golden_u2.py
``ukernel_*.npz``: the iteration of ``kernel`` (pyscf/cc/umpcc_lowlevel.py:48-89) — the loop is
``oracle/cc_oracle.umpcc_kernel`` (restated: what precedes the loop in the reference needs a PySCF mean field and an
undefined global) around the reference's unmodified ``updated_amp`` (:174-357), ``run_diis`` / ``amplitudes_to_vector`` /
``vector_to_amplitudes`` (:15-46, :92-98) and ``lib.diis.DIIS``; the two ``ccsd`` vector helpers they call are the
restatements of ``oracle/cc_oracle.py`` (ccsd.py itself cannot be imported). Start: t1 = 0 (:171).
Recorded: iteration count, (DE, res) per iteration, final t1a / t1b, strided samples and norm of the DIIS vector of the
first iteration and of the final t2 blocks; and ``compute_CC2_energy`` (:100-127) on t1 / t2 drawn from ``cc2_seed``
(t1 = 0.05 N(0,1) (oa,va), then t2 = 0.1 N(0,1) (va,oa,va,oa), numpy default_rng).
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import cc_oracle, ref_shim                  # noqa: E402
from u_problem import make_u_problem, reference_state   # noqa: E402

CASES = [(3, 2, 8, 13, 4), (1, 3, 7, 5, 16), (6, 5, 20, 31, 2), (12, 10, 40, 70, 3)]     # oa, ob, nmo, naux, seed
STRIDE = 37


def main():
    u = ref_shim.load_reference_unrestricted()
    _, diis = ref_shim.load_reference()
    for oa, ob, n, L, seed in CASES:
        p = make_u_problem(oa, ob, n, L, seed=seed)
        m = reference_state(p)
        m.t1 = [np.zeros_like(m.t1[0]), np.zeros_like(m.t1[1])]
        m.diis, m.nmo, m.nocc = True, n, (oa, ob)
        first = {}

        def run_diis(myll, t1, t2, adiis):
            if not first:
                first["vec"] = u.amplitudes_to_vector(t1, t2).copy()
            return u.run_diis(myll, t1, t2, adiis)

        with contextlib.redirect_stdout(io.StringIO()):
            DE, count, hist = cc_oracle.umpcc_kernel(m, u.updated_amp, run_diis, lambda: diis.DIIS())
        rng = np.random.default_rng(seed + 100)
        va = n - oa
        t1e = 0.05 * rng.standard_normal((oa, va))
        t2e = 0.1 * rng.standard_normal((va, oa, va, oa))
        out = dict(shape=np.array([oa, ob, n, L]), seed=seed, iterations=count, DE=DE, history=np.array(hist, dtype=float),
                   t1a=m.t1[0], t1b=m.t1[1], first_vector_samples=first["vec"][::STRIDE].copy(),
                   first_vector_norm=np.linalg.norm(first["vec"]), first_vector_size=first["vec"].size, stride=STRIDE,
                   cc2_seed=seed + 100, cc2_energy=u.compute_CC2_energy(reference_state(p), t1e, t2e))
        for name, t in zip(("t2aa", "t2ab", "t2bb"), m.t2):
            out[name + "_samples"] = np.ascontiguousarray(t).reshape(-1)[::STRIDE].copy()
            out[name + "_norm"] = np.linalg.norm(t)
            out[name + "_shape"] = np.array(t.shape)
        path = os.path.join(HERE, f"ukernel_{oa}_{ob}_{n}_{L}.npz")
        np.savez_compressed(path, **out)
        print(path, "iterations =", count, "DE =", DE)


if __name__ == "__main__":
    main()
