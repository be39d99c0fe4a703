"""Regenerate tests/golden/umpcc_*.npz by running the REFERENCE's own unrestricted low-level function.

Run in the build container only (needs /root/reference):
    python tests/golden/generate_golden_u.py
Loads ``pyscf/cc/umpcc_lowlevel.py`` unmodified through ``oracle/ref_shim.py`` and records ``updated_amp`` applied
``steps`` times to the seeded problems of ``tests/u_problem.py`` (res, dE, both t1 blocks, the three t2 blocks as
norm + strided samples).
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim                               # noqa: E402
from u_problem import make_u_problem, reference_state      # noqa: E402

CASES = [(3, 2, 8, 13, 1, 2), (6, 5, 20, 31, 2, 3), (12, 10, 52, 90, 3, 2), (17, 16, 49, 64, 4, 1)]   # oa, ob, nmo, naux, seed, steps
STRIDE = 89


def main():
    ref = ref_shim.load_reference_unrestricted()
    for (oa, ob, n, L, seed, steps) in CASES:
        p = make_u_problem(oa, ob, n, L, seed=seed)
        m = reference_state(p)
        with contextlib.redirect_stdout(io.StringIO()):
            for _ in range(steps):
                res, dE, t1, t2 = ref.updated_amp(m)
        out = dict(shape=np.array([oa, ob, n, L]), seed=seed, steps=steps, stride=STRIDE, res=res, dE=dE,
                   t1a=t1[0], t1b=t1[1])
        for k, name in enumerate(("t2aa", "t2ab", "t2bb")):
            out[name + "_samples"] = t2[k].reshape(-1)[::STRIDE].copy()
            out[name + "_norm"] = np.linalg.norm(t2[k])
        path = os.path.join(HERE, f"umpcc_{oa}_{ob}_{n}_{L}.npz")
        np.savez_compressed(path, **out)
        print(path, "res", res, "dE", dE)


if __name__ == "__main__":
    main()
