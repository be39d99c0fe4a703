"""GPU parity of the unrestricted low-level residual (row f4) against fixtures produced by the reference's own
updated_amp (tests/golden/umpcc_*.npz, generate_golden_u.py)."""
import glob
import os

import numpy as np
import pytest

from conftest import GOLDEN
from u_problem import make_u_problem

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "umpcc_*.npz"))), ids=os.path.basename)
def test_updated_amp_against_reference_fixture(path):
    from pyscf_mpcc_b200 import umpcc_lowlevel as U
    from pyscf_mpcc_b200.tensor_engine import make_cuda_engine
    g = np.load(path)
    oa, ob, n, L = (int(x) for x in g["shape"])
    p = make_u_problem(oa, ob, n, L, seed=int(g["seed"]))
    eng, keep = make_cuda_engine(max(oa, ob), n - min(oa, ob), L)
    st = U.UMPCCState(eng, p["a"], p["b"])
    st.set_t1(p["a"]["t1"], p["b"]["t1"])
    for _ in range(int(g["steps"])):
        res, dE, t1, t2 = U.updated_amp(st)
    assert abs(dE - float(g["dE"])) <= 1e-8 and abs(res - float(g["res"])) <= 1e-10
    assert np.abs(eng.numpy(t1[0]) - g["t1a"]).max() <= 1e-10 and np.abs(eng.numpy(t1[1]) - g["t1b"]).max() <= 1e-10
    for k, name in enumerate(("t2aa", "t2ab", "t2bb")):
        x = eng.numpy(t2[k])
        assert np.abs(x.reshape(-1)[::int(g["stride"])] - g[name + "_samples"]).max() <= 1e-10
        assert abs(np.linalg.norm(x) - float(g[name + "_norm"])) <= 1e-10 * max(1.0, float(g[name + "_norm"]))
    assert eng.gemms >= 40
