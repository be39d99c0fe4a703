"""CPU check of the unrestricted low-level transcription (row f4): pyscf_mpcc_b200/umpcc_lowlevel.updated_amp
through the NumPy stand-in backend against the REFERENCE's own updated_amp (pyscf/cc/umpcc_lowlevel.py:174-357,
executed unmodified through oracle/ref_shim.py) and against the committed fixtures."""
import contextlib
import glob
import io
import os

import numpy as np
import pytest

from conftest import GOLDEN
from np_backend import NumpyBackend
from oracle import ref_shim
from pyscf_mpcc_b200 import umpcc_lowlevel as U
from pyscf_mpcc_b200.tensor_engine import Engine
from u_problem import make_u_problem, reference_state

needs_ref = pytest.mark.skipif(not ref_shim.full_tree_available(), reason="full reference tree not present")


def ours(p, steps=1, eng=None):
    eng = eng or Engine(NumpyBackend([0.0], [1.0]))
    st = U.UMPCCState(eng, p["a"], p["b"])
    st.set_t1(p["a"]["t1"], p["b"]["t1"])
    out = None
    for _ in range(steps):
        out = U.updated_amp(st)
    res, dE, t1, t2 = out
    return res, dE, [eng.numpy(x) for x in t1], [eng.numpy(x) for x in t2]


@needs_ref
@pytest.mark.parametrize("case", [(3, 2, 8, 13), (4, 4, 9, 10), (1, 3, 7, 5)])
def test_updated_amp_against_reference_code(case):
    oa, ob, n, L = case
    p = make_u_problem(oa, ob, n, L, seed=sum(case))
    ref = ref_shim.load_reference_unrestricted()
    m = reference_state(p)
    with contextlib.redirect_stdout(io.StringIO()):
        for _ in range(2):
            res_r, dE_r, t1_r, t2_r = ref.updated_amp(m)
    res, dE, t1, t2 = ours(p, steps=2)
    assert abs(res - res_r) <= 1e-12 and abs(dE - dE_r) <= 1e-12
    for x, y in zip(t1 + t2, list(t1_r) + list(t2_r)):
        assert np.abs(x - y).max() <= 1e-12


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "umpcc_*.npz"))), ids=os.path.basename)
def test_updated_amp_against_fixture(path):
    g = np.load(path)
    oa, ob, n, L = (int(x) for x in g["shape"])
    p = make_u_problem(oa, ob, n, L, seed=int(g["seed"]))
    res, dE, t1, t2 = ours(p, steps=int(g["steps"]))
    assert abs(res - float(g["res"])) <= 1e-12 and abs(dE - float(g["dE"])) <= 1e-12
    assert np.abs(t1[0] - g["t1a"]).max() <= 1e-12 and np.abs(t1[1] - g["t1b"]).max() <= 1e-12
    for k, name in enumerate(("t2aa", "t2ab", "t2bb")):
        assert np.abs(t2[k].reshape(-1)[::int(g["stride"])] - g[name + "_samples"]).max() <= 1e-12
        assert abs(np.linalg.norm(t2[k]) - float(g[name + "_norm"])) <= 1e-11
