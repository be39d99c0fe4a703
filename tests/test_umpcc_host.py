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


# ---- the rest of the prototype: vector layout, run_diis, kernel loop, compute_CC2_energy ----------------------------
def _kernel_state(p, eng=None):
    eng = eng or Engine(NumpyBackend([0.0], [1.0]))
    return eng, U.UMPCCState(eng, p["a"], p["b"])


@needs_ref
@pytest.mark.parametrize("case", [(3, 2, 8, 13), (4, 4, 9, 10), (1, 3, 7, 5), (2, 1, 4, 6)])
def test_vector_layout_and_loop_against_reference_code(case):
    """amplitudes_to_vector / vector_to_amplitudes / run_diis / kernel (umpcc_lowlevel.py:15-98) against the reference's
    own functions (ccsd's s4 helpers restated in the oracle), iteration by iteration."""
    from oracle import cc_oracle
    oa, ob, n, L = case
    p = make_u_problem(oa, ob, n, L, seed=sum(case))
    ref = ref_shim.load_reference_unrestricted()
    _, diis = ref_shim.load_reference()
    # layout round trip on the blocks one residual evaluation produces
    m = reference_state(p)
    with contextlib.redirect_stdout(io.StringIO()):
        _, _, t1_r, t2_r = ref.updated_amp(m)
    vec_r = ref.amplitudes_to_vector(t1_r, t2_r)
    back_r = ref.vector_to_amplitudes(vec_r, n, (oa, ob))
    eng, st = _kernel_state(p)
    st.set_t1(p["a"]["t1"], p["b"]["t1"])
    _, _, t1, t2 = U.updated_amp(st)
    vec = U.amplitudes_to_vector(t1, t2, eng=eng)
    assert np.array_equal(np.asarray(vec), vec_r) or np.abs(np.asarray(vec) - vec_r).max() <= 1e-13
    back = U.vector_to_amplitudes(vec, n, (oa, ob), eng=eng)
    for x, y in zip(back[0] + back[1], list(back_r[0]) + list(back_r[1])):
        assert x.shape == y.shape and np.abs(eng.numpy(x) - y).max() <= 1e-13
    # the loop
    m = reference_state(p)
    m.t1 = [np.zeros_like(m.t1[0]), np.zeros_like(m.t1[1])]
    m.diis, m.nmo, m.nocc = True, n, (oa, ob)
    with contextlib.redirect_stdout(io.StringIO()):
        DE_r, count_r, hist_r = cc_oracle.umpcc_kernel(m, ref.updated_amp, ref.run_diis, lambda: diis.DIIS())
    eng, st = _kernel_state(p)
    DE = U.kernel(st, verbose=False)
    assert st.iterations == count_r and abs(DE - DE_r) <= 1e-12
    assert np.abs(np.array(st.history) - np.array(hist_r, dtype=float)).max() <= 1e-12
    for x, y in zip(st.t1 + st.t2, list(m.t1) + list(m.t2)):
        assert np.abs(eng.numpy(x) - y).max() <= 1e-12
    # a vector of the spin-orbital length is rejected like in the reference (:36-38)
    nso, vso = oa + ob, 2 * n - oa - ob
    with pytest.raises(RuntimeError, match="GCCSD"):
        U.vector_to_amplitudes(np.zeros(nso * vso + nso * (nso - 1) // 2 * vso * (vso - 1) // 2), n, (oa, ob), eng=eng)


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "ukernel_*.npz"))), ids=os.path.basename)
def test_kernel_loop_and_energy_against_fixture(path):
    g = np.load(path)
    oa, ob, n, L = (int(x) for x in g["shape"])
    if oa * (n - oa) > 60:
        pytest.skip("NumPy stand-in: small cases only")
    p = make_u_problem(oa, ob, n, L, seed=int(g["seed"]))
    eng, st = _kernel_state(p)
    first = {}
    orig = U.amplitudes_to_vector

    def spy(t1, t2, out=None, *, eng):
        v = orig(t1, t2, out, eng=eng)
        first.setdefault("vec", np.array(eng.be.download(v)))
        return v

    U.amplitudes_to_vector = spy
    try:
        DE = U.kernel(st, verbose=False)
    finally:
        U.amplitudes_to_vector = orig
    stride = int(g["stride"])
    assert st.iterations == int(g["iterations"]) and abs(DE - float(g["DE"])) <= 1e-10
    assert np.abs(np.array(st.history) - g["history"]).max() <= 1e-10
    assert first["vec"].size == int(g["first_vector_size"])
    assert np.abs(first["vec"][::stride] - g["first_vector_samples"]).max() <= 1e-12
    assert abs(np.linalg.norm(first["vec"]) - float(g["first_vector_norm"])) <= 1e-11
    assert np.abs(eng.numpy(st.t1[0]) - g["t1a"]).max() <= 1e-10 and np.abs(eng.numpy(st.t1[1]) - g["t1b"]).max() <= 1e-10
    for k, name in enumerate(("t2aa", "t2ab", "t2bb")):
        t = eng.numpy(st.t2[k])
        assert t.shape == tuple(g[name + "_shape"])
        assert np.abs(t.reshape(-1)[::stride] - g[name + "_samples"]).max() <= 1e-10
        assert abs(np.linalg.norm(t) - float(g[name + "_norm"])) <= 1e-10
    rng = np.random.default_rng(int(g["cc2_seed"]))
    va = n - oa
    t1e = 0.05 * rng.standard_normal((oa, va))
    t2e = 0.1 * rng.standard_normal((va, oa, va, oa))
    assert abs(U.compute_CC2_energy(st, eng.tensor(t1e), eng.tensor(t2e)) - float(g["cc2_energy"])) <= 1e-12
