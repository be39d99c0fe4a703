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


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "ukernel_*.npz"))), ids=os.path.basename)
def test_kernel_loop_and_energy_on_device_against_fixture(path):
    """kernel / run_diis / the s4 vector layout / compute_CC2_energy of the unrestricted prototype on the device
    (mpccb_take_2d, mpccb_takebak_2d, mpccb_vec_diis_push) against the loop run around the reference's own functions
    (tests/golden/generate_golden_u2.py)."""
    from pyscf_mpcc_b200 import umpcc_lowlevel as U
    from pyscf_mpcc_b200.tensor_engine import make_cuda_engine
    g = np.load(path)
    oa, ob, n, L = (int(x) for x in g["shape"])
    p = make_u_problem(oa, ob, n, L, seed=int(g["seed"]))
    eng, keep = make_cuda_engine(max(oa, ob), n - min(oa, ob), L)
    st = U.UMPCCState(eng, p["a"], p["b"])
    first = {}
    orig = U.amplitudes_to_vector

    def spy(t1, t2, out=None, *, eng):
        v = orig(t1, t2, out, eng=eng)
        first.setdefault("vec", eng.be.download(v).copy())
        return v

    U.amplitudes_to_vector = spy
    try:
        DE = U.kernel(st, verbose=False)
    finally:
        U.amplitudes_to_vector = orig
    stride = int(g["stride"])
    assert st.iterations == int(g["iterations"]) and abs(DE - float(g["DE"])) <= 1e-8
    hist = np.array(st.history)
    assert np.abs(hist[:, 1] - g["history"][:, 1]).max() <= 1e-8
    np.testing.assert_allclose(hist[:, 2], g["history"][:, 2], rtol=1e-6, atol=1e-12)
    assert first["vec"].size == int(g["first_vector_size"])
    assert np.abs(first["vec"][::stride] - g["first_vector_samples"]).max() <= 1e-10
    assert abs(np.linalg.norm(first["vec"]) - float(g["first_vector_norm"])) <= 1e-10 * max(1.0, float(g["first_vector_norm"]))
    assert np.abs(eng.numpy(st.t1[0]) - g["t1a"]).max() <= 1e-10 and np.abs(eng.numpy(st.t1[1]) - g["t1b"]).max() <= 1e-10
    for k, name in enumerate(("t2aa", "t2ab", "t2bb")):
        t = eng.numpy(st.t2[k])
        assert t.shape == tuple(g[name + "_shape"])
        assert np.abs(t.reshape(-1)[::stride] - g[name + "_samples"]).max() <= 1e-10
    rng = np.random.default_rng(int(g["cc2_seed"]))
    va = n - oa
    t1e = 0.05 * rng.standard_normal((oa, va))
    t2e = 0.1 * rng.standard_normal((va, oa, va, oa))
    assert abs(U.compute_CC2_energy(st, eng.tensor(t1e), eng.tensor(t2e)) - float(g["cc2_energy"])) <= 1e-10
