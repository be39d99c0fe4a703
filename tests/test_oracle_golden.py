"""The oracle (oracle/mpcc_oracle.py) against the fixtures the REFERENCE produced
(tests/golden/generate_golden.py ran pyscf/mpcc/dfrmpcc_lowlevel.py + pyscf/lib/diis.py unmodified)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN, golden_cases
from oracle.mpcc_oracle import (LowLevelOracle, PulayDIIS, amps_to_vector, tril_pack, tril_unpack_symmetric,
                                vector_to_amps, flops_per_updated_amps)
from pyscf_mpcc_b200.synthetic import SyntheticMF

STRIDE = 97


def load(path):
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    return g, o, v, L, float(g["scale"])


@pytest.mark.parametrize("path", golden_cases(), ids=os.path.basename)
@pytest.mark.parametrize("mode", ["blas", "einsum"])
def test_kernel_trace_matches_reference(path, mode):
    g, o, v, L, scale = load(path)
    if mode == "einsum" and o * v > 400:
        pytest.skip("un-optimised einsum only on the small cases (minutes otherwise)")
    orc = LowLevelOracle(SyntheticMF(o, v, L, seed=0, scale=scale), mode=mode)
    e = orc.kernel()
    assert orc.iterations == int(g["iterations"])
    tr = np.array(orc.history)
    np.testing.assert_allclose(tr[:, 1], g["trace_dE"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(tr[:, 2], g["trace_res"], rtol=1e-9, atol=1e-13)
    assert abs(e - g["trace_dE"][-1]) < 1e-12
    np.testing.assert_allclose(orc.t1, g["final_t1"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(orc.t2.reshape(-1)[::STRIDE], g["final_t2_samples"], rtol=0, atol=1e-12)
    assert abs(np.linalg.norm(orc.t2) - g["final_t2_norm"]) < 1e-11
    if "final_t2" in g:
        np.testing.assert_allclose(orc.t2, g["final_t2"], rtol=0, atol=1e-12)


@pytest.mark.parametrize("path", golden_cases(), ids=os.path.basename)
@pytest.mark.parametrize("tag", ["zero", "rand"])
def test_updated_amps_known_answers(path, tag):
    g, o, v, L, scale = load(path)
    orc = LowLevelOracle(SyntheticMF(o, v, L, seed=0, scale=scale), mode="blas")
    res, dE, t1n, t2 = orc.updated_amps(g[f"kat_{tag}_t1_in"])
    assert abs(res - g[f"kat_{tag}_res"]) < 1e-12 * max(1.0, abs(float(g[f"kat_{tag}_res"])))
    assert abs(dE - g[f"kat_{tag}_dE"]) < 1e-12
    np.testing.assert_allclose(t1n, g[f"kat_{tag}_t1_new"], rtol=0, atol=1e-12)
    assert t2.shape == (v, o, v, o)
    np.testing.assert_allclose(np.ascontiguousarray(t2).reshape(-1)[::STRIDE], g[f"kat_{tag}_t2_samples"], atol=1e-12)
    vec = amps_to_vector(t1n, t2, o, v)
    np.testing.assert_allclose(vec[::STRIDE], g[f"kat_{tag}_vec_samples"], atol=1e-12)
    assert abs(np.linalg.norm(vec) - g[f"kat_{tag}_vec_norm"]) < 1e-11


def test_diis_known_answer_sequence():
    g = np.load(os.path.join(GOLDEN, "diis_kat.npz"))
    d = PulayDIIS()
    for x, want in zip(g["inputs"], g["outputs"]):
        got = d.update(x.copy())
        np.testing.assert_allclose(got, want, rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(d.B, g["H"], rtol=1e-10, atol=1e-14)


def test_diis_jacobi_fixture_like_reference_test():
    """pyscf/lib/test/test_diis.py:21-60: make_ab(16), seed 1, Jacobi + DIIS reaches |ax-b| < 1e-6 in 20 steps."""
    np.random.seed(1)
    n = 16
    a = np.random.random((n, n)) + np.eye(n) * 2
    b = np.random.random(n)
    ad = np.diag(a).copy()
    x = np.zeros(n)
    d = PulayDIIS()
    for _ in range(20):
        x = (b - (a - np.diag(ad)) @ x) / ad
        x = d.update(x)
    assert np.linalg.norm(a @ x - b) < 1e-6


def test_pack_unpack_roundtrip_and_order():
    rng = np.random.default_rng(3)
    for n in (1, 2, 5, 12):
        m = rng.standard_normal((n, n))
        p = tril_pack(m)
        r, c = np.tril_indices(n)
        np.testing.assert_array_equal(p, m[r, c])            # NPdpack_tril order
        full = tril_unpack_symmetric(p)
        np.testing.assert_array_equal(full, full.T)
        np.testing.assert_array_equal(np.tril(full), np.tril(m))
    assert list(tril_pack(np.arange(9.0).reshape(3, 3))) == [0, 3, 4, 6, 7, 8]   # numpy_helper.py:262-263


def test_vector_roundtrip_keeps_lower_triangle_only():
    o, v = 3, 4
    rng = np.random.default_rng(0)
    t1 = rng.standard_normal((o, v))
    t2 = rng.standard_normal((v, o, v, o))
    vec = amps_to_vector(t1, t2, o, v)
    t1b, t2b = vector_to_amps(vec, o, v)
    np.testing.assert_array_equal(t1, t1b)
    assert t2b.shape == (o, o, v, v)
    m = np.ascontiguousarray(t2.transpose(0, 2, 1, 3)).reshape(o * v, o * v)
    s = np.tril(m) + np.tril(m, -1).T
    np.testing.assert_array_equal(t2b, s.reshape(o, v, o, v).transpose(0, 2, 1, 3))


def test_flop_formula_target_shape():
    assert abs(flops_per_updated_amps(64, 512, 1536) - 6.799e12) < 2e9
    from pyscf_mpcc_b200.flops import updated_amps_flops
    assert updated_amps_flops(21, 243, 654) == flops_per_updated_amps(21, 243, 654)


def test_config2_fixture_from_the_reference_itself():
    """BASELINE.json configs[1]'s shape (21, 243, 654): the fixture was produced by the reference's own files
    (tests/golden/generate_golden_config2.py, about 40 s of np.einsum per call); the GEMM-shaped oracle reproduces
    its iteration count, trace and amplitudes."""
    from pyscf_mpcc_b200.synthetic import default_scale
    g = np.load(os.path.join(GOLDEN, "config2_mpcc_21_243_654.npz"))
    o, v, L = (int(x) for x in g["shape"])
    assert (o, v, L) == (21, 243, 654) and float(g["scale"]) == default_scale(o, v)
    orc = LowLevelOracle(SyntheticMF(o, v, L, seed=0, scale=float(g["scale"])), mode="blas")
    e = orc.kernel()
    assert orc.iterations == int(g["iterations"]) == 5
    tr = np.array(orc.history)
    np.testing.assert_allclose(tr[:, 1], g["trace_dE"], rtol=0, atol=1e-11)
    np.testing.assert_allclose(tr[:, 2], g["trace_res"], rtol=1e-8, atol=1e-13)
    np.testing.assert_allclose(orc.t1, g["final_t1"], rtol=0, atol=1e-12)
    s = int(g["sample_stride"])
    np.testing.assert_allclose(orc.t2.reshape(-1)[::s], g["final_t2_samples"], rtol=0, atol=1e-12)
    assert abs(np.linalg.norm(orc.t2) - g["final_t2_norm"]) < 1e-10
    assert abs(e - g["trace_dE"][-1]) < 1e-11


@pytest.mark.parametrize("shape", [(5, 19, 116, 0.1, 4), (12, 64, 192, 0.1, 7), (21, 30, 70, 0.1, 32)])
def test_packed_form_equals_plain(shape):
    """``updated_amps_packed`` (slab by slab, DIIS vector written directly — the form that produced the full-size
    trace fixture) against ``updated_amps`` + ``amps_to_vector``, and a whole solve through it against the fixture
    the reference produced."""
    o, v, L, scale, slab = shape
    mf = SyntheticMF(o, v, L, seed=0, scale=scale)
    orc = LowLevelOracle(mf, mode="blas")
    t1 = 0.05 * np.random.default_rng(2).standard_normal((o, v))
    res, dE, t1n, t2 = orc.updated_amps(t1)
    res2, dE2, t1n2, vec2, ss = orc.updated_amps_packed(t1, slab=slab)
    assert abs(res - res2) < 1e-14 and abs(dE - dE2) < 1e-13
    np.testing.assert_allclose(t1n2, t1n, rtol=0, atol=1e-14)
    np.testing.assert_allclose(vec2, amps_to_vector(t1n, t2, o, v), rtol=0, atol=1e-14)
    assert abs(ss - float(np.vdot(t2, t2))) < 1e-12 * max(1.0, ss)
    g = np.load(os.path.join(GOLDEN, f"mpcc_{o}_{v}_{L}.npz"))
    t1, ad, err, count, tr = np.zeros((o, v)), PulayDIIS(), np.inf, 0, []
    while err > 1e-6 and count < 50:
        res, dE, t1n, vec, _ = orc.updated_amps_packed(t1, slab=slab)
        t1 = np.array(ad.update(vec)[:o * v]).reshape(o, v)
        count += 1
        err = res
        tr.append((dE, res))
    assert count == int(g["iterations"])
    np.testing.assert_allclose(np.array(tr)[:, 0], g["trace_dE"], rtol=0, atol=1e-12)
    np.testing.assert_allclose(t1, g["final_t1"], rtol=0, atol=1e-12)
