"""Full-size (BASELINE.json configs[2]: nocc=64, nvir=512, naux=1536) checks: the iteration trace against the
fixture the oracle produced on the same inputs (tests/golden/generate_trace_config3.py), size-independent
properties, and an oracle comparison on a slab of rows (the oracle finishes those in seconds)."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def big():
    import torch
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import host_blocks, default_scale
    o, v, L = 64, 512, 1536
    # the inputs of the trace fixture and of bench.py: SyntheticMF's stream, drawn in chunks on the host
    Loo, Lov, Lvv, eo, ev = host_blocks(o, v, L, seed=0, scale=default_scale(o, v))
    ll = MPCC_LL(blocks=(Loo, Lov, Lvv, np.diag(eo), np.diag(ev), np.concatenate([eo, ev])))
    ll.verbose_print = False
    return ll


def test_fullsize_trace_against_oracle_fixture(big):
    """kernel()'s loop at the shape the metric is quoted on, iteration by iteration against the oracle's trace:
    identical iteration count, |dE| <= 1e-8 Eh, res to 1e-6 relative, amplitude and DIIS-vector norms, final t1 to
    1e-10 (north_star's tolerances)."""
    import torch
    from pyscf_mpcc_b200.diis import DIIS
    g = np.load(os.path.join(GOLDEN, "trace_64_512_1536.npz"))
    ll = big
    o, v = ll.nocc, ll.nvir
    assert tuple(int(x) for x in g["shape"]) == (o, v, ll.naux)
    ll.t1 = np.zeros((o, v))
    ad = DIIS()
    err, count = np.inf, 0
    while err > 1e-6 and count < 50:
        res, dE, t1n, t2 = ll.updated_amps()
        assert count < len(g["trace_dE"]), "more iterations than the oracle needed"
        assert abs(dE - g["trace_dE"][count]) <= 1e-8
        assert abs(res - g["trace_res"][count]) <= 1e-6 * g["trace_res"][count] + 1e-12
        assert abs(np.linalg.norm(t1n) - g["trace_t1_new_norm"][count]) <= 1e-10
        t2n = float(torch.linalg.vector_norm(ll._T).item())
        assert abs(t2n - g["trace_t2_norm"][count]) <= 1e-9 * g["trace_t2_norm"][count]
        vec = ll.amplitudes_to_vector(t1n, t2)
        vn = float(torch.linalg.vector_norm(vec.packed_device()).item())
        assert abs(vn - g["trace_vec_norm"][count]) <= 1e-9 * g["trace_vec_norm"][count]
        if count == 0:
            assert np.abs(t1n - g["it1_t1_new"]).max() <= 1e-10
        ll.t1, ll.t2 = ll.vector_to_amplitudes(ad.update(vec))
        assert abs(np.linalg.norm(ll.t1) - g["trace_t1_next_norm"][count]) <= 1e-10
        if ad.last_coeff is not None:
            c = g["diis_coeff"][count][:len(ad.last_coeff)]
            np.testing.assert_allclose(ad.last_coeff[1:], c[1:], rtol=0, atol=1e-6)
        count += 1
        err = res
    assert count == int(g["iterations"]) and bool(g["converged"])
    assert np.abs(ll.t1 - g["final_t1"]).max() <= 1e-10
    del ad
    torch.cuda.empty_cache()


def test_fullsize_slab_against_oracle_and_symmetry(big):
    from oracle.mpcc_oracle import LowLevelOracle
    ll = big
    o, v, L = ll.nocc, ll.nvir, ll.naux
    t1 = 0.01 * np.random.default_rng(0).standard_normal((o, v))
    ll.t1 = t1
    res, dE, t1n, t2 = ll.updated_amps()
    assert np.isfinite(res) and np.isfinite(dE) and dE < 0
    orc = LowLevelOracle(blocks=(ll.Loo, ll.Lov, ll.Lvv, ll.foo, ll.fvv, ll.mo_energy), mode="blas")
    a0, a1 = 200, 203
    _, _, _, t2_slab = orc.updated_amps(t1, a_slab=(a0, a1))          # [a,i,b,j] rows a0..a1
    import torch
    T = ll._T.view(v, v, o, o)
    got = T[a0:a1].permute(0, 2, 1, 3).cpu().numpy()                  # -> [a,i,b,j]
    assert np.abs(got - t2_slab).max() <= 1e-10
    # t2[a,b,i,j] == t2[b,a,j,i] (eris is a Gram matrix): sampled blocks
    for (a, b) in [(0, 511), (17, 300), (255, 256), (400, 3)]:
        assert (T[a, b] - T[b, a].T).abs().max().item() <= 1e-12
    # Yvo rows of the slab
    Y = ll.debug_buffer("Y").reshape(v, o, L)
    Lov = orc.Lov
    y_ref = np.einsum("aibj,Ljb->aiL", t2_slab, Lov, optimize=True)
    assert np.abs(Y[a0:a1] - y_ref).max() <= 1e-10


def test_fullsize_iteration_converges_like_small_cases(big):
    ll = big
    ll.con_tol, ll.max_its = 1e-6, 50
    e = ll.kernel()
    assert 3 <= ll.iterations <= 12 and ll.history[-1][2] < 1e-6
    # energy lags amplitudes by one update: re-evaluating at the converged t1 changes dE below 1e-6
    assert abs(ll.energy() - e) < 1e-5
    # DIIS vector of the converged amplitudes reproduces itself: unpack(pack) keeps t1 exactly
    t1 = ll.t1.copy()
    vec = ll._adiis._xprev
    assert np.array_equal(vec[:ll.nov].cpu().numpy().reshape(ll.nocc, ll.nvir), t1)
