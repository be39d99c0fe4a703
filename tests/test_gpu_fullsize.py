"""Full-size (BASELINE.json configs[2]: nocc=64, nvir=512, naux=1536) checks through size-independent
properties plus an oracle comparison on a slab of rows (the oracle finishes those in seconds)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def big():
    import torch
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import device_blocks, default_scale
    o, v, L = 64, 512, 1536
    Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
    ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])))
    ll.verbose_print = False
    return ll


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
