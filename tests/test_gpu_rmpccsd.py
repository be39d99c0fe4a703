"""GPU parity of the CCSD-type residual (rows f1/f2) against the fixtures the reference's own
rmpccsd_slow.update_amps / energy produced (tests/golden/rmpccsd_*.npz, generate_golden_cc.py)."""
import glob
import os
import types

import numpy as np
import pytest

from conftest import GOLDEN
from cc_problem import make_problem

pytestmark = pytest.mark.gpu


def _device_problem(p):
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    o = p["o"]
    ll = MPCC_LL(blocks=(p["Loo"], p["Lov"], p["Lvv"], p["fock"][:o, :o].copy(), p["fock"][o:, o:].copy(),
                         np.concatenate([p["eo"], p["ev"]])))
    return ll, ll.high_level_eris(fock=p["fock"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rmpccsd_[0-9]*.npz"))), ids=os.path.basename)
def test_update_amps_and_energy_against_reference_fixture(path):
    from pyscf_mpcc_b200 import rmpccsd
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    ll, eris = _device_problem(p)
    cc = types.SimpleNamespace(cc2=False)
    t1n, t2n = rmpccsd.update_amps(cc, p["t1"], p["t2"], eris, list(g["act_hole"]), list(g["act_particle"]),
                                   list(g["idx_singles"]), list(g["idx_doubles"]))
    e = rmpccsd.energy(cc, p["t1"], p["t2"], eris)
    scale = max(1.0, float(np.abs(g["t2new_samples"]).max()))
    assert np.abs(t1n - g["t1new"]).max() <= 1e-10 * max(1.0, float(np.abs(g["t1new"]).max()))
    assert np.abs(t2n.reshape(-1)[::int(g["stride"])] - g["t2new_samples"]).max() <= 1e-10 * scale
    assert abs(np.linalg.norm(t2n) - float(g["t2new_norm"])) <= 1e-10 * float(g["t2new_norm"])
    assert abs(e - float(g["energy"])) <= 1e-8


def test_pp_ladder_slabs_agree():
    """The DF-direct particle-particle ladder gives the same t2 whether W is built in one slab or many."""
    from pyscf_mpcc_b200 import rmpccsd
    p = make_problem(6, 14, 40, seed=2)
    ll, eris = _device_problem(p)
    eng = eris.eng
    t1, t2 = eng.tensor(p["t1"]), eng.tensor(p["t2"])
    tau = rmpccsd.make_tau(eng, t1, t2)
    one = rmpccsd._pp_ladder(eng, eris, t1, tau, eng.zeros(t2.shape))
    many = rmpccsd._pp_ladder(eng, eris, t1, tau, eng.zeros(t2.shape), slab_bytes=8 * 14 ** 3 * 3)
    ref = np.einsum("abcd,ijcd->ijab",
                    np.einsum("Lac,Lbd->abcd", p["Lvv"], p["Lvv"])
                    - np.einsum("kdac,kb->abcd", np.einsum("Lkd,Lac->kdac", p["Lov"], p["Lvv"]), p["t1"])
                    - np.einsum("kcbd,ka->abcd", np.einsum("Lkc,Lbd->kcbd", p["Lov"], p["Lvv"]), p["t1"]),
                    p["t2"] + np.einsum("ia,jb->ijab", p["t1"], p["t1"]), optimize=True)
    assert np.abs(eng.numpy(one) - ref).max() <= 1e-11 and np.abs(eng.numpy(many) - ref).max() <= 1e-11


@pytest.mark.parametrize("case", [((7, 5, 3), (2, 0, 1)), ((4, 9, 6, 5), (3, 1, 0, 2)), ((33, 70), (1, 0)),
                                  ((3, 40, 2, 37), (1, 3, 0, 2)), ((5, 6, 7, 8, 3), (4, 2, 0, 3, 1)), ((11,), (0,)),
                                  ((2, 3, 4, 5, 3, 2), (5, 0, 3, 1, 4, 2)), ((64, 65, 3), (1, 0, 2))])
def test_permute_axpby_kernel(case):
    """mpccb_permute_axpby (generic and tiled variants) against numpy transpose, with accumulate."""
    import torch
    from pyscf_mpcc_b200 import _cabi
    from pyscf_mpcc_b200.tensor_engine import CudaBackend
    dims, perm = case
    ctx = _cabi.Context(2, 3, 4, 0, 3, 0)
    ctx.call("mpccb_set_stream", torch.cuda.current_stream().cuda_stream)
    be = CudaBackend(ctx, torch.device("cuda", 0))
    rng = np.random.default_rng(len(dims))
    a = rng.standard_normal(dims)
    c0 = rng.standard_normal(tuple(dims[k] for k in perm))
    src, dst = be.upload(a), be.upload(c0)
    be.permute_axpby(src, dst, dims, perm, 1.5, -0.5)
    assert np.abs(be.download(dst).reshape(c0.shape) - (1.5 * a.transpose(perm) - 0.5 * c0)).max() <= 4e-16 * 8   # FMA contraction
    be.permute_axpby(src, dst, dims, perm, 1.0, 0.0)
    assert np.array_equal(be.download(dst).reshape(c0.shape), a.transpose(perm))


@pytest.mark.parametrize("mnk", [(200, 300, 1), (256, 512, 1024), (1000, 130, 96), (64, 64, 4096), (129, 1025, 250)])
def test_gemm_nt_dispatch(mnk):
    """mpccb_gemm_nt: TMA-fed kernel for the large aligned products, generic DMMA kernel otherwise."""
    import torch
    from pyscf_mpcc_b200 import _cabi
    m, n, k = mnk
    ctx = _cabi.Context(2, 3, 4, 0, 3, 0)
    ctx.call("mpccb_set_stream", torch.cuda.current_stream().cuda_stream)
    g = torch.Generator(device="cuda").manual_seed(m + n + k)
    a = torch.randn(m, k, dtype=torch.float64, device="cuda", generator=g)
    b = torch.randn(n, k, dtype=torch.float64, device="cuda", generator=g)
    c = torch.full((m, n), float("nan"), dtype=torch.float64, device="cuda")
    ctx.call("mpccb_gemm_nt", a.data_ptr(), b.data_ptr(), c.data_ptr(), m, n, k)
    torch.cuda.synchronize()
    ref = a @ b.T
    assert torch.isfinite(c).all()
    assert (c - ref).abs().max().item() <= 1e-12 * max(1.0, ref.abs().max().item()) * k ** 0.5


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rmpccsd_cc2_*.npz"))), ids=os.path.basename)
def test_cc2_update_amps_against_reference_fixture(path):
    """cc.cc2 = True (rmpccsd_slow.py:115-133) on the device against the reference's own output."""
    from pyscf_mpcc_b200 import rmpccsd
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    ll, eris = _device_problem(p)
    t1n, t2n = rmpccsd.update_amps(types.SimpleNamespace(cc2=True), p["t1"], p["t2"], eris, list(g["act_hole"]),
                                   list(g["act_particle"]), list(g["idx_singles"]), list(g["idx_doubles"]))
    scale = max(1.0, float(np.abs(g["t2new_samples"]).max()))
    assert np.abs(t1n - g["t1new"]).max() <= 1e-10 * max(1.0, float(np.abs(g["t1new"]).max()))
    assert np.abs(t2n.reshape(-1)[::int(g["stride"])] - g["t2new_samples"]).max() <= 1e-10 * scale
    assert abs(np.linalg.norm(t2n) - float(g["t2new_norm"])) <= 1e-10 * float(g["t2new_norm"])


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rmpccsd_kernel_*.npz"))), ids=os.path.basename)
def test_driver_loop_on_device_against_fixture(path):
    """RMPCCSD.kernel on the GPU (device-resident amplitudes, VectorDIIS on packed device vectors) against the loop of
    ccsd.py:42-150 run around the reference's update_amps / energy / lib.diis.DIIS: identical cycle counts,
    |E_corr| <= 1e-8, amplitudes to 1e-10."""
    from pyscf_mpcc_b200 import rmpccsd
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    ll, eris = _device_problem(p)
    cc = rmpccsd.RMPCCSD(ll, fock=p["fock"])
    e, t1, t2 = cc.kernel(list(g["act_hole"]), list(g["act_particle"]), list(g["idx_singles"]), list(g["idx_doubles"]),
                          cc2=bool(g["cc2"]))
    hist = np.array(cc.history)
    assert cc.converged == bool(g["converged"]) and cc.cycles == len(g["history"])
    assert np.abs(hist[:, 1] - g["history"][:, 1]).max() <= 1e-8
    np.testing.assert_allclose(hist[:, 3], g["history"][:, 3], rtol=1e-5, atol=1e-10)
    assert abs(e - float(g["e_corr"])) <= 1e-8 and abs(cc.emp2 - float(g["emp2"])) <= 1e-10
    eng = eris.eng
    assert np.abs(eng.numpy(t1) - g["t1"]).max() <= 1e-10
    t2h = eng.numpy(t2)
    assert np.abs(t2h.reshape(-1)[::int(g["stride"])] - g["t2_samples"]).max() <= 1e-10
    assert abs(np.linalg.norm(t2h) - float(g["t2_norm"])) <= 1e-9


def test_vector_layouts_on_device():
    """amplitudes_to_vector / vector_to_amplitudes of the CCSD-type stage (ccsd.py:719-735) against the oracle."""
    from oracle import cc_oracle
    p = make_problem(5, 9, 12, seed=3)
    ll, eris = _device_problem(p)
    eng = eris.eng
    vec = eng.amplitudes_to_vector(eng.tensor(p["t1"]), eng.tensor(p["t2"]))
    want = cc_oracle.amplitudes_to_vector(p["t1"], p["t2"])
    assert np.array_equal(eng.be.download(vec), want)
    t1, t2 = eng.vector_to_amplitudes(vec, 5, 9)
    t1o, t2o = cc_oracle.vector_to_amplitudes(want, 14, 5)
    assert np.array_equal(eng.numpy(t1), t1o) and np.array_equal(eng.numpy(t2), t2o)


@pytest.mark.parametrize("cc2", [False, True])
def test_update_amps_with_a_t2_that_is_not_pair_symmetric(cc2):
    """The reference's update_amps on a t2 with t2[ijab] != t2[jiba] (fixture rmpccsd_asym_*.npz): the halved
    particle-particle ladder needs that symmetry, so the residual checks it and falls back to the full ladder."""
    from pyscf_mpcc_b200 import rmpccsd
    g = np.load(os.path.join(GOLDEN, "rmpccsd_asym_4_9_21.npz"))
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    t2 = p["t2"] + float(g["eps"]) * np.random.default_rng(int(g["noise_seed"])).standard_normal(p["t2"].shape)
    ll, eris = _device_problem(p)
    t1n, t2n = rmpccsd.update_amps(types.SimpleNamespace(cc2=cc2), p["t1"], t2, eris, list(g["act_hole"]),
                                   list(g["act_particle"]), list(g["idx_singles"]), list(g["idx_doubles"]))
    tag = "cc2" if cc2 else "ccsd"
    assert np.abs(t1n - g[f"t1new_{tag}"]).max() <= 1e-10
    assert np.abs(t2n - g[f"t2new_{tag}"]).max() <= 1e-10
