"""GPU parity tests (-m gpu): the CUDA path through the C ABI / MPCC_LL drop-in against
(1) the fixtures the reference produced (tests/golden) and (2) the oracle on the same seeded inputs.
Tolerances are north_star's: |dE| <= 1e-8 Eh, amplitudes max-abs <= 1e-10, identical iteration counts."""
import ctypes
import os

import numpy as np
import pytest

from conftest import golden_cases

pytestmark = pytest.mark.gpu

E_TOL = 1e-8
AMP_TOL = 1e-10
STRIDE = 97


def _solver(o, v, L, scale, seed=0, **kw):
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    mf = SyntheticMF(o, v, L, seed=seed, scale=scale)
    ll = MPCC_LL(mf, **kw)
    ll.verbose_print = False
    return mf, ll


def _load(path):
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    return g, o, v, L, float(g["scale"])


@pytest.mark.parametrize("path", golden_cases(), ids=os.path.basename)
def test_kernel_against_reference_fixture(path):
    g, o, v, L, scale = _load(path)
    mf, ll = _solver(o, v, L, scale)
    ll.con_tol, ll.max_its = 1e-6, 50
    e = ll.kernel()
    assert ll.iterations == int(g["iterations"])
    tr = np.array(ll.history)
    assert np.abs(tr[:, 1] - g["trace_dE"]).max() <= E_TOL
    np.testing.assert_allclose(tr[:, 2], g["trace_res"], rtol=1e-6, atol=1e-12)
    assert abs(e - g["trace_dE"][-1]) <= E_TOL
    assert np.abs(ll.t1 - g["final_t1"]).max() <= AMP_TOL
    t2 = np.asarray(ll.t2)
    assert t2.shape == (o, o, v, v)                       # (o,o,v,v) after DIIS, as the reference leaves it
    assert np.abs(t2.reshape(-1)[::STRIDE] - g["final_t2_samples"]).max() <= AMP_TOL
    assert abs(np.linalg.norm(t2) - g["final_t2_norm"]) <= 1e-9
    if "final_t2" in g:
        assert np.abs(t2 - g["final_t2"]).max() <= AMP_TOL


@pytest.mark.parametrize("path", golden_cases(), ids=os.path.basename)
@pytest.mark.parametrize("tag", ["zero", "rand"])
def test_updated_amps_known_answers(path, tag):
    g, o, v, L, scale = _load(path)
    mf, ll = _solver(o, v, L, scale)
    ll.t1 = g[f"kat_{tag}_t1_in"]
    res, dE, t1n, t2 = ll.updated_amps()
    assert abs(dE - g[f"kat_{tag}_dE"]) <= E_TOL
    assert abs(res - g[f"kat_{tag}_res"]) <= 1e-10 * max(1.0, float(g[f"kat_{tag}_res"]))
    assert np.abs(t1n - g[f"kat_{tag}_t1_new"]).max() <= AMP_TOL
    assert t2.shape == (v, o, v, o)                       # (v,o,v,o) from updated_amps
    t2h = np.asarray(t2)
    assert np.abs(t2h.reshape(-1)[::STRIDE] - g[f"kat_{tag}_t2_samples"]).max() <= AMP_TOL
    vec = ll.amplitudes_to_vector(t1n, t2).numpy()
    assert np.abs(vec[::STRIDE] - g[f"kat_{tag}_vec_samples"]).max() <= AMP_TOL
    assert abs(np.linalg.norm(vec) - g[f"kat_{tag}_vec_norm"]) <= 1e-9
    # energy() alias = second return value; update_amps alias
    assert abs(ll.energy() - dE) <= 1e-13
    assert ll.update_amps.__func__ is ll.updated_amps.__func__


@pytest.mark.parametrize("split", [2, 3, 5])
def test_yvo_split_k_matches_oracle(split):
    """Split-K form of the TMA Yvo GEMM (used when a rank's slab has few tiles) against the oracle."""
    from oracle.mpcc_oracle import LowLevelOracle
    o, v, L = 32, 70, 129
    mf, ll = _solver(o, v, L, 0.05, seed=4, yvo_split=split)
    orc = LowLevelOracle(mf, mode="blas")
    t1 = 0.03 * np.random.default_rng(12).standard_normal((o, v))
    ll.t1 = t1
    res, dE, t1n, t2 = ll.updated_amps()
    res_o, dE_o, t1n_o, t2_o = orc.updated_amps(t1)
    Y = np.einsum("aibj,Ljb->aiL", t2_o, orc.Lov, optimize=True).reshape(v * o, L)
    assert np.abs(ll.debug_buffer("Y").reshape(v * o, L) - Y).max() <= 1e-12 * max(1.0, np.abs(Y).max())
    assert abs(dE - dE_o) <= E_TOL and np.abs(t1n - t1n_o).max() <= AMP_TOL


@pytest.mark.parametrize("variant", ["default", "wide", "unfused", "unfused_no_sym", "unfused_no_tma", "no_tma", "no_sym",
                                     "wide_no_sym"])
def test_t2_kernel_variants_agree_with_fixture(variant):
    """All forms of the eris -> t2 step (128x64 two-CTA tiles, 128x128 tiles, unfused GEMM + in-place pass,
    generic non-TMA kernel) reproduce the reference fixture."""
    g, o, v, L, scale = _load([p for p in golden_cases() if "12_64_192" in p][0])
    kw = {"default": {}, "wide": {"t2_wide_tiles": True}, "unfused": {"t2_unfused": True}, "unfused_no_sym": {"t2_unfused": True, "t2_sym": False},
          "unfused_no_tma": {"t2_unfused": True, "use_tma": False}, "no_tma": {"use_tma": False},
          "no_sym": {"t2_sym": False}, "wide_no_sym": {"t2_sym": False, "t2_wide_tiles": True}}[variant]
    mf, ll = _solver(o, v, L, scale, **kw)
    e = ll.kernel()
    assert ll.iterations == int(g["iterations"]) and abs(e - g["trace_dE"][-1]) <= E_TOL
    assert np.abs(ll.t1 - g["final_t1"]).max() <= AMP_TOL
    assert np.abs(np.asarray(ll.t2).reshape(-1)[::STRIDE] - g["final_t2_samples"]).max() <= AMP_TOL


@pytest.mark.parametrize("shape", [(5, 19, 116, 0.1), (7, 33, 65, 0.15), (16, 48, 150, 0.1), (33, 70, 129, 0.05),
                                   (64, 66, 130, 0.05), (80, 90, 100, 0.04), (128, 130, 64, 0.03), (131, 140, 48, 0.03),
                                   (48, 50, 70, 0.05), (100, 102, 40, 0.03)])
def test_intermediates_against_oracle(shape):
    """Every stage of updated_amps vs the oracle on the same seeded inputs (odd sizes, occupied counts on
    both sides of the fused-epilogue limit of 128, badly tiling nocc = 80 (unfused path), naux not a multiple of the K
    tile; nocc = 48, 80, 100: Yvo through power-of-two pieces of the occupied index, 32+16, 64+16, 64+32+4)."""
    from oracle.mpcc_oracle import LowLevelOracle
    o, v, L, scale = shape
    mf, ll = _solver(o, v, L, scale, seed=3)
    orc = LowLevelOracle(mf, mode="blas")
    t1 = 0.03 * np.random.default_rng(11).standard_normal((o, v))
    ll.t1 = t1
    res, dE, t1n, t2 = ll.updated_amps()
    res_o, dE_o, t1n_o, t2_o = orc.updated_amps(t1)
    Lov, Lvv, Loo = orc.Lov, orc.Lvv, orc.Loo
    Xvo = np.einsum("Lab,ib->Lai", Lvv, t1, optimize=True)
    Xoo = np.einsum("Lia,ja->Lij", Lov, t1, optimize=True)
    Joo = Xoo + Loo
    Jvo = Xvo + Lov.transpose(0, 2, 1) - np.einsum("Lij,ja->Lai", Joo, t1, optimize=True)
    Y = np.einsum("aibj,Ljb->Lai", t2_o, Lov, optimize=True)
    B = ll.debug_buffer
    ldk = (L + 1) // 2 * 2

    def close(a, b, name):
        scale_ = max(1.0, np.abs(b).max())
        assert np.abs(a - b).max() <= 1e-12 * scale_, name

    close(B("Xov").reshape(L, o, v), Xvo.transpose(0, 2, 1), "Xvo")
    close(B("Xoo").reshape(L, o, o), Xoo, "Xoo")
    close(B("Joo").reshape(L, o, o), Joo, "Joo")
    close(B("X"), np.einsum("Lii->L", Xoo), "X")
    close(B("JK").reshape(v * o, ldk)[:, :L], Jvo.transpose(1, 2, 0).reshape(v * o, L), "Jvo")
    close(B("T").reshape(v, v, o, o), t2_o.transpose(0, 2, 1, 3), "t2")
    close(B("Y").reshape(v * o, L), Y.transpose(1, 2, 0).reshape(v * o, L), "Yvo")
    assert abs(dE - dE_o) <= E_TOL and abs(res - res_o) <= 1e-10 * max(1, res_o)
    assert np.abs(t1n - t1n_o).max() <= AMP_TOL
    assert np.abs(np.asarray(t2) - t2_o).max() <= AMP_TOL


def test_dropin_factory_and_print_format(capsys):
    """mpcc.MPCC(mf): construction runs the solve with con_tol 1e-6 / max_its 50 (pyscf/mpcc/mpcc.py:13-17)
    and prints the reference's progress lines (dfrmpcc_lowlevel.py:95,97)."""
    import pyscf_mpcc_b200
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    g = np.load(golden_cases()[0])
    o, v, L = (int(x) for x in g["shape"])
    obj = pyscf_mpcc_b200.MPCC(SyntheticMF(o, v, L, seed=0, scale=float(g["scale"])))
    out = capsys.readouterr().out.strip().splitlines()
    assert obj.lowlevel.con_tol == 1e-6 and obj.lowlevel.max_its == 50
    assert out[-1] == "In Kernel!"
    its = [l for l in out if l.startswith("It ")]
    assert len(its) == int(g["iterations"])
    k = len(its)
    assert its[-1] == f"It {k} Energy progress {g['trace_dE'][-1]:.6e} residual progress {g['trace_res'][-1]:.6e}"
    assert abs(obj.kernel() - g["trace_dE"][-1]) <= E_TOL and obj.with_df is not None


def test_reference_seam_accepts_the_module():
    """The reference's own solver object only needs `lowlevel.MPCC_LL(mf)` + attribute pokes (mpcc.py:13-17)."""
    import pyscf_mpcc_b200.lowlevel as module
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    mf = SyntheticMF(4, 10, 30, seed=0, scale=0.3)
    low = module.MPCC_LL(mf)
    low.verbose_print = False
    low.con_tol = 1e-6
    low.max_its = 50
    e = low.kernel()
    assert abs(e - (-7.1793901067873866e-01)) <= E_TOL and low.iterations == 9
    for name in ("mf", "with_df", "t1", "t2", "diis", "nao", "nocc", "nvir", "naux", "mo_energy", "mo_coeff",
                 "nvir_pair", "Loo", "Lov", "Lvo", "Lvv", "foo", "fvv", "eia", "D"):
        assert hasattr(low, name), name
    assert low.Lov.shape == (30, 4, 10) and low.Lvo.shape == (30, 10, 4) and low.D.shape == (10, 4, 10, 4)
    np.testing.assert_array_equal(low.Lvo, low.Lov.transpose(0, 2, 1))


def test_diis_off_and_max_its():
    """diis=False keeps t2 (v,o,v,o) and does not converge on these problems: loop stops at max_its."""
    from oracle.mpcc_oracle import LowLevelOracle
    mf, ll = _solver(5, 19, 116, 0.1)
    ll.diis = False
    ll.max_its = 7
    e = ll.kernel()
    orc = LowLevelOracle(mf, mode="blas"); orc.diis = False; orc.max_its = 7
    e_o = orc.kernel()
    assert ll.iterations == 7 == orc.iterations
    assert abs(e - e_o) <= E_TOL and np.abs(ll.t1 - orc.t1).max() <= AMP_TOL
    assert ll.t2.shape == (19, 5, 19, 5)
    assert np.abs(np.asarray(ll.t2) - orc.t2).max() <= AMP_TOL


def test_host_pointer_cabi_call():
    """mpccb_updated_amps_host: the reference call with host arrays (t1 in; t1_new, res, dE out)."""
    from oracle.mpcc_oracle import LowLevelOracle
    mf, ll = _solver(8, 40, 120, 0.1)
    orc = LowLevelOracle(mf, mode="blas")
    t1 = 0.02 * np.random.default_rng(2).standard_normal((8, 40))
    t1n = np.empty_like(t1)
    res, dE = ctypes.c_double(), ctypes.c_double()
    ll._ctx.call("mpccb_updated_amps_host", t1.ctypes.data, t1n.ctypes.data, ctypes.byref(res), ctypes.byref(dE))
    r_o, e_o, t_o, _ = orc.updated_amps(t1)
    assert abs(dE.value - e_o) <= E_TOL and abs(res.value - r_o) <= 1e-10 and np.abs(t1n - t_o).max() <= AMP_TOL


@pytest.mark.parametrize("mnk", [(1, 1, 1), (17, 5, 3), (128, 128, 16), (129, 257, 70), (300, 64, 1000), (64, 300, 33)])
def test_dmma_gemm_against_fp64_matmul(mnk):
    import torch
    from pyscf_mpcc_b200 import _cabi
    m, n, k = mnk
    ctx = _cabi.Context(2, 3, 4, 0, 3, 0)
    g = torch.Generator(device="cuda").manual_seed(m * 7 + n)
    a = torch.randn(m, k, dtype=torch.float64, device="cuda", generator=g)
    b = torch.randn(n, k, dtype=torch.float64, device="cuda", generator=g)
    c = torch.full((m, n), float("nan"), dtype=torch.float64, device="cuda")
    ctx.call("mpccb_set_stream", torch.cuda.current_stream().cuda_stream)
    ctx.call("mpccb_gemm_nt", a.data_ptr(), b.data_ptr(), c.data_ptr(), m, n, k)
    torch.cuda.synchronize()
    ref = a @ b.T
    assert torch.isfinite(c).all()
    assert (c - ref).abs().max().item() <= 1e-12 * max(1.0, ref.abs().max().item()) * max(1, k) ** 0.5


@pytest.mark.parametrize("shape", [(5, 19, 116), (16, 48, 150), (70, 90, 250)])
def test_device_ao2mo_with_rotated_orbitals(shape):
    """compute_three_center_ints on the device (unpack_tril + two GEMM passes + slicing, :46-69) with a
    non-trivial mo_coeff, several with_df.loop() blocks (naux > 240) and a ragged last block."""
    from oracle.mpcc_oracle import LowLevelOracle
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    o, v, L = shape
    mf = SyntheticMF(o, v, L, seed=5, scale=0.1)
    q, _ = np.linalg.qr(np.random.default_rng(9).standard_normal((o + v, o + v)))
    mf.mo_coeff = q
    mf.get_fock = lambda: q @ np.diag(mf.mo_energy) @ q.T
    ll = MPCC_LL(mf)
    ll.verbose_print = False
    orc = LowLevelOracle(mf, mode="blas")
    for name in ("Loo", "Lov", "Lvv"):
        ref = getattr(orc, name)
        assert np.abs(getattr(ll, name) - ref).max() <= 1e-12 * max(1.0, np.abs(ref).max()), name
    assert np.abs(ll.foo - orc.foo).max() <= 1e-12 and np.abs(ll.fvv - orc.fvv).max() <= 1e-12
    e = ll.kernel()
    e_o = orc.kernel()
    assert ll.iterations == orc.iterations and abs(e - e_o) <= E_TOL
    assert np.abs(ll.t1 - orc.t1).max() <= AMP_TOL


@pytest.mark.parametrize("shape", [(1, 1, 1, 0.3), (1, 3, 4, 0.3), (2, 1, 3, 0.3), (3, 2, 1, 0.3), (1, 40, 17, 0.2),
                                   (6, 300, 20, 0.05), (40, 41, 33, 0.05)])
def test_degenerate_and_ragged_shapes(shape):
    """Smallest possible problems (one occupied / one virtual / one auxiliary function), shapes far from
    any tile multiple: full kernel() against the oracle (iteration count, energy, amplitudes)."""
    from oracle.mpcc_oracle import LowLevelOracle
    o, v, L, scale = shape
    mf, ll = _solver(o, v, L, scale, seed=7)
    ll.max_its = 12
    e = ll.kernel()
    orc = LowLevelOracle(mf, mode="blas")
    orc.max_its = 12
    e_o = orc.kernel()
    assert ll.iterations == orc.iterations
    assert abs(e - e_o) <= E_TOL and np.abs(ll.t1 - orc.t1).max() <= AMP_TOL
    assert np.abs(np.asarray(ll.t2) - orc.t2).max() <= AMP_TOL
    for (k, de, r), (k2, de2, r2) in zip(ll.history, orc.history):
        assert abs(de - de2) <= E_TOL and abs(r - r2) <= 1e-9 * max(1.0, r2)


def test_errors_are_loud():
    """Bad arguments raise (int status + message through the C ABI), nothing falls back silently."""
    from pyscf_mpcc_b200 import _cabi
    with pytest.raises(_cabi.MpccError):
        _cabi.Context(0, 4, 4, 0, 4, 0)                      # nocc must be positive
    with pytest.raises(_cabi.MpccError):
        _cabi.Context(2, 4, 4, 3, 2, 0)                      # empty slab
    ctx = _cabi.Context(2, 4, 4, 0, 4, 0)
    with pytest.raises(_cabi.MpccError):
        ctx.call("mpccb_dress", None)                        # workspace/integrals not bound
    with pytest.raises(_cabi.MpccError):
        ctx.call("mpccb_set_option", b"no_such_option", 1)
    import pyscf_mpcc_b200

    class UHF:                                               # the factory rejects unrestricted mean fields
        pass
    with pytest.raises(RuntimeError, match="UMPCC"):
        pyscf_mpcc_b200.MPCC(UHF())
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    mf = SyntheticMF(2, 3, 4)
    mf.with_df = None
    with pytest.raises(RuntimeError):
        pyscf_mpcc_b200.MPCC(mf)


@pytest.mark.parametrize("path", golden_cases(), ids=os.path.basename)
def test_column_owned_diis_vector_reproduces_fixture(path):
    """The balanced multi-GPU form of the DIIS vector (mpccb_bind_diis_columns), forced on one rank: same
    iteration counts, energies and amplitudes as the reference, and the same flat vector after the gather."""
    g, o, v, L, scale = _load(path)
    mf, ll = _solver(o, v, L, scale, diis_layout="columns")
    e = ll.kernel()
    assert ll.iterations == int(g["iterations"])
    tr = np.array(ll.history)
    assert np.abs(tr[:, 1] - g["trace_dE"]).max() <= E_TOL
    np.testing.assert_allclose(tr[:, 2], g["trace_res"], rtol=1e-6, atol=1e-12)
    assert np.abs(ll.t1 - g["final_t1"]).max() <= AMP_TOL
    t2 = np.asarray(ll.t2)
    assert np.abs(t2.reshape(-1)[::STRIDE] - g["final_t2_samples"]).max() <= AMP_TOL
    assert abs(np.linalg.norm(t2) - g["final_t2_norm"]) <= 1e-9
    ll.t1 = g["kat_rand_t1_in"]
    res, dE, t1n, t2h = ll.updated_amps()
    vec = ll.amplitudes_to_vector(t1n, t2h).numpy()
    assert np.abs(vec[::STRIDE] - g["kat_rand_vec_samples"]).max() <= AMP_TOL
    assert abs(np.linalg.norm(vec) - g["kat_rand_vec_norm"]) <= 1e-9


@pytest.mark.parametrize("shape", [(6, 5, 9, 0.1), (16, 48, 40, 0.1), (33, 70, 30, 0.05), (64, 66, 40, 0.05)])
def test_column_owned_vector_equals_row_owned_vector(shape):
    """Bitwise: the column-owned vector, scattered to reference order, is the row-owned vector (nocc > nvir,
    blocks that straddle matrix rows, nocc not a multiple of the 8 x 4 thread tile)."""
    o, v, L, scale = shape
    t1 = 0.03 * np.random.default_rng(5).standard_normal((o, v))
    vecs = []
    for layout in ("rows", "columns"):
        mf, ll = _solver(o, v, L, scale, seed=2, diis_layout=layout)
        ll.t1 = t1
        res, dE, t1n, t2h = ll.updated_amps()
        vecs.append(ll.amplitudes_to_vector(t1n, t2h).numpy())
    assert vecs[0].shape == vecs[1].shape == (o * v + (o * v) * (o * v + 1) // 2,)
    assert np.array_equal(vecs[0], vecs[1])


def test_config2_kernel_against_the_references_own_run():
    """BASELINE.json configs[1]'s shape (benzene/cc-pVTZ: 21, 243, 654): iteration count, per-iteration dE / res,
    final t1 and t2 samples against the fixture the REFERENCE's files produced (generate_golden_config2.py)."""
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "config2_mpcc_21_243_654.npz"))
    o, v, L = (int(x) for x in g["shape"])
    mf, ll = _solver(o, v, L, float(g["scale"]))
    ll.con_tol, ll.max_its = 1e-6, 50
    e = ll.kernel()
    assert ll.iterations == int(g["iterations"])
    tr = np.array(ll.history)
    assert np.abs(tr[:, 1] - g["trace_dE"]).max() <= E_TOL
    np.testing.assert_allclose(tr[:, 2], g["trace_res"], rtol=1e-6, atol=1e-12)
    assert abs(e - g["trace_dE"][-1]) <= E_TOL
    assert np.abs(ll.t1 - g["final_t1"]).max() <= AMP_TOL
    t2 = np.asarray(ll.t2)
    s = int(g["sample_stride"])
    assert np.abs(t2.reshape(-1)[::s] - g["final_t2_samples"]).max() <= AMP_TOL
    assert abs(np.linalg.norm(t2) - g["final_t2_norm"]) <= 1e-9
    assert abs(t2.sum() - g["final_t2_sum"]) <= 1e-8
    # first iteration from t1 = 0: t1_new, (v,o,v,o) t2 and the DIIS vector
    ll.t1 = np.zeros((o, v))
    res, dE, t1n, t2a = ll.updated_amps()
    assert np.abs(t1n - g["it1_t1_new"]).max() <= AMP_TOL
    t2h = np.asarray(t2a)
    assert np.abs(t2h.reshape(-1)[::s] - g["it1_t2_samples"]).max() <= AMP_TOL
    vec = ll.amplitudes_to_vector(t1n, t2a).numpy()
    assert np.abs(vec[::s] - g["it1_vec_samples"]).max() <= AMP_TOL
    assert abs(np.linalg.norm(vec) - g["it1_vec_norm"]) <= 1e-9


def test_the_references_solver_object_drives_this_module(capsys):
    """pyscf/mpcc/mpcc.py:5-17 executed UNMODIFIED (staged copy, oracle/make_ref.py): ``MPCC(mf, module)`` builds
    ``module.MPCC_LL(mf)``, sets con_tol / max_its from outside and calls ``kernel()``. Passing
    ``pyscf_mpcc_b200.lowlevel`` makes the reference's own solver object run on the GPU."""
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("no reference files (neither /root/reference nor oracle/_ref)")
    import pyscf_mpcc_b200.lowlevel as b200_lowlevel
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    ref_mpcc = ref_shim.load_reference_solver()
    g, o, v, L, scale = _load([p for p in golden_cases() if "8_40_120" in p][0])
    mf = SyntheticMF(o, v, L, seed=0, scale=scale)
    obj = ref_mpcc.MPCC(mf, b200_lowlevel)                      # the reference's class, our module
    assert type(obj).__module__ == "pyscf.mpcc.mpcc"
    ll = obj.lowlevel
    assert isinstance(ll, b200_lowlevel.MPCC_LL)
    assert ll.con_tol == 1e-6 and ll.max_its == 50 and obj._scf is mf
    assert ll.iterations == int(g["iterations"])
    assert abs(ll.history[-1][1] - g["trace_dE"][-1]) <= E_TOL
    assert np.abs(ll.t1 - g["final_t1"]).max() <= AMP_TOL
    lines = [x for x in capsys.readouterr().out.splitlines() if x.startswith("It ")]
    assert len(lines) == ll.iterations
    # and the reference's own low-level module behind the same object gives the same numbers
    ref_ll, _ = ref_shim.load_reference()
    obj_ref = ref_mpcc.MPCC(mf, ref_ll)
    assert abs(obj_ref.lowlevel.t1 - ll.t1).max() <= AMP_TOL


def test_to_gpu_of_a_constructed_reference_object():
    """``lowlevel.to_gpu(ref_ll)`` / ``MPCC_LL.from_cpu`` (PySCF's ``obj.to_gpu()`` convention,
    pyscf/lib/misc.py:1508-1556): the reference's own CPU object, constructed by its unmodified file, hands its MO
    blocks and outside pokes to the GPU solver, which then reproduces the reference's ``kernel()``."""
    from oracle import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("no reference files (neither /root/reference nor oracle/_ref)")
    import pyscf_mpcc_b200.lowlevel as b200_lowlevel
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    ref_ll, _ = ref_shim.load_reference()
    g, o, v, L, scale = _load([p for p in golden_cases() if "5_19_116" in p][0])
    mf = SyntheticMF(o, v, L, seed=0, scale=scale)
    cpu = ref_ll.MPCC_LL(mf)
    cpu.con_tol, cpu.max_its = 1e-6, 50
    gpu = b200_lowlevel.to_gpu(cpu)
    assert isinstance(gpu, b200_lowlevel.MPCC_LL) and gpu.mf is mf and gpu.with_df is mf.with_df
    assert (gpu.nocc, gpu.nvir, gpu.naux) == (o, v, L) and gpu.con_tol == 1e-6 and gpu.max_its == 50
    gpu.verbose_print = False
    dE = gpu.kernel()
    assert gpu.iterations == int(g["iterations"])
    assert abs(dE - g["trace_dE"][-1]) <= E_TOL
    assert np.abs(gpu.t1 - g["final_t1"]).max() <= AMP_TOL
    # a t1 already present on the CPU object is carried over: one more residual evaluation from the converged t1
    cpu.t1 = gpu.t1.copy()
    again = b200_lowlevel.MPCC_LL.from_cpu(cpu)
    res, dE2, _, _ = again.updated_amps()
    res_ref, dE_ref, _, _ = cpu.updated_amps()
    assert abs(dE2 - dE_ref) <= E_TOL and abs(res - res_ref) <= 1e-10


@pytest.mark.parametrize("layout", ["rows", "columns"])
def test_host_vector_round_trip(layout):
    """vector_to_amplitudes(amplitudes_to_vector(t1, t2).numpy()): a HOST vector is the reference-order flat vector,
    whatever the device-side ownership of DIIS vectors is (:162-176)."""
    from oracle.mpcc_oracle import vector_to_amps
    o, v, L = 6, 11, 20
    mf, ll = _solver(o, v, L, 0.1, seed=2, diis_layout=layout)
    ll.t1 = 0.05 * np.random.default_rng(3).standard_normal((o, v))
    res, dE, t1n, t2 = ll.updated_amps()
    flat = ll.amplitudes_to_vector(t1n, t2).numpy()
    assert flat.shape == (o * v + o * v * (o * v + 1) // 2,)
    t1b, t2b = ll.vector_to_amplitudes(flat)
    t1o, t2o = vector_to_amps(flat, o, v)
    assert np.array_equal(t1b, t1o)
    assert np.array_equal(np.asarray(t2b), t2o)
    with pytest.raises(ValueError):
        ll.vector_to_amplitudes(flat[:-1])


def test_extrapolated_t2_handle_goes_stale_instead_of_changing():
    """run_diis returns a t2 that aliases the DIIS extrapolation buffer; the next update overwrites it in place. The
    reference returns a fresh array (diis.py:268-275), so an old handle must raise, not silently show new values."""
    from pyscf_mpcc_b200.diis import DIIS
    o, v, L = 5, 9, 16
    mf, ll = _solver(o, v, L, 0.1, seed=1)
    ll.t1 = np.zeros((o, v))
    ad = DIIS()
    handles = []
    for _ in range(3):
        res, dE, t1n, t2 = ll.updated_amps()
        ll.t1, t2x = ll.run_diis(t1n, t2, ad)
        handles.append(t2x)
    assert np.asarray(handles[-1]).shape == (o, o, v, v)
    with pytest.raises(RuntimeError, match="stale"):
        np.asarray(handles[0])
    with pytest.raises(RuntimeError, match="stale"):
        np.asarray(handles[1])


def test_fast_div_within_one_ulp_of_ieee_division():
    """The fused t2 epilogue replaces the IEEE division of :137 by MUFU.RCP64H + Newton + residual correction
    (gemm_tma.cuh fast_div / fast_div_batch). Swept against numpy's x / y over the denominator range of the
    fixtures (eo+eo-ev-ev in [-10, -1.6]), near-degenerate denominators, and a wide exponent range."""
    import torch
    mf, ll = _solver(4, 6, 8, 0.1)
    rng = np.random.default_rng(0)
    n = 1 << 21
    dens = [-(1.6 + 8.4 * rng.random(n)),                             # fixture range
            -(10.0 ** rng.uniform(-9, -1, n)),                        # near-degenerate orbital energies
            np.ldexp(1.0 + rng.random(n), rng.integers(-300, 300, n)) * rng.choice([-1.0, 1.0], n)]
    nums = [rng.standard_normal(n) * 10.0 ** rng.uniform(-12, 2, n),
            rng.standard_normal(n) * 1e-3,
            np.ldexp(1.0 + rng.random(n), rng.integers(-300, 300, n)) * rng.choice([-1.0, 1.0], n)]
    worst = 0
    for x, y in zip(nums, dens):
        ok = np.isfinite(x / y) & (np.abs(x / y) > 1e-290) & (np.abs(x / y) < 1e290)
        x, y = x[ok], y[ok]
        xd, yd = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
        qd = torch.empty_like(xd)
        ll._ctx.call("mpccb_test_fast_div", xd.data_ptr(), yd.data_ptr(), qd.data_ptr(), x.size)
        torch.cuda.synchronize()
        q = qd.cpu().numpy()
        want = x / y
        ulps = np.abs(q.view(np.int64) - want.view(np.int64))
        assert np.all(np.sign(q) == np.sign(want))
        worst = max(worst, int(ulps.max()))
    assert worst <= 1, f"fast_div is {worst} ulp away from IEEE division"


def test_near_degenerate_orbital_energies():
    """HOMO-LUMO gap of 1e-3 Eh: denominators down to 2e-3, amplitudes ~500x larger; one residual evaluation
    against the oracle to 1e-10 relative to the largest amplitude."""
    from oracle.mpcc_oracle import LowLevelOracle
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    o, v, L = 16, 40, 64
    mf = SyntheticMF(o, v, L, seed=6, scale=0.05)
    Loo, Lov, Lvv = mf.mo_blocks()
    eo = np.sort(np.concatenate([np.linspace(-1.0, -0.6, o - 2), [-0.5005, -0.5]]))
    ev = np.sort(np.concatenate([[-0.499, -0.4985], np.linspace(0.3, 2.0, v - 2)]))
    blocks = (Loo, Lov, Lvv, np.diag(eo), np.diag(ev), np.concatenate([eo, ev]))
    ll = MPCC_LL(blocks=blocks)
    orc = LowLevelOracle(blocks=blocks, mode="blas")
    t1 = 0.01 * np.random.default_rng(1).standard_normal((o, v))
    ll.t1 = t1
    res, dE, t1n, t2 = ll.updated_amps()
    res_o, dE_o, t1n_o, t2_o = orc.updated_amps(t1)
    scale = np.abs(t2_o).max()
    assert scale > 0.5                                     # dominated by the 2e-3 denominators
    assert np.abs(np.asarray(t2) - t2_o).max() <= 1e-10 * scale
    assert np.abs(t1n - t1n_o).max() <= 1e-10 * max(1.0, np.abs(t1n_o).max())
    assert abs(dE - dE_o) <= 1e-8 * max(1.0, abs(dE_o))
