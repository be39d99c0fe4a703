"""CPU checks of the CCSD-residual transcription (rows f1/f2): pyscf_mpcc_b200/rmpccsd.py driven through a
NumPy stand-in backend against the REFERENCE's own rmpccsd_slow.update_amps / energy executed unmodified
through oracle/ref_shim.py (build container), and against the committed fixture tests/golden/rmpccsd_*.npz."""
import glob
import os
import types

import numpy as np
import pytest

from conftest import GOLDEN
from cc_problem import make_problem, reference_eris
from np_backend import NumpyBackend
from oracle import ref_shim
from pyscf_mpcc_b200 import rmpccsd
from pyscf_mpcc_b200.tensor_engine import Engine

needs_ref = pytest.mark.skipif(not ref_shim.full_tree_available(), reason="full reference tree not present")


def _ours(p, act_h, act_p, idx_s, idx_d):
    eng = Engine(NumpyBackend(p["eo"], p["ev"]))
    eris = rmpccsd.DFERIs.from_numpy(eng, p["Loo"], p["Lov"], p["Lvv"], p["fock"])
    cc = types.SimpleNamespace(cc2=False)
    t1n, t2n = rmpccsd.update_amps(cc, p["t1"], p["t2"], eris, act_h, act_p, idx_s, idx_d)
    return t1n, t2n, rmpccsd.energy(cc, p["t1"], p["t2"], eris), eng


@needs_ref
@pytest.mark.parametrize("case", [(3, 5, 11, [0, 2], [1, 3], [0, 2], [1, 5, 10, 12]),
                                  (4, 7, 13, [1], [0, 2, 6], [], []),
                                  (2, 6, 9, [0, 1], [0, 1, 2, 3, 4, 5], [1], list(range(15)))])
def test_update_amps_against_reference_code(case):
    o, v, L, act_h, act_p, idx_s, idx_d = case
    p = make_problem(o, v, L, seed=o * 10 + v)
    slow, _ = ref_shim.load_reference_cc()
    cc = types.SimpleNamespace(cc2=False)
    er = reference_eris(p)
    t1r, t2r = slow.update_amps(cc, p["t1"], p["t2"], er, act_h, act_p, idx_s, idx_d)
    e_r = slow.energy(cc, p["t1"], p["t2"], er)
    t1n, t2n, e, eng = _ours(p, act_h, act_p, idx_s, idx_d)
    assert np.abs(t1n - t1r).max() <= 1e-12
    assert np.abs(t2n - t2r).max() <= 1e-12
    assert abs(e - e_r) <= 1e-12
    assert eng.gemms > 50


def test_class_numbering_matches_get_index_tuples():
    o, v = 5, 7
    act_h, act_p = [1, 3], [0, 2, 5]
    singles, doubles = rmpccsd.get_index_tuples(act_h, act_p, o, v)
    ah = np.zeros(o, int); ah[act_h] = 1
    ap = np.zeros(v, int); ap[act_p] = 1
    for s, ix in enumerate(singles):
        m = np.zeros((o, v), bool); m[ix] = True
        cls = 2 * ah[:, None] + ap[None, :]
        assert np.array_equal(m, cls == rmpccsd._SINGLES_BITS[s])
    cls = (8 * ah[:, None, None, None] + 4 * ah[None, :, None, None] + 2 * ap[None, None, :, None] + ap[None, None, None, :])
    seen = set()
    for d, ix in enumerate(doubles):
        m = np.zeros((o, o, v, v), bool); m[ix] = True
        assert np.array_equal(m, cls == rmpccsd._DOUBLES_BITS[d]), d
        seen.add(rmpccsd._DOUBLES_BITS[d])
    assert seen == set(range(16))


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rmpccsd_[0-9]*.npz"))), ids=os.path.basename)
def test_update_amps_against_fixture(path):
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    t1n, t2n, e, _ = _ours(p, list(g["act_hole"]), list(g["act_particle"]), list(g["idx_singles"]), list(g["idx_doubles"]))
    assert np.abs(t1n - g["t1new"]).max() <= 1e-12
    assert np.abs(t2n.reshape(-1)[::int(g["stride"])] - g["t2new_samples"]).max() <= 1e-12
    assert abs(np.linalg.norm(t2n) - float(g["t2new_norm"])) <= 1e-11 and abs(e - float(g["energy"])) <= 1e-12


def test_engine_rejects_what_it_cannot_do():
    eng = Engine(NumpyBackend([0.0], [1.0]))
    a = eng.tensor(np.ones((2, 3)))
    with pytest.raises(NotImplementedError):
        eng.einsum("ij,ij->ij", a, a)            # batch indices
    with pytest.raises(NotImplementedError):
        eng.einsum("ij,k->k", a, eng.tensor(np.ones(4)))   # index summed inside one operand
    with pytest.raises(ValueError):
        eng.einsum("ij,jk->ik", a, a)            # extent mismatch


def _engine(p):
    eng = Engine(NumpyBackend(p["eo"], p["ev"]))
    return eng, rmpccsd.DFERIs.from_numpy(eng, p["Loo"], p["Lov"], p["Lvv"], p["fock"])


@needs_ref
@pytest.mark.parametrize("case", [(3, 5, 11, [0, 2], [1, 3], [0, 2], [1, 5, 10, 12]), (4, 7, 13, [1], [0, 2, 6], [], [])])
def test_cc2_branch_against_reference_code(case):
    """cc.cc2 = True: rmpccsd_slow.update_amps :115-133 executed unmodified vs the transcription."""
    o, v, L, act_h, act_p, idx_s, idx_d = case
    p = make_problem(o, v, L, seed=o * 10 + v)
    slow, _ = ref_shim.load_reference_cc()
    cc = types.SimpleNamespace(cc2=True)
    t1r, t2r = slow.update_amps(cc, p["t1"], p["t2"], reference_eris(p), act_h, act_p, idx_s, idx_d)
    eng, eris = _engine(p)
    t1n, t2n = rmpccsd.update_amps(cc, p["t1"], p["t2"], eris, act_h, act_p, idx_s, idx_d)
    assert np.abs(t1n - t1r).max() <= 1e-12 and np.abs(t2n - t2r).max() <= 1e-12


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rmpccsd_cc2_*.npz"))), ids=os.path.basename)
def test_cc2_branch_against_fixture(path):
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    eng, eris = _engine(p)
    t1n, t2n = rmpccsd.update_amps(types.SimpleNamespace(cc2=True), p["t1"], p["t2"], eris, list(g["act_hole"]),
                                   list(g["act_particle"]), list(g["idx_singles"]), list(g["idx_doubles"]))
    assert np.abs(t1n - g["t1new"]).max() <= 1e-12
    assert np.abs(t2n.reshape(-1)[::int(g["stride"])] - g["t2new_samples"]).max() <= 1e-12
    assert abs(np.linalg.norm(t2n) - float(g["t2new_norm"])) <= 1e-11


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(GOLDEN, "rmpccsd_kernel_*.npz"))), ids=os.path.basename)
def test_driver_loop_against_fixture(path):
    """RMPCCSD.kernel (rmpccsd_slow.py:221-258 -> ccsd.kernel, ccsd.py:42-150) with DIIS on packed vectors: cycle
    count, per-cycle energies and norms, converged amplitudes against the loop run around the reference's functions."""
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    if o * v > 150:
        pytest.skip("NumPy stand-in: small cases only")
    p = make_problem(o, v, L, seed=int(g["seed"]))
    eng, eris = _engine(p)
    cc = rmpccsd.RMPCCSD(eris=eris)
    e, t1, t2 = cc.kernel(list(g["act_hole"]), list(g["act_particle"]), list(g["idx_singles"]), list(g["idx_doubles"]),
                          cc2=bool(g["cc2"]))
    hist = np.array(cc.history)
    assert cc.converged == bool(g["converged"]) and cc.cycles == len(g["history"])
    np.testing.assert_allclose(hist[:, 1], g["history"][:, 1], rtol=0, atol=1e-10)
    np.testing.assert_allclose(hist[:, 3], g["history"][:, 3], rtol=1e-6, atol=1e-11)
    assert abs(e - float(g["e_corr"])) <= 1e-10 and abs(cc.emp2 - float(g["emp2"])) <= 1e-12
    assert np.abs(eng.numpy(t1) - g["t1"]).max() <= 1e-10
    t2h = eng.numpy(t2)
    assert np.abs(t2h.reshape(-1)[::int(g["stride"])] - g["t2_samples"]).max() <= 1e-10
    # MBPT2 one-shot and the mutually-exclusive guard (:229-235)
    e2, _, _ = cc.kernel([], [], [], [], mbpt2=True)
    assert abs(e2 - float(g["emp2"])) <= 1e-12
    with pytest.raises(RuntimeError):
        cc.kernel([], [], [], [], mbpt2=True, cc2=True)


class _CountingBackend(NumpyBackend):
    """NumPy stand-in that counts the elements moved by permute passes (what the layout planner minimises)."""

    def __init__(self, *a):
        super().__init__(*a)
        self.permuted = 0

    def permute_axpby(self, src, dst, dims, perm, alpha, beta):
        n = 1
        for d in dims:
            n *= d
        self.permuted += n
        return super().permute_axpby(src, dst, dims, perm, alpha, beta)


def test_layout_planner_leaves_the_large_operand_in_place_and_caches_integral_copies():
    """'kcad,id->akic' (rmpccsd_slow.py Wvoov build): the o v^3 integral block already is [(k,c,a)][d]; only the
    o^2 v^2 result may be permuted. And a permuted copy of a const tensor is made once. Values unchanged."""
    rng = np.random.default_rng(3)
    o, v = 3, 7
    be = _CountingBackend(np.zeros(o), np.ones(v))
    eng = Engine(be)
    ovvv = eng.tensor(rng.standard_normal((o, v, v, v)))
    t1 = eng.tensor(rng.standard_normal((o, v)))
    want = np.einsum("kcad,id->akic", eng.numpy(ovvv), eng.numpy(t1))
    be.permuted = 0
    got = eng.numpy(eng.einsum("kcad,id->akic", ovvv, t1))
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=1e-12)
    assert be.permuted <= o * o * v * v + o * v          # never the o v^3 operand
    # 'kdac,ikcd->ia' needs the block as [a][(k,d,c)] whatever the roles: a copy, but only once when it is const
    t2 = eng.tensor(rng.standard_normal((o, o, v, v)))
    want = np.einsum("kdac,ikcd->ia", eng.numpy(ovvv), eng.numpy(t2))
    ovvv.const = True
    be.permuted = 0
    np.testing.assert_allclose(eng.numpy(eng.einsum("kdac,ikcd->ia", ovvv, t2)), want, rtol=1e-12, atol=1e-12)
    first = be.permuted
    be.permuted = 0
    np.testing.assert_allclose(eng.numpy(eng.einsum("kdac,ikcd->ia", ovvv, t2)), want, rtol=1e-12, atol=1e-12)
    assert first >= o * v ** 3 and be.permuted <= first - o * v ** 3
    # planner off: same numbers
    eng.plan_layouts = False
    np.testing.assert_allclose(eng.numpy(eng.einsum("kcad,id->akic", ovvv, t1)),
                               np.einsum("kcad,id->akic", eng.numpy(ovvv), eng.numpy(t1)), rtol=1e-12, atol=1e-12)


class _RankBackend(NumpyBackend):
    """One simulated rank of a group: ``all_reduce`` only records the buffer (the test sums the ranks' buffers)."""

    def __init__(self, eo, ev, rank, world):
        super().__init__(eo, ev)
        self.rank, self.world, self.reduced = rank, world, []

    def rank_and_size(self, group):
        return self.rank, self.world

    def all_reduce(self, buf, group):
        self.reduced.append(buf.copy())


@pytest.mark.parametrize("world,asym", [(2, 0.0), (3, 0.0), (8, 0.0), (3, 0.05)])
def test_sharded_ladder_and_ring_terms_sum_to_the_one_rank_result(world, asym):
    """Several ranks (rmpccsd.update_amps with cc.group): every rank fills its slabs of the particle-particle ladder
    and its slab of the ring / exchange terms into one [a,b,i,j] accumulator that is all-reduced. Simulated rank by
    rank on the CPU: the accumulators add up to an R whose symmetrised transpose is exactly what the one-rank code
    adds to t2new for those terms; checked through the final amplitudes by linearity (the terms outside R are
    replicated, the division by the denominators is linear)."""
    p = make_problem(5, 11, 23, seed=9)
    o, v = 5, 11
    t2_in = p["t2"] + asym * np.random.default_rng(2).standard_normal(p["t2"].shape)   # asym > 0: the full-ladder path

    def run(rank, world):
        be = _RankBackend(p["eo"], p["ev"], rank, world)
        eng = Engine(be)
        eris = rmpccsd.DFERIs.from_numpy(eng, p["Loo"], p["Lov"], p["Lvv"], p["fock"])
        cc = types.SimpleNamespace(cc2=False, group="g" if world > 1 else None)
        t1n, t2n = rmpccsd.update_amps(cc, p["t1"], t2_in, eris, [], [], [], [])
        return t1n, t2n, be.reduced

    t1_ref, t2_ref, _ = run(0, 1)
    parts = [run(r, world) for r in range(world)]
    # every rank's t2new = D^-1 (replicated terms + sym(R_r)); sum over ranks = D^-1 (world * replicated + sym(R));
    # and with all ranks' R summed in place of its own, each rank would return t2_ref
    t2_sum = sum(t2n for _, t2n, _ in parts)
    t2_rep = (t2_sum - t2_ref) / (world - 1)            # D^-1 (replicated terms) if and only if sum_r R_r == R
    assert all(len(red) == (2 if asym else 1) for _, _, red in parts)
    for t1n, _, _ in parts:
        np.testing.assert_allclose(t1n, t1_ref, rtol=1e-11, atol=1e-12)          # singles are replicated
    R_total = sum(red[-1] for _, _, red in parts).reshape(v, v, o, o)           # asym: [ladder R, ring R2]
    R_once = sum(red[0] for _, _, red in parts).reshape(v, v, o, o) if asym else 0.0
    # reference R: ladder + ring terms of the reference code = what the one-rank result contains beyond the
    # replicated terms; compare sym(R_total) / D with t2_ref - t2_rep
    eia = p["eo"][:, None] - p["ev"][None, :]
    D = eia[:, None, :, None] + eia[None, :, None, :]
    symR = R_total.transpose(2, 3, 0, 1) + R_total.transpose(3, 2, 1, 0)
    if asym:
        symR = symR + R_once.transpose(2, 3, 0, 1)
    # a wrong or missing share Delta = R_total - R would show up here as Delta * world / (world - 1)
    np.testing.assert_allclose(t2_ref - t2_rep, symR / D, rtol=1e-9, atol=1e-11)


@pytest.mark.parametrize("cc2", [False, True])
def test_update_amps_with_a_t2_that_is_not_pair_symmetric(cc2):
    """update_amps is a public function of the reference (rmpccsd_slow.py:73) and assumes nothing about t2; a class
    overwrite with a class list that is not closed under (i,a) <-> (j,b) produces a t2 with t2[ijab] != t2[jiba].
    The halved particle-particle ladder needs that symmetry: it is checked, and the full ladder is used otherwise."""
    slow, _ = ref_shim.load_reference_cc()
    p = make_problem(4, 9, 21, seed=11)
    t2 = p["t2"] + 0.05 * np.random.default_rng(5).standard_normal(p["t2"].shape)
    assert np.abs(t2 - t2.transpose(1, 0, 3, 2)).max() > 0.1
    cc = types.SimpleNamespace(cc2=cc2)
    want1, want2 = slow.update_amps(cc, p["t1"], t2, reference_eris(p), [0, 2], [1, 3], [0], [1, 5])
    eng, eris = _engine(p)
    got1, got2 = rmpccsd.update_amps(cc, p["t1"], t2, eris, [0, 2], [1, 3], [0], [1, 5])
    np.testing.assert_allclose(got1, want1, rtol=1e-11, atol=1e-12)
    np.testing.assert_allclose(got2, want2, rtol=1e-11, atol=1e-12)
