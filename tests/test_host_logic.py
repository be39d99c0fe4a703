"""Host-side logic of the package that needs no GPU: factory dispatch, sharding arithmetic,
the DIIS bookkeeping (driven through a NumPy stand-in for the device kernels), flop counts."""
import ctypes
import os
import types

import numpy as np
import pytest

from conftest import GOLDEN
import pyscf_mpcc_b200
from pyscf_mpcc_b200 import sharding, flops
from pyscf_mpcc_b200.diis import DIIS, AmplitudeVector, solve_diis_system
from pyscf_mpcc_b200.synthetic import SyntheticMF, pack_lower


def test_factory_rejects_unrestricted_and_non_df():
    class UHF: pass
    class GHF: pass
    class RHF: with_df = None
    with pytest.raises(RuntimeError, match="UMPCC is not implemented"):
        pyscf_mpcc_b200.MPCC(UHF())
    with pytest.raises(RuntimeError, match="GMPCC is not implemented"):
        pyscf_mpcc_b200.MPCC(GHF())
    with pytest.raises(RuntimeError, match="Chosen RMPCC variant is not implemented"):
        pyscf_mpcc_b200.MPCC(RHF())


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU path"):
        pyscf_mpcc_b200.MPCC(SyntheticMF(2, 3, 5))


def test_synthetic_generator_contract():
    mf = SyntheticMF(4, 10, 30, seed=0, scale=0.3)
    assert mf.mol.nao == 14 and mf.mol.nelec == (4, 4) and mf.with_df.get_naoaux() == 30
    blocks = list(mf.with_df.loop())
    assert blocks[0].shape == (30, 14 * 15 // 2) and blocks[0].flags.c_contiguous
    np.testing.assert_array_equal(blocks[0][3], pack_lower(mf._full[3]))
    assert np.all(np.diff(mf.mo_energy[:4]) > 0) and mf.mo_energy[3] < -0.5 < 0.3 < mf.mo_energy[4]
    np.testing.assert_array_equal(mf.get_fock(), np.diag(mf.mo_energy))
    # first draws are part of the fixture contract (tests/golden was generated from them)
    assert abs(mf._full[0, 0, 0] - 2 * 0.3 / np.sqrt(30) * np.random.default_rng(0).standard_normal()) < 1e-15
    assert [b.shape[0] for b in SyntheticMF(2, 3, 500).with_df.loop()] == [240, 240, 20]


@pytest.mark.parametrize("v,p", [(512, 8), (243, 4), (19, 2), (7, 7), (10, 3)])
def test_slab_bounds_partition(v, p):
    spans = [sharding.slab_bounds(v, p, r) for r in range(p)]
    assert spans[0][0] == 0 and spans[-1][1] == v
    assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
    sizes = [e - b for b, e in spans]
    assert max(sizes) - min(sizes) <= 1 and min(sizes) >= 1
    o = 5
    segs = [sharding.packed_segment(o, b, e) for b, e in spans]
    nov = o * v
    assert segs[0][0] == 0 and segs[-1][1] == nov * (nov + 1) // 2
    assert sum(sharding.local_vector_length(o, v, b, e) - nov for b, e in spans) == nov * (nov + 1) // 2
    with pytest.raises(ValueError):
        sharding.slab_bounds(3, 4, 0)


def test_solve_diis_system_matches_reference_branches():
    h = np.array([[0.0, 1.0], [1.0, 2.5]])
    np.testing.assert_allclose(solve_diis_system(h), [-2.5, 1.0])
    # singular B (two identical error vectors) takes the pseudo-inverse branch (diis.py:257-259)
    h = np.array([[0, 1, 1], [1, 1e-3, 1e-3], [1, 1e-3, 1e-3]], dtype=float)
    c = solve_diis_system(h)
    assert abs(c[1] + c[2] - 1.0) < 1e-9 and abs(c[1] - c[2]) < 1e-9


class _FakeCtx:
    """NumPy stand-in for mpccb_diis_push / mpccb_diis_combine so the ring/B-matrix logic of
    pyscf_mpcc_b200.diis.DIIS can be pinned to the reference's DIIS fixture without a GPU."""

    def __init__(self):
        self.mem = {}
        self.next = 1000

    def alloc(self, arr):
        self.next += 16
        self.mem[self.next] = arr
        return self.next

    def call(self, name, *a):
        if name == "mpccb_diis_push":
            t1p, xprev, xpl, xpc, n_prev, xs, es, elist, nd, slot, first, count_t1, dots = a
            x = self.mem[t1p]
            if not first:
                if n_prev:                       # xprev as a combination of the stored vectors (read before xs is written)
                    prev = np.zeros(len(x))
                    for k in range(n_prev):
                        prev += self.mem[xpl[k]] * xpc[k]
                else:
                    prev = self.mem[xprev].copy()
            self.mem[xs][:] = x
            if not first:
                self.mem[es][:] = x - prev
                d = self.mem[dots]
                for k in range(nd):
                    ek = self.mem[es] if k == slot else self.mem[elist[k]]
                    d[k] = float(np.dot(self.mem[es], ek))
        elif name == "mpccb_diis_combine":
            xl, coeff, nd, out, n = a
            acc = np.zeros(n)
            for k in range(nd):
                acc += self.mem[xl[k]] * coeff[k]
            self.mem[out][:] = acc
        else:
            raise AssertionError(name)


class _FakeTensor:
    def __init__(self, ctx, arr):
        self.ctx, self.arr = ctx, arr
        self.ptr = ctx.alloc(arr)

    def data_ptr(self):
        return self.ptr

    def __getitem__(self, s):
        return self

    def numel(self):
        return self.arr.size

    def numpy(self):
        return self.arr

    def copy_(self, other, non_blocking=False):
        self.arr[:len(other.arr)] = other.arr
        return self


@pytest.mark.parametrize("implicit", [True, False])
def test_device_diis_bookkeeping_against_reference_fixture(monkeypatch, implicit):
    g = np.load(os.path.join(GOLDEN, "diis_kat.npz"))
    n = g["inputs"].shape[1]
    ctx = _FakeCtx()
    ll = types.SimpleNamespace(_ctx=ctx, _vec_len_local=n, device="cpu", nov=n, _count_t1=True, world_size=1,
                               _sync=lambda: None, _diis_implicit=implicit)
    ll._dots_dev = _FakeTensor(ctx, np.zeros(7))
    ll._dots_host = _FakeTensor(ctx, np.zeros(7))
    d = DIIS()
    monkeypatch.setattr(DIIS, "_alloc", lambda self, k, dev: _FakeTensor(ctx, np.zeros(k)))
    import pyscf_mpcc_b200.diis as diis_mod
    monkeypatch.setattr(diis_mod.torch, "empty", lambda k, dtype=None, device=None: _FakeTensor(ctx, np.zeros(k)))
    for x, want in zip(g["inputs"], g["outputs"]):
        out = d.update(AmplitudeVector(ll, vec=None, t1_dev=_FakeTensor(ctx, x.copy())))
        np.testing.assert_allclose(out.packed_device().arr, want, rtol=1e-9, atol=1e-11)
        np.testing.assert_allclose(out.t1_dev.arr, want, rtol=1e-9, atol=1e-11)      # nov == n in this stand-in
    np.testing.assert_allclose(d._H, g["H"], rtol=1e-10, atol=1e-14)
    assert d._head == 3 and d.get_num_vec() == 6          # 9 pushes into a ring of 6 wrapped once


def test_flop_and_byte_counts():
    assert abs(flops.updated_amps_flops(64, 512, 1536) / 6.799e12 - 1) < 1e-3
    assert flops.eris_flops(64, 512, 1536, 512) == 2.0 * 1536 * 32768 ** 2
    n = 32768 + 32768 * 32769 // 2
    assert flops.diis_bytes(64, 512, 6) == 8.0 * n * 16


@pytest.mark.parametrize("nranks,v", [(1, 7), (2, 8), (2, 9), (3, 10), (4, 13), (8, 16), (8, 37)])
def test_mirror_scheme_covers_every_block_exactly_once(nranks, v):
    """Rule of the symmetric eris/t2 scheme (include/mpcc_b200.h, mpccb_bind_t2_peers): of every pair of blocks
    {(a,b),(b,a)} exactly one is computed, by the rank that owns its first index, and that rank mirrors it; the
    work is balanced to within one slab-half. Pure host function of the library (no GPU)."""
    import ctypes
    from pyscf_mpcc_b200 import _cabi, sharding
    lib = _cabi.load()
    bounds = [sharding.slab_bounds(v, nranks, r)[0] for r in range(nranks)] + [v]
    arr = (ctypes.c_int * (nranks + 1))(*bounds)
    owner = lambda x: max(r for r in range(nranks) if bounds[r] <= x)
    work = [0] * nranks
    for a in range(v):
        for b in range(v):
            qa, qb = owner(a), owner(b)
            fa = lib.mpccb_sym_query(nranks, qa, arr, 1, a, b)
            fb = lib.mpccb_sym_query(nranks, qb, arr, 1, b, a)
            assert (fa >> 8) == qb and (fb >> 8) == qa
            if a == b:
                assert fa & 1 and not fa & 2
            else:
                assert (fa & 1) + (fb & 1) == 1, (a, b)          # exactly one of the pair is computed ...
                assert bool(fa & 1) == bool(fa & 2) and bool(fb & 1) == bool(fb & 2)   # ... and mirrored by its owner
            work[qa] += fa & 1
            # without peer mapping every cross-slab block is computed locally and nothing is mirrored across slabs
            f0 = lib.mpccb_sym_query(nranks, qa, arr, 0, a, b)
            assert (f0 & 1) == (1 if qa != qb else int(b >= a)) and bool(f0 & 2) == (qa == qb and b > a)
    if v % nranks == 0 and (v // nranks) % 2 == 0:
        assert max(work) - min(work) <= v // nranks          # balanced: P/2 slab-blocks each (+- the diagonal)


@pytest.mark.parametrize("o,v,L,p", [(64, 512, 1536, 2), (64, 512, 1536, 8), (80, 848, 2224, 4), (32, 256, 768, 8),
                                     (5, 13, 40, 2), (3, 7, 11, 7)])
def test_balanced_bounds(o, v, L, p):
    """Cost-balanced slabs: a partition of [0, nvir) in tile granularity whose modelled spread is no worse than
    the equal split's (and wider slabs come first, since packed rows get longer with a)."""
    b = sharding.balanced_bounds(o, v, L, p)
    assert len(b) == p + 1 and b[0] == 0 and b[-1] == v
    widths = [hi - lo for lo, hi in zip(b[:-1], b[1:])]
    assert all(w >= 1 for w in widths)
    equal = [sharding.slab_bounds(v, p, r)[0] for r in range(p)] + [v]
    if b != equal:
        gran = 128 // o if 128 % o == 0 else 1
        assert all(x % gran == 0 for x in b[:-1])
        assert widths[0] >= widths[-1]
        spread = lambda bb: max(sharding.modelled_cost(o, v, L, bb)) / (sum(sharding.modelled_cost(o, v, L, bb)) / p)
        assert spread(b) <= spread(equal) + 1e-12
    assert sharding.balanced_bounds(o, v, L, p, diis=False) == equal
    assert sharding.balanced_bounds(o, v, L, 1) == [0, v]


@pytest.mark.parametrize("o,v,p", [(3, 7, 2), (4, 10, 3), (5, 13, 4), (8, 12, 2), (6, 5, 2), (2, 9, 8)])
def test_column_owned_diis_vector_is_a_partition_of_the_packed_triangle(o, v, p):
    """Column ownership (mpccb_bind_diis_columns): over the ranks every packed element appears exactly once, the
    values read from the local slabs are the reference vector's, and the shares are close to 1/P."""
    rng = np.random.default_rng(o * 100 + v)
    nov = o * v
    m = rng.standard_normal((nov, nov))
    m = m + m.T                                                   # (ai|bj)-type symmetry: T[a,b,i,j] = T[b,a,j,i]
    T = m.reshape(v, o, v, o).transpose(0, 2, 1, 3).copy()         # [a,b,i,j] buffer
    ref_vec = T.reshape(nov, nov)[np.tril_indices(nov)]           # dfrmpcc_lowlevel.py:162-168 on that buffer
    seen = np.zeros(ref_vec.size, dtype=int)
    sizes = []
    for r in range(p):
        b0, b1 = sharding.slab_bounds(v, p, r)
        prefix = sharding.column_owned_prefix(o, v, b0, b1)
        loc, ref = sharding.column_owned_positions(o, v, b0, b1)
        assert prefix[-1] == loc.size == ref.size
        seen[ref] += 1
        slab = T[b0:b1].reshape(-1)
        assert np.array_equal(slab[loc], ref_vec[ref])
        sizes.append(loc.size)
    assert np.all(seen == 1)
    assert sum(sizes) == nov * (nov + 1) // 2


def test_column_owned_shares_are_balanced_at_the_target_shape():
    o, v = 64, 512
    for p in (2, 8):
        sizes = [int(sharding.column_owned_prefix(o, v, *sharding.slab_bounds(v, p, r))[-1]) for r in range(p)]
        assert max(sizes) / (sum(sizes) / p) < 1.003
        rows = [sharding.local_vector_length(o, v, *sharding.slab_bounds(v, p, r)) - o * v for r in range(p)]
        assert max(rows) / (sum(rows) / p) > 1.4                 # what the row-owned split costs


def test_host_blocks_equal_synthetic_mf():
    """The chunked generator behind bench.py and the full-size trace fixture draws the same stream as SyntheticMF."""
    from pyscf_mpcc_b200.synthetic import SyntheticMF, host_blocks
    o, v, L = 6, 21, 50
    mf = SyntheticMF(o, v, L, seed=5, scale=0.07)
    for chunk in (1, 7, 64):
        Loo, Lov, Lvv, eo, ev = host_blocks(o, v, L, seed=5, scale=0.07, chunk=chunk)
        for got, want in zip((Loo, Lov, Lvv), mf.mo_blocks()):
            assert np.array_equal(got, want)
        assert np.array_equal(np.concatenate([eo, ev]), mf.mo_energy)


def test_staged_reference_files_match_the_reference_tree():
    """oracle/make_ref.py stages the five hot-path files byte for byte (build container only)."""
    import hashlib
    from oracle import make_ref
    if not os.path.isdir(make_ref.SOURCE_ROOT):
        pytest.skip("reference tree not present")
    manifest = make_ref.stage()
    for rel, digest in manifest.items():
        with open(os.path.join(make_ref.SOURCE_ROOT, rel), "rb") as f:
            assert hashlib.sha256(f.read()).hexdigest() == digest
    assert make_ref.staged()


def test_direct_lapack_eigh_is_bit_identical_to_scipy_eigh():
    """The DIIS solve calls dsyevr directly (pyscf_mpcc_b200.diis._eigh); the reference goes through
    scipy.linalg.eigh (pyscf/lib/diis.py:251-265). Same routine, same arguments: identical bits, including the
    nearly singular B matrices of a converged solve where the |w| < 1e-14 filter decides the branch."""
    import scipy.linalg
    from pyscf_mpcc_b200.diis import _eigh
    rng = np.random.default_rng(1)
    for n in range(2, 8):
        for decay in (1, 3, 5, 8, 10):
            for _ in range(10):
                e = rng.standard_normal((n - 1, 50)) * np.logspace(0, -decay, n - 1)[:, None]
                h = np.zeros((n, n))
                h[0, 1:] = h[1:, 0] = 1
                h[1:, 1:] = e @ e.T
                w0, v0 = scipy.linalg.eigh(h)
                w1, v1 = _eigh(h.copy())
                assert np.array_equal(w0, w1) and np.array_equal(v0, v1)


@pytest.mark.parametrize("nranks,v", [(2, 8), (3, 10), (4, 13), (5, 11), (8, 16), (8, 37)])
def test_subset_exchange_sends_every_slab_to_every_rank_that_computes_with_it(nranks, v):
    """mpccb_bind_jk_subset stores a slab's dressed integrals only into the ranks given by mpccb_jk_subset_needs.
    Every block (a,b) a rank computes under the mirror scheme reads the JK rows of the slabs of a (its own) and of b:
    the owner of b must be sending to it. And nobody receives a slab it never reads (that is the saving)."""
    import ctypes
    from pyscf_mpcc_b200 import _cabi, sharding
    lib = _cabi.load()
    bounds = [sharding.slab_bounds(v, nranks, r)[0] for r in range(nranks)] + [v]
    arr = (ctypes.c_int * (nranks + 1))(*bounds)
    owner = lambda x: max(r for r in range(nranks) if bounds[r] <= x)
    used = set()
    for a in range(v):
        for b in range(v):
            q = owner(a)
            if lib.mpccb_sym_query(nranks, q, arr, 1, a, b) & 1:          # rank q computes block (a,b)
                assert lib.mpccb_jk_subset_needs(nranks, owner(b), q) == 1, (a, b)
                used.add((owner(b), q))
    for src in range(nranks):
        assert lib.mpccb_jk_subset_needs(nranks, src, src) == 1
        for dst in range(nranks):
            if lib.mpccb_jk_subset_needs(nranks, src, dst):
                assert (src, dst) in used, (src, dst)
        assert sum(lib.mpccb_jk_subset_needs(nranks, src, d) for d in range(nranks)) == nranks // 2 + 1
