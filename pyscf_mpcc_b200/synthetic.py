"""Synthetic density-fitted mean-field objects for the DF-MPCC hot path.

``SyntheticMF`` is the duck-typed stand-in for a PySCF ``RHF(...).density_fit()``
object: it offers exactly the attributes the reference low-level solver reads
from ``mf`` (``/root/reference/pyscf/mpcc/dfrmpcc_lowlevel.py:10-35``):
``mol.nao``, ``mol.nelec``, ``mo_energy``, ``mo_coeff``, ``get_fock()`` and
``with_df.get_naoaux()/loop()`` (packed lower-triangular Cholesky rows in
blocks of ``blockdim`` = 240, ``pyscf/df/df.py:76,194-237``).

The random draw order is part of the fixture contract: the
golden vectors under ``tests/golden/`` were produced from this generator, so
any change to the order or the distributions invalidates them.
"""
from __future__ import annotations

import types

import numpy as np


def pack_lower(mat: np.ndarray) -> np.ndarray:
    """Row-major lower triangle (incl. diagonal) of a square matrix as 1-D.

    Same element order as ``NPdpack_tril`` (``pyscf/lib/np_helper/pack_tril.c:99-107``).
    """
    n = mat.shape[0]
    rows, cols = np.tril_indices(n)
    return np.ascontiguousarray(mat[rows, cols])


class SyntheticDF:
    """Minimal ``with_df`` object: holds the packed (naux, nao_pair) tensor."""

    blockdim = 240

    def __init__(self, cderi: np.ndarray):
        self._cderi = cderi

    def get_naoaux(self) -> int:
        return int(self._cderi.shape[0])

    def loop(self, blksize=None):
        step = int(blksize or self.blockdim)
        for lo in range(0, self._cderi.shape[0], step):
            yield np.asarray(self._cderi[lo:lo + step], order="C")


class SyntheticMF:
    """Seeded synthetic RHF+DF problem of shape (nocc, nvir, naux).

    ``L[k]`` is a symmetric (nao, nao) matrix with N(0,1) entries scaled by
    ``scale/sqrt(naux)``; occupied energies are sorted U(-2,-0.5), virtual
    energies sorted U(0.3,3); MO coefficients are the identity, so the Fock
    matrix is ``diag(mo_energy)``.
    """

    def __init__(self, nocc: int, nvir: int, naux: int, seed: int = 0, scale: float = 0.1):
        rng = np.random.default_rng(seed)
        nao = nocc + nvir
        self.mol = types.SimpleNamespace(nao=nao, nelec=(nocc, nocc))
        full = rng.standard_normal((naux, nao, nao)) * scale / np.sqrt(naux)
        full = full + full.transpose(0, 2, 1)
        rows, cols = np.tril_indices(nao)
        self.with_df = SyntheticDF(np.ascontiguousarray(full[:, rows, cols]))
        e_occ = np.sort(rng.uniform(-2.0, -0.5, nocc))
        e_vir = np.sort(rng.uniform(0.3, 3.0, nvir))
        self.mo_energy = np.concatenate([e_occ, e_vir])
        self.mo_coeff = np.eye(nao)
        # kept so callers that only want the MO blocks can skip the AO->MO step
        self._full = full
        self.nocc, self.nvir, self.naux = nocc, nvir, naux

    def get_fock(self) -> np.ndarray:
        return np.diag(self.mo_energy)

    def mo_blocks(self):
        """(Loo, Lov, Lvv) C-contiguous copies; valid because mo_coeff = I."""
        o = self.nocc
        f = self._full
        return (np.ascontiguousarray(f[:, :o, :o]),
                np.ascontiguousarray(f[:, :o, o:]),
                np.ascontiguousarray(f[:, o:, o:]))


def host_blocks(nocc: int, nvir: int, naux: int, seed: int = 0, scale: float | None = None, chunk: int = 64):
    """(Loo, Lov, Lvv, eps_o, eps_v) of ``SyntheticMF(nocc, nvir, naux, seed, scale)`` — the same random stream, hence
    bit-identical MO blocks (``mo_coeff`` = identity) — drawn ``chunk`` auxiliary rows at a time, so that the full
    (naux, nao, nao) tensor and its packed copy never exist: at (64, 512, 1536) the peak is the 3.6 GB of the blocks
    instead of 12 GB. ``numpy.random.Generator.standard_normal`` fills its output sequentially, so chunked draws
    continue the stream exactly (``tests/test_host_logic.py::test_host_blocks_equal_synthetic_mf``). These are the
    inputs of ``bench.py`` and of the full-size trace fixture: reproducible on the CPU (oracle) and on every rank."""
    if scale is None:
        scale = default_scale(nocc, nvir)
    rng = np.random.default_rng(seed)
    nao = nocc + nvir
    Loo = np.empty((naux, nocc, nocc))
    Lov = np.empty((naux, nocc, nvir))
    Lvv = np.empty((naux, nvir, nvir))
    for lo in range(0, naux, chunk):
        hi = min(lo + chunk, naux)
        full = rng.standard_normal((hi - lo, nao, nao)) * scale / np.sqrt(naux)      # same rounding as SyntheticMF
        full = full + full.transpose(0, 2, 1)
        Loo[lo:hi] = full[:, :nocc, :nocc]
        Lov[lo:hi] = full[:, :nocc, nocc:]
        Lvv[lo:hi] = full[:, nocc:, nocc:]
    e_occ = np.sort(rng.uniform(-2.0, -0.5, nocc))
    e_vir = np.sort(rng.uniform(0.3, 3.0, nvir))
    return Loo, Lov, Lvv, e_occ, e_vir


def default_scale(nocc: int, nvir: int) -> float:
    """Scale that keeps |dE| O(1) for large shapes."""
    return 0.1 * float(np.sqrt(76.0 / (nocc + nvir)))


def device_blocks(nocc, nvir, naux, seed=0, scale=None, device="cuda", symmetric=True):
    """Generate (Loo, Lov, Lvv, eps_o, eps_v) directly on ``device`` with torch's
    Philox generator — throughput-only inputs for shapes whose host generator
    would need tens of GB ("use those only for throughput, not
    for parity")."""
    import torch
    if scale is None:
        scale = default_scale(nocc, nvir)
    g = torch.Generator(device=device)
    g.manual_seed(int(seed))
    amp = scale / float(np.sqrt(naux))
    kw = dict(device=device, dtype=torch.float64, generator=g)
    Loo = torch.randn(naux, nocc, nocc, **kw) * amp
    Loo = Loo + Loo.transpose(1, 2)
    Lov = torch.randn(naux, nocc, nvir, **kw) * (amp * np.sqrt(2.0))
    Lvv = torch.randn(naux, nvir, nvir, **kw) * amp
    if symmetric:
        # in place: Lvv + Lvv^T without a second full-size temporary per L
        for lo in range(0, naux, 64):
            blk = Lvv[lo:lo + 64]
            blk.add_(blk.transpose(1, 2).clone())
    eo = torch.sort(torch.rand(nocc, **kw) * 1.5 - 2.0).values
    ev = torch.sort(torch.rand(nvir, **kw) * 2.7 + 0.3).values
    return Loo.contiguous(), Lov.contiguous(), Lvv, eo, ev
