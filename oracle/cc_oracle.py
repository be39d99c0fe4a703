"""TEST INFRASTRUCTURE — CPU restatement of the CCSD-type driver loop the high-level MPCC stage runs.

``pyscf/cc/rmpccsd_slow.py::RMPCCSD.ccsd`` (:223-258) hands the iteration to ``pyscf/cc/ccsd.py::kernel`` (:42-150),
which cannot be imported here (it pulls libcc / ao2mo / h5py). The loop is restated below, line by line, around
callables that ARE the reference's own code when the full tree is present: ``update_amps`` / ``energy`` of
``rmpccsd_slow.py`` and ``lib.diis.DIIS`` of ``pyscf/lib/diis.py``, all executed unmodified through
``oracle/ref_shim.py`` (``tests/golden/generate_golden_cc2.py``). Only ``tests/`` may import this file.

Restated: ``kernel`` ccsd.py:42-150; ``amplitudes_to_vector`` / ``vector_to_amplitudes`` ccsd.py:719-735;
``run_diis`` ccsd.py:1269-1276; ``init_amps`` rmpccsd_slow.py:206-219.
"""
from __future__ import annotations

import numpy as np


def amplitudes_to_vector(t1, t2):
    """ccsd.py:719-726: [t1 | pack_tril(t2.transpose(0,2,1,3).reshape(nov, nov))]."""
    o, v = t1.shape
    nov = o * v
    m = t2.transpose(0, 2, 1, 3).reshape(nov, nov)
    r, c = np.tril_indices(nov)
    return np.concatenate([t1.ravel(), m[r, c]])


def vector_to_amplitudes(vec, nmo, nocc):
    """ccsd.py:728-735: symmetric unpack, (o,v,o,v) -> transpose(0,2,1,3), C-contiguous."""
    nvir = nmo - nocc
    nov = nocc * nvir
    t1 = vec[:nov].copy().reshape(nocc, nvir)
    r, c = np.tril_indices(nov)
    full = np.zeros((nov, nov))
    full[c, r] = vec[nov:]
    full[r, c] = vec[nov:]
    t2 = full.reshape(nocc, nvir, nocc, nvir).transpose(0, 2, 1, 3)
    return t1, np.asarray(t2, order="C")


def init_amps(fock, ovov, nocc):
    """rmpccsd_slow.py:206-219."""
    mo_e = fock.diagonal().real
    eia = mo_e[:nocc, None] - mo_e[None, nocc:]
    eijab = eia[:, None, :, None] + eia[None, :, None, :]
    t1 = fock[:nocc, nocc:] / eia
    t2 = ovov.transpose(0, 2, 1, 3) / eijab
    emp2 = 2 * np.einsum("ijab,iajb", t2, ovov) - np.einsum("ijab,ibja", t2, ovov)
    return emp2, t1, t2


def ccsd_kernel(update_amps, energy, diis_factory, t1, t2, nmo, nocc, max_cycle=50, tol=1e-7, tolnormt=1e-5,
                diis_start_cycle=0, diis_start_energy_diff=1e9, iterative_damping=1.0):
    """ccsd.py:42-150 with ``update_amps(t1, t2) -> (t1new, t2new)`` and ``energy(t1, t2) -> float`` closed over the
    eris and the active-space arguments. ``diis_factory()`` returns a ``lib.diis.DIIS``-like object or None.
    Returns (converged, eccsd, t1, t2, history) with history rows (cycle, eccsd, dE, normt)."""
    eold = 0
    eccsd = energy(t1, t2)
    adiis = diis_factory()
    converged = False
    history = []
    for istep in range(max_cycle):
        t1new, t2new = update_amps(t1, t2)
        tmpvec = amplitudes_to_vector(t1new, t2new)
        tmpvec -= amplitudes_to_vector(t1, t2)
        normt = np.linalg.norm(tmpvec)
        if iterative_damping < 1.0:
            alpha = iterative_damping
            t1new = (1 - alpha) * t1 + alpha * t1new
            t2new *= alpha
            t2new += (1 - alpha) * t2
        t1, t2 = t1new, t2new
        if adiis is not None and istep >= diis_start_cycle and abs(eccsd - eold) < diis_start_energy_diff:
            vec = amplitudes_to_vector(t1, t2)
            t1, t2 = vector_to_amplitudes(adiis.update(vec), nmo, nocc)
        eold, eccsd = eccsd, energy(t1, t2)
        history.append((istep + 1, eccsd, eccsd - eold, normt))
        if abs(eccsd - eold) < tol and normt < tolnormt:
            converged = True
            break
    return converged, eccsd, t1, t2, history


# ---------------------------------------------------------------------------------------------------------------
# Unrestricted low-level prototype (row f4): the vector helpers its run_diis needs and the loop of its kernel().
# pyscf/cc/umpcc_lowlevel.py:15-46 calls ccsd.amplitudes_to_vector_s4 / vector_to_amplitudes_s4 (pyscf/cc/ccsd.py
# :705-757), which go through lib.take_2d / lib.takebak_2d (pyscf/lib/numpy_helper.py:478-541, C helpers of
# libnp_helper that cannot be loaded here). Restated in NumPy below; ref_shim installs them in the fake pyscf.cc.ccsd
# so that the reference's own amplitudes_to_vector / vector_to_amplitudes / run_diis run unmodified on top of them.
# tests/test_oracle_vs_reference.py executes the source text of the three ccsd.py functions themselves (extracted from
# the reference file at test time, with these take_2d / takebak_2d) against the restatement.
# ---------------------------------------------------------------------------------------------------------------
def take_2d(a, idx, idy, out=None):
    """a[idx[:,None], idy] for a 2-D array (numpy_helper.py:478-504)."""
    a = np.asarray(a, order="C")
    idx, idy = np.asarray(idx, dtype=np.int64), np.asarray(idy, dtype=np.int64)
    res = np.ndarray((len(idx), len(idy)), dtype=a.dtype, buffer=out)
    res[:] = a[idx[:, None], idy]
    return res


def takebak_2d(out, a, idx, idy, thread_safe=True):
    """out[idx[:,None], idy] += a (numpy_helper.py:506-541); the index sets used here hold no duplicates."""
    idx, idy = np.asarray(idx, dtype=np.int64), np.asarray(idy, dtype=np.int64)
    out[idx[:, None], idy] += np.asarray(a).reshape(len(idx), len(idy))
    return out


def amplitudes_to_vector_s4(t1, t2, out=None):
    """[t1 | t2 viewed as (nocc^2, nvir^2), rows i>j, columns a>b] (ccsd.py:738-748). ``t2`` only needs nocc^2 nvir^2
    elements: the unrestricted prototype hands in its (v,o,v,o) blocks and the reshape reinterprets them."""
    nocc, nvir = t1.shape
    nov = nocc * nvir
    size = nov + nocc * (nocc - 1) // 2 * nvir * (nvir - 1) // 2
    vector = np.ndarray(size, t1.dtype, buffer=out)
    vector[:nov] = t1.ravel()
    otril = np.tril_indices(nocc, k=-1)
    vtril = np.tril_indices(nvir, k=-1)
    take_2d(t2.reshape(nocc ** 2, nvir ** 2), otril[0] * nocc + otril[1], vtril[0] * nvir + vtril[1], out=vector[nov:])
    return vector


def _unpack_4fold(c2vec, nocc, nvir, anti_symm=True):
    """(ccsd.py:705-717): the packed block goes to (i>j, a>b) and (j,i | b,a) as it is, to (i,j | b,a) and (j,i | a,b)
    with the opposite sign."""
    t2 = np.zeros((nocc ** 2, nvir ** 2), dtype=c2vec.dtype)
    if nocc > 1 and nvir > 1:
        t2tril = c2vec.reshape(nocc * (nocc - 1) // 2, nvir * (nvir - 1) // 2)
        otril = np.tril_indices(nocc, k=-1)
        vtril = np.tril_indices(nvir, k=-1)
        takebak_2d(t2, t2tril, otril[0] * nocc + otril[1], vtril[0] * nvir + vtril[1])
        takebak_2d(t2, t2tril, otril[1] * nocc + otril[0], vtril[1] * nvir + vtril[0])
        if anti_symm:
            t2tril = -t2tril
        takebak_2d(t2, t2tril, otril[0] * nocc + otril[1], vtril[1] * nvir + vtril[0])
        takebak_2d(t2, t2tril, otril[1] * nocc + otril[0], vtril[0] * nvir + vtril[1])
    return t2.reshape(nocc, nocc, nvir, nvir)


def vector_to_amplitudes_s4(vector, nmo, nocc):
    """(ccsd.py:750-757)"""
    nvir = nmo - nocc
    nov = nocc * nvir
    size = nov + nocc * (nocc - 1) // 2 * nvir * (nvir - 1) // 2
    t1 = vector[:nov].copy().reshape(nocc, nvir)
    t2 = _unpack_4fold(vector[nov:size], nocc, nvir)
    return t1, t2


def umpcc_kernel(myll, updated_amp, run_diis, diis_factory, tol=1e-6, maxcount=100):
    """The loop of ``kernel`` in pyscf/cc/umpcc_lowlevel.py:48-89, around the callables of the reference module:
    ``compute_three_center_ints`` and ``initialize_t1`` (:60-63) cannot run here (PySCF mean-field objects, an undefined
    global ``mol``) — the caller provides ``myll`` with the three-index tensors in place and ``myll.t1`` zero, which is
    what ``initialize_t1`` leaves (:171); tol / maxcount are the literals of :53-54. Returns (DE, count, history)."""
    err, count, DE = np.inf, 0, None
    history = []
    adiis = diis_factory()
    while err > tol and count < maxcount:
        res, DE, t1_new, t2_new = updated_amp(myll)
        if myll.diis:
            t1, t2 = run_diis(myll, t1_new, t2_new, adiis)
        else:
            t1, t2 = t1_new, t2_new
        myll.t1 = t1
        myll.t2 = t2
        count += 1
        err = res
        history.append((count, DE, res))
    return DE, count, history
