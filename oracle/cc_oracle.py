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
