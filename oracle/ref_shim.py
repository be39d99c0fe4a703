"""TEST INFRASTRUCTURE — loads the reference's own hot-path files UNMODIFIED.

Runs ``/root/reference/pyscf/mpcc/dfrmpcc_lowlevel.py`` and
``/root/reference/pyscf/lib/diis.py`` exactly as written by installing a tiny
fake ``pyscf`` / ``pyscf.lib`` package in ``sys.modules`` (real PySCF cannot be
imported in this image: no h5py, no compiled ``libnp_helper``).

``/root/reference`` exists only in the build container, never on the GPU box.
There the five hot-path files staged by ``oracle/make_ref.py`` under ``oracle/_ref/``
(git-ignored, shipped with the snapshot) are executed instead — the same bytes.
Callers: ``tests/golden/generate_golden*.py`` (run here, output committed), the tests
that re-validate the oracle or drive the reference's solver object through the GPU
module, and the ``--impl reference`` / ``cpu_baseline`` legs of ``bench.py``.
Nothing under ``pyscf_mpcc_b200/`` imports this.

The fake helpers restate:
  * ``lib.pack_tril``   — ``pyscf/lib/numpy_helper.py:256-294`` + ``np_helper/pack_tril.c:99-107``
  * ``lib.unpack_tril`` — ``numpy_helper.py:297-358`` + ``pack_tril.c:24-37,58-69`` (symmetric fill)
  * ``lib.direct_sum``  — ``numpy_helper.py:1060-1127`` (only the two strings MPCC_LL uses)
  * ``lib.StreamObject``— ``pyscf/lib/misc.py:578-612``
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np

_STAGED_ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _pick_root() -> str:
    """MPCC_REFERENCE_ROOT, else the reference tree of the build container, else the staged copy (oracle/make_ref.py)."""
    env = os.environ.get("MPCC_REFERENCE_ROOT")
    if env:
        return env
    for root in ("/root/reference", _STAGED_ROOT):
        if os.path.isfile(os.path.join(root, "pyscf", "mpcc", "dfrmpcc_lowlevel.py")):
            return root
    return "/root/reference"


REFERENCE_ROOT = _pick_root()


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "pyscf", "mpcc", "dfrmpcc_lowlevel.py"))


def full_tree_available() -> bool:
    """True when the whole reference tree is there (build container), not only the staged hot-path files."""
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "pyscf", "cc", "rmpccsd_slow.py"))


def _tril_pack(mat, axis=-1, out=None):
    mat = np.asarray(mat)
    n = mat.shape[0]
    dst = np.ndarray(n * (n + 1) // 2, mat.dtype, buffer=out)
    r, c = np.tril_indices(n)
    dst[:] = mat[r, c]
    return dst


def _tril_unpack(tril, filltriu=1, axis=-1, out=None):
    tril = np.asarray(tril)
    if tril.ndim == 1:
        n = int(np.sqrt(2 * tril.size))
        full = np.zeros((n, n), tril.dtype)
        r, c = np.tril_indices(n)
        full[c, r] = tril
        full[r, c] = tril
        return full
    cnt, npair = tril.shape
    n = int(np.sqrt(2 * npair))
    full = np.zeros((cnt, n, n), tril.dtype)
    r, c = np.tril_indices(n)
    full[:, c, r] = tril
    full[:, r, c] = tril
    return full


def _direct_sum(spec, *ops):
    if spec == "a-i->ia":
        ev, eo = ops
        return np.ascontiguousarray((np.asarray(ev)[:, None] - np.asarray(eo)[None, :]).T)
    if spec == "-ia-jb->aibj":
        x, y = ops
        four = -x[:, :, None, None] - y[None, None, :, :]        # [i,a,j,b]
        return np.einsum("iajb->aibj", four)
    raise NotImplementedError(spec)


class _MemFile(dict):
    """Stand-in for ``misc.H5TmpFile`` (diis.py:123-139 spills vectors >= 1e7 doubles)."""

    def __init__(self, *a, **k):
        super().__init__()

    def flush(self):
        pass

    def create_dataset(self, key, shape, dtype):
        self[key] = np.empty(shape, dtype)
        return self[key]


_loaded = None
_loaded_cc = None
_loaded_solver = None


def load_reference_solver():
    """The reference's solver-object module ``pyscf/mpcc/mpcc.py`` (class ``MPCC(mf, lowlevel)``, :3-17), executed
    unmodified. ``MPCC(mf, module)`` instantiates ``module.MPCC_LL(mf)``, pokes ``con_tol``/``max_its`` and runs
    ``kernel()`` — the plug-in seam the GPU module is a drop-in for."""
    load_reference()
    if isinstance(_loaded_solver, Exception) or _loaded_solver is None:
        raise _loaded_solver if _loaded_solver is not None else RuntimeError("reference solver module not loaded")
    return _loaded_solver


def _direct_sum_general(spec, *ops):
    """lib.direct_sum for the strings the reference uses (numpy_helper.py:1060-1127): the two MPCC_LL ones
    plus 'ia,jb->ijab' of the CCSD denominators (rmpccsd_slow.py:168)."""
    if spec == "ia,jb->ijab":
        a, b = ops
        return a[:, None, :, None] + b[None, :, None, :]
    return _direct_sum(spec, *ops)


_loaded_u = None


def load_reference_unrestricted():
    """pyscf/cc/umpcc_lowlevel.py executed from the reference tree (its ``updated_amp`` is the oracle of row f4)."""
    load_reference()
    if isinstance(_loaded_u, Exception) or _loaded_u is None:
        raise _loaded_u if _loaded_u is not None else RuntimeError("unrestricted reference not loaded")
    return _loaded_u


def load_reference_cc():
    """(rmpccsd_slow module, rintermediates module) executed from the reference tree."""
    load_reference()
    if isinstance(_loaded_cc, Exception):
        raise _loaded_cc
    return _loaded_cc


def load_reference():
    """Return (lowlevel_module, diis_module) executed from the reference tree."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise FileNotFoundError(f"reference tree not found under {REFERENCE_ROOT}")
    os.environ["PYTHONBREAKPOINT"] = "0"       # kernel() ends in breakpoint() (dfrmpcc_lowlevel.py:98)
    saved = {k: v for k, v in sys.modules.items() if k == "pyscf" or k.startswith("pyscf.")}

    def mod(name):
        m = types.ModuleType(name)
        sys.modules[name] = m
        return m

    pyscf = mod("pyscf"); pyscf.__path__ = []
    pyscf.__config__ = mod("pyscf.__config__")
    lib = mod("pyscf.lib"); lib.__path__ = []; pyscf.lib = lib
    logger = mod("pyscf.lib.logger"); lib.logger = logger
    logger.INFO = 4
    logger.debug = logger.debug1 = logger.warn = lambda *a, **k: None
    misc = mod("pyscf.lib.misc"); lib.misc = misc

    class StreamObject:
        stdout = sys.stdout
        verbose = 0

    misc.StreamObject = StreamObject
    misc.prange = lambda lo, hi, step: ((p, min(p + step, hi)) for p in range(lo, hi, step))
    misc.H5TmpFile = _MemFile
    misc.to_gpu = None
    nh = mod("pyscf.lib.numpy_helper"); lib.numpy_helper = nh
    for m in (nh, lib):
        m.pack_tril, m.unpack_tril, m.direct_sum = _tril_pack, _tril_unpack, _direct_sum
        m.einsum = lambda s, *t, **k: np.einsum(s, *t, optimize=True)
        m.SYMMETRIC, m.HERMITIAN = 3, 1
    lib.StreamObject = StreamObject

    def run(name, rel):
        spec = importlib.util.spec_from_file_location(name, os.path.join(REFERENCE_ROOT, "pyscf", rel))
        m = importlib.util.module_from_spec(spec)
        sys.modules[name] = m
        spec.loader.exec_module(m)
        return m

    diis = run("pyscf.lib.diis", "lib/diis.py")
    lib.diis = diis
    lowlevel = run("pyscf.mpcc.dfrmpcc_lowlevel", "mpcc/dfrmpcc_lowlevel.py")
    global _loaded_solver
    try:
        _loaded_solver = run("pyscf.mpcc.mpcc", "mpcc/mpcc.py")
    except Exception as exc:
        _loaded_solver = exc
    # CCSD-type residual of the high-level stage (rows f1/f2): rintermediates.py and rmpccsd_slow.py, unmodified.
    # Everything they import besides numpy is faked: lib.einsum/direct_sum/dot, an empty ao2mo, a stub ccsd.CCSD.
    global _loaded_cc
    try:
        lib.dot = lambda a, b, *r, **k: np.dot(a, b)
        lib.direct_sum = nh.direct_sum = _direct_sum_general
        lib.linalg_helper = mod("pyscf.lib.linalg_helper")
        logger.info = logger.note = logger.timer = lambda *a, **k: None
        logger.Logger = type("Logger", (), {"__init__": lambda self, *a, **k: None})
        ao2mo = mod("pyscf.ao2mo"); pyscf.ao2mo = ao2mo
        ao2mo.outcore = types.SimpleNamespace(general_iofree=None)     # default argument of _ChemistsERIs (:992)
        cc = mod("pyscf.cc"); cc.__path__ = []; pyscf.cc = cc
        ccsd = mod("pyscf.cc.ccsd"); cc.ccsd = ccsd
        ccsd.CCSD = type("CCSD", (StreamObject,), {"_keys": set(), "__init__": lambda self, *a, **k: None})
        imd = run("pyscf.cc.rintermediates", "cc/rintermediates.py")
        cc.rintermediates = imd
        slow = run("pyscf.cc.rmpccsd_slow", "cc/rmpccsd_slow.py")
        _loaded_cc = (slow, imd)
        # unrestricted low-level prototype (row f4): only the module-level function updated_amp is used
        global _loaded_u
        try:
            ao2mo._ao2mo = types.SimpleNamespace()
            mod("pyscf.ao2mo._ao2mo")
            pyscf.df = mod("pyscf.df")
            mp = mod("pyscf.mp"); mp.__path__ = []; pyscf.mp = mp
            ump2 = mod("pyscf.mp.ump2"); mp.ump2 = ump2
            ump2.MP2 = type("MP2", (StreamObject,), {"_keys": set(), "__init__": lambda self, *a, **k: None})
            ump2.make_rdm1 = ump2.make_rdm2 = None
            dfn = mod("pyscf.mp.dfump2_native"); mp.dfump2_native = dfn
            dfn.DFMP2 = type("DFMP2", (), {})
            pyscf.scf = mod("pyscf.scf")                      # `from pyscf import scf, cc` at module level (:631)
            # ccsd.py cannot be imported (libcc): the two vector helpers the prototype's run_diis calls are the
            # restatements of oracle/cc_oracle.py (pinned to the text of ccsd.py:705-757 in test_oracle_vs_reference.py)
            from . import cc_oracle
            ccsd.amplitudes_to_vector_s4 = cc_oracle.amplitudes_to_vector_s4
            ccsd.vector_to_amplitudes_s4 = cc_oracle.vector_to_amplitudes_s4
            _loaded_u = run("pyscf.cc.umpcc_lowlevel", "cc/umpcc_lowlevel.py")
        except Exception as exc:
            _loaded_u = exc
    except Exception as exc:                     # keep the low-level oracle usable even if this part fails
        _loaded_cc = exc
    # leave the fake package installed only under private aliases
    for k in [k for k in sys.modules if k == "pyscf" or k.startswith("pyscf.")]:
        sys.modules["_mpcc_refshim_" + k] = sys.modules.pop(k)
    sys.modules.update(saved)
    _loaded = (lowlevel, diis)
    return _loaded
