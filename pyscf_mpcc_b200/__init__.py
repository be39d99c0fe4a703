"""pyscf_mpcc_b200 — B200-native DF-MPCC amplitude iteration.

Factory functions follow ``pyscf/mpcc/__init__.py:3-17``: unrestricted and
generalised mean fields are rejected, restricted ones must carry ``with_df``.
PySCF itself is optional: the SCF type test uses PySCF classes when importable
and falls back to the class name otherwise (synthetic mean fields are plain RHF-like).
"""
from __future__ import annotations

__version__ = "0.1.0"


def _scf_kind(mf) -> str:
    names = {k.__name__ for k in type(mf).__mro__}
    if "UHF" in names or "UKS" in names:
        return "U"
    if "GHF" in names or "GKS" in names:
        return "G"
    return "R"


def MPCC(mf, **kwargs):
    kind = _scf_kind(mf)
    if kind == "U":
        raise RuntimeError("UMPCC is not implemented")
    if kind == "G":
        raise RuntimeError("GMPCC is not implemented")
    return RMPCC(mf, **kwargs)


def RMPCC(mf, **kwargs):
    from . import dfrmpcc
    if getattr(mf, "with_df", None):
        return dfrmpcc.RMPCC(mf, **kwargs)
    raise RuntimeError("Chosen RMPCC variant is not implemented")
