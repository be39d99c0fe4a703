"""Restricted density-fitted MPCC bound to the B200 low-level module
(counterpart of ``pyscf/mpcc/dfrmpcc.py:4-12``)."""
from __future__ import annotations

from . import lowlevel as _b200_lowlevel
from . import mpcc as _mpcc


class RMPCC(_mpcc.MPCC):
    def __init__(self, mf, **lowlevel_kwargs):
        with_df = getattr(mf, "with_df", None)
        if not with_df:
            raise RuntimeError("Chosen RMPCC variant is not implemented")   # same text as pyscf/mpcc/__init__.py:17
        super().__init__(mf, _b200_lowlevel, **lowlevel_kwargs)
        self.with_df = with_df
        self._keys = set(self._keys) | {"with_df"}
