"""Solver object ("high level") of the MPCC drop-in.

Mirror of ``pyscf/mpcc/mpcc.py:3-17``: the constructor receives the mean-field
object and a *low-level module*, instantiates ``lowlevel.MPCC_LL(mf)``, sets
``con_tol = 1e-6`` / ``max_its = 50`` from outside and runs ``kernel()``
immediately — construction is the solve. The later ``kernel()`` call users make
(``examples/mpcc/01-dfrmpcc.py:26``) resolves to the base-class no-op
(``pyscf/lib/misc.py:604-612``) in the reference; here it returns the stored
energy so the call is still harmless and a bit more useful.

Because the seam is a module, the reference's own class can also be pointed at
the GPU path without this file: ``pyscf.mpcc.mpcc.MPCC.__init__(obj, mf,
pyscf_mpcc_b200.lowlevel)``.
"""
from __future__ import annotations

import sys


class StreamObject:
    """The sliver of ``pyscf.lib.StreamObject`` (``pyscf/lib/misc.py:578-698``) that MPCC touches."""
    verbose = 0
    stdout = sys.stdout
    _keys = {"output", "verbose", "stdout", "max_memory"}

    def kernel(self, *args, **kwargs):
        pass

    def set(self, *args, **kwargs):
        for k, v in kwargs.items():
            setattr(self, k, v)
        return self

    def run(self, *args, **kwargs):
        self.set(**kwargs)
        self.kernel(*args)
        return self


class MPCC(StreamObject):
    CON_TOL = 1e-6
    MAX_ITS = 50

    def __init__(self, mf, lowlevel, **lowlevel_kwargs):
        self.mol = getattr(mf, "mol", None)
        self._scf = mf
        solver = lowlevel.MPCC_LL(mf, **lowlevel_kwargs)
        solver.con_tol = self.CON_TOL
        solver.max_its = self.MAX_ITS
        self.lowlevel = solver
        self.e_corr = solver.kernel()

    def kernel(self, *args, **kwargs):
        return self.e_corr
