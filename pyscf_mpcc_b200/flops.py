"""Algorithmic FP64 operation counts of the DF-MPCC iteration."""


def updated_amps_flops(o: int, v: int, L: int) -> float:
    """One ``updated_amps`` (reference dfrmpcc_lowlevel.py:119-151)."""
    nov = o * v
    return (4.0 * L * nov * nov      # eris (:135) + Yvo (:138)
            + 2.0 * nov * nov        # t2 . Fov (:139)
            + 6.0 * L * v * v * o    # Xvo, Jvv's Lov.t1 part, Jvv.Yvo
            + 14.0 * L * o * o * v   # Xoo, Joo.t1, Xvo.Joo, Xoo.t1, (.)Joo, Fov 2nd term, Joo.Yvo
            + 8.0 * L * o * v)       # X, Jvo.X, Fov 1st term, dE


def eris_flops(o: int, v: int, L: int, na: int) -> float:
    """The fused eris/t2 GEMM for a slab of ``na`` virtual indices (M = na*o, N = nov, K = L)."""
    return 2.0 * L * (na * o) * (o * v)


def yvo_flops(o: int, v: int, L: int, na: int) -> float:
    return 2.0 * L * (na * o) * (o * v)


def diis_bytes(o: int, v: int, nd: int) -> float:
    """HBM bytes of one DIIS update with ``nd`` stored vectors: push reads the t2
    row prefixes + xprev + (nd-1) old error vectors and writes x and e; combine
    reads nd x vectors and writes xprev."""
    nov = o * v
    n = nov + nov * (nov + 1) // 2
    return 8.0 * n * ((2 + (nd - 1)) + 2 + nd + 1)
