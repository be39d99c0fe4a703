"""Host-side partitioning of the DF-MPCC doubles work over GPUs.

The (nov x nov) amplitude matrix is sharded by its first virtual index ``a``:
rank r owns a in [a_begin, a_end), i.e. rows [a_begin*o, a_end*o) of the matrix
in [a,b,i,j] storage order. With that split the eris/t2/Yvo contractions and the
t2 part of the DIIS vector are rank-local; per iteration only ``nov + 1``
doubles (partial singles residual + energy) and <= 6 DIIS dot products are
all-reduced.
"""
from __future__ import annotations


def slab_bounds(nvir: int, world_size: int, rank: int):
    """Contiguous, near-equal split of the virtual index. Every rank gets >= 1."""
    if world_size > nvir:
        raise ValueError(f"more ranks ({world_size}) than virtual orbitals ({nvir})")
    base, extra = divmod(nvir, world_size)
    begin = rank * base + min(rank, extra)
    end = begin + base + (1 if rank < extra else 0)
    return begin, end


def tril_offset(row: int) -> int:
    """Start of ``row`` in a row-major packed lower triangle (pack_tril.c:99-107)."""
    return row * (row + 1) // 2


def packed_segment(nocc: int, a_begin: int, a_end: int):
    """[start, stop) of the packed-lower-triangle elements owned by a slab."""
    return tril_offset(a_begin * nocc), tril_offset(a_end * nocc)


def local_vector_length(nocc: int, nvir: int, a_begin: int, a_end: int) -> int:
    """t1 part (replicated) + the slab's packed t2 rows."""
    s0, s1 = packed_segment(nocc, a_begin, a_end)
    return nocc * nvir + (s1 - s0)
