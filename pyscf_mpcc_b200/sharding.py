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


# Hardware ratio behind the cost model of balanced_bounds: bytes the DIIS bookkeeping moves per packed
# element and iteration with a full subspace (push: T, xprev, 5 older error vectors read, x and e written;
# combine: 6 x read, xprev written = 16 doubles), over the HBM rate, against the 3 * naux FP64 flops per
# (nov x nov) matrix element of the two big contractions (2 for Yvo, 1 for the half eris Gram matrix that
# the mirror scheme computes) at the DMMA rate. Both rates are the ones measured on this pool's B200
# (DESIGN.md §5): 6.5e12 B/s and 3.4e13 flop/s.
_DIIS_BYTES_PER_ELEM = 16 * 8
_HBM_BYTES_PER_S = 6.5e12
_FP64_FLOP_PER_S = 3.4e13


def balanced_bounds(nocc: int, nvir: int, naux: int, world_size: int, diis: bool = True):
    """Slab bounds [b_0=0, .., b_P=nvir] that equalise the per-rank time of one iteration.

    The contractions cost the same for every virtual index a, but row r of the packed lower triangle that
    the DIIS vector holds (dfrmpcc_lowlevel.py:162-168) has r+1 elements, so with equal slabs the last rank
    streams (2P-1) times the DIIS bytes of the first one and everybody waits for it at the all-reduce of the
    B-matrix dots. In units of "one virtual index worth of GEMM time" the cumulative cost up to index x is
    C(x) = x + rho * x^2 / nvir with rho = DIIS time / GEMM time for the whole problem; the bounds are the
    points where C crosses r/P of its total, rounded to the tile granularity of the eris/Yvo kernels
    (128-row tiles hold 128/nocc virtual indices). Falls back to the equal split when slabs get too small
    for the rounding to make sense."""
    if world_size > nvir:
        raise ValueError(f"more ranks ({world_size}) than virtual orbitals ({nvir})")
    equal = [slab_bounds(nvir, world_size, r)[0] for r in range(world_size)] + [nvir]
    gran = 128 // nocc if (nocc <= 128 and 128 % nocc == 0) else 1
    if world_size == 1 or not diis or nvir < 8 * gran * world_size:
        return equal
    t_diis = _DIIS_BYTES_PER_ELEM * 0.5 / _HBM_BYTES_PER_S          # per element of the full matrix (half is packed)
    t_gemm = 3.0 * naux / _FP64_FLOP_PER_S
    rho = t_diis / t_gemm
    total = nvir * (1.0 + rho)
    bounds = [0]
    for r in range(1, world_size):
        target = total * r / world_size
        # x + rho x^2 / nvir = target
        x = (-1.0 + (1.0 + 4.0 * rho * target / nvir) ** 0.5) * nvir / (2.0 * rho)
        b = int(round(x / gran)) * gran
        b = max(b, bounds[-1] + gran)
        bounds.append(b)
    bounds.append(nvir)
    if any(hi <= lo for lo, hi in zip(bounds[:-1], bounds[1:])):
        return equal
    return bounds


def modelled_cost(nocc: int, nvir: int, naux: int, bounds):
    """Per-rank cost of ``bounds`` in the units of balanced_bounds (for tests and DESIGN.md)."""
    rho = (_DIIS_BYTES_PER_ELEM * 0.5 / _HBM_BYTES_PER_S) / (3.0 * naux / _FP64_FLOP_PER_S)
    return [(hi - lo) + rho * (hi * hi - lo * lo) / nvir for lo, hi in zip(bounds[:-1], bounds[1:])]


def tril_count_below(f, nov):
    """Number of packed-lower-triangle elements of the (nov x nov) matrix whose flat row-major index is < f
    (f scalar or int64 array): full rows r' < r contribute r'+1 each, row r the columns < min(c, r+1)."""
    import numpy as np
    f = np.asarray(f, dtype=np.int64)
    r, c = f // nov, f % nov
    return r * (r + 1) // 2 + np.minimum(c, r + 1)


def column_owned_prefix(nocc: int, nvir: int, b_begin: int, b_end: int):
    """Layout of the column-owned DIIS vector of a slab (include/mpcc_b200.h, mpccb_bind_diis_columns).

    The slab [b_begin, b_end) of the [a,b,i,j] buffer also holds, read as t2[A=b, B=a, I=j, J=i], every
    element of the reference's (nov x nov) matrix whose second virtual index B lies in the slab. The local
    block k = (B - b_begin) * nvir + A covers the flat reference indices [F0, F0 + o^2), F0 = (A*nvir + B) * o^2;
    the packed triangle keeps tril_count_below(F0 + o^2) - tril_count_below(F0) of them. Returns the int64
    prefix sums (na*nvir + 1 entries; the last one is the length of the slab's t2 part of the vector)."""
    import numpy as np
    o, v = nocc, nvir
    nov = o * v
    B = np.arange(b_begin, b_end, dtype=np.int64)[:, None]
    A = np.arange(v, dtype=np.int64)[None, :]
    F0 = (A * v + B) * (o * o)
    n = tril_count_below(F0 + o * o, nov) - tril_count_below(F0, nov)
    prefix = np.zeros(n.size + 1, dtype=np.int64)
    np.cumsum(n.reshape(-1), out=prefix[1:])
    return prefix


def column_owned_positions(nocc: int, nvir: int, b_begin: int, b_end: int):
    """Host restatement of the element walk of k_diis_push_cols (tests, and the reference-order gather on
    the host): for every element of the slab's column-owned vector, in storage order, (index into the flat
    local slab buffer, index into the reference packed triangle)."""
    import numpy as np
    o, v = nocc, nvir
    nov = o * v
    prefix = column_owned_prefix(o, v, b_begin, b_end)
    local_idx = np.empty(int(prefix[-1]), dtype=np.int64)
    ref_idx = np.empty(int(prefix[-1]), dtype=np.int64)
    t = np.arange(o * o, dtype=np.int64)
    I, J = t // o, t % o                      # reference block order (I, J); stored at local [i=J][j=I]
    for k in range((b_end - b_begin) * v):
        if prefix[k + 1] == prefix[k]:
            continue
        bl, A = divmod(k, v)
        f = (A * v + b_begin + bl) * o * o + t
        keep = (f % nov) <= (f // nov)
        r, c = f[keep] // nov, f[keep] % nov
        ref_idx[prefix[k]:prefix[k + 1]] = r * (r + 1) // 2 + c
        local_idx[prefix[k]:prefix[k + 1]] = k * o * o + J[keep] * o + I[keep]
    return local_idx, ref_idx
