"""TEST INFRASTRUCTURE — NumPy stand-in for the tensor engine's backend (tests only; the product backend is
``pyscf_mpcc_b200.tensor_engine.CudaBackend``). It lets the CPU suite check the *transcription* of the reference
formulas in ``pyscf_mpcc_b200/rmpccsd.py`` (index strings, signs, permutations, class masks) without a GPU; the
CUDA kernels themselves are checked by the ``-m gpu`` tests against the same reference numbers."""
import numpy as np


class NumpyBackend:
    def __init__(self, eps_o, eps_v):
        self.eo = np.asarray(eps_o, dtype=float)
        self.ev = np.asarray(eps_v, dtype=float)

    def empty(self, n):
        return np.full(int(n), np.nan)

    def zeros(self, n):
        return np.zeros(int(n))

    def upload(self, a):
        return np.array(a, dtype=float).reshape(-1)

    def upload_int(self, a):
        return np.array(a, dtype=np.int32).reshape(-1)

    def download(self, buf):
        return np.array(buf)

    def copy(self, buf):
        return np.array(buf)

    def permute_axpby(self, src, dst, dims, perm, alpha, beta):
        val = alpha * src[:int(np.prod(dims))].reshape(dims).transpose(perm).reshape(-1)
        if beta == 0.0:
            dst[:val.size] = val
        else:
            dst[:val.size] = val + beta * dst[:val.size]

    def gemm_nt(self, a, b, c, m, n, k):
        c[:m * n] = (a[:m * k].reshape(m, k) @ b[:n * k].reshape(n, k).T).reshape(-1)

    def gemv(self, a, x, y, rows, k, trans, alpha, beta):
        mat = a[:rows * k].reshape(k, rows).T if trans else a[:rows * k].reshape(rows, k)
        val = alpha * (mat @ x[:k])
        y[:rows] = val + (beta * y[:rows] if beta != 0.0 else 0.0)

    def gemm_nt_mapped(self, a, b, c, m, n, k, mmap, nmap, alpha, beta):
        prod = a[:m * k].reshape(m, k) @ b[:n * k].reshape(n, k).T
        mi, ni = np.arange(m), np.arange(n)
        off = ((mi // mmap[0]) * mmap[1] + (mi % mmap[0]) * mmap[2])[:, None] \
            + ((ni // nmap[0]) * nmap[1] + (ni % nmap[0]) * nmap[2])[None, :]
        assert np.unique(off).size == off.size, "output map is not injective"
        old = c[off] if beta != 0.0 else 0.0
        c[off] = alpha * prod + beta * old

    def div_denominators(self, buf, rank):
        o, v = self.eo.size, self.ev.size
        eia = self.eo[:, None] - self.ev[None, :]
        if rank == 2:
            buf[:] = (buf.reshape(o, v) / eia).reshape(-1)
        else:
            buf[:] = (buf.reshape(o, o, v, v) / (eia[:, None, :, None] + eia[None, :, None, :])).reshape(-1)

    def div_general(self, buf, dims, evecs, signs):
        den = 0.0
        for d, (e, sgn) in enumerate(zip(evecs, signs)):
            shape = [1] * len(dims)
            shape[d] = dims[d]
            den = den + sgn * np.asarray(e)[:dims[d]].reshape(shape)
        buf[:int(np.prod(dims))] = (buf[:int(np.prod(dims))].reshape(dims) / den).reshape(-1)

    def class_overwrite(self, new, old, rank, act_h, act_p, mask):
        o, v = self.eo.size, self.ev.size
        if rank == 2:
            cls = 2 * act_h[:, None] + act_p[None, :]
        else:
            cls = (8 * act_h[:, None, None, None] + 4 * act_h[None, :, None, None]
                   + 2 * act_p[None, None, :, None] + act_p[None, None, None, :])
        sel = ((mask >> cls.reshape(-1)) & 1).astype(bool)
        new[sel] = old[sel]

    def dot(self, x, y, n):
        return float(np.dot(x[:n], y[:n]))

    def take_2d(self, a, lda, idx, idy, out):
        out[:len(idx) * len(idy)] = a[(idx.astype(np.int64)[:, None] * lda + idy.astype(np.int64)[None, :]).reshape(-1)]

    def takebak_2d(self, out, ldo, a, idx, idy, alpha=1.0):
        pos = (idx.astype(np.int64)[:, None] * ldo + idy.astype(np.int64)[None, :]).reshape(-1)
        assert len(np.unique(pos)) == len(pos)          # the kernel's contract: no duplicate targets
        out[pos] += alpha * a[:len(pos)]

    def pack_tril(self, mat, n, out):
        r, c = np.tril_indices(n)
        out[:r.size] = mat[:n * n].reshape(n, n)[r, c]

    def unpack_tril(self, packed, n, out):
        r, c = np.tril_indices(n)
        full = np.zeros((n, n))
        full[c, r] = packed[:r.size]
        full[r, c] = packed[:r.size]
        out[:n * n] = full.reshape(-1)

    def vec_diis_push(self, x, n, xprev, prev_x, prev_c, x_slot, e_slot, e_list, nd, slot, first):
        if first:
            x_slot[:n] = x[:n]
            return None
        if prev_x:
            prev = np.zeros(n)
            for xk, ck in zip(prev_x, prev_c):
                prev += xk[:n] * ck
        else:
            prev = np.array(xprev[:n])
        x_slot[:n] = x[:n]
        e_slot[:n] = x[:n] - prev
        return np.array([float(np.dot(e_slot[:n], e_slot[:n] if k == slot else e_list[k][:n])) for k in range(nd)])

    def combine(self, xs, coeffs, out, n):
        acc = np.zeros(n)
        for xk, ck in zip(xs, coeffs):
            acc += xk[:n] * ck
        out[:n] = acc
