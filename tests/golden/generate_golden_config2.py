"""Golden vectors at BASELINE.json configs[1]'s shape (benzene/cc-pVTZ: nocc=21, nvir=243, naux=654), produced by
the REFERENCE's own code.

Run in the build container only (needs /root/reference; about 50 s of single-threaded ``np.einsum`` per
``updated_amps`` call, about ten minutes in total):
    python tests/golden/generate_golden_config2.py
``pyscf/mpcc/dfrmpcc_lowlevel.py`` and ``pyscf/lib/diis.py`` run unmodified through ``oracle/ref_shim.py`` on the
seeded synthetic problem ``SyntheticMF(21, 243, 654, seed=0, scale=default_scale(21, 243))`` (PySCF itself cannot be
imported here, so the tensors are "shape-equivalent synthetic"). Recorded: the iteration count of
``kernel()``'s loop, dE and res of every iteration, the final t1, and norm / sum / strided samples of the final
(o,o,v,v) t2 and of the first iteration's (v,o,v,o) t2 and DIIS vector.
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_shim                                             # noqa: E402
from pyscf_mpcc_b200.synthetic import SyntheticMF, default_scale        # noqa: E402

SHAPE = (21, 243, 654)
SAMPLE_STRIDE = 9973


def main():
    ll, diis = ref_shim.load_reference()
    o, v, L = SHAPE
    scale = default_scale(o, v)
    mf = SyntheticMF(o, v, L, seed=0, scale=scale)
    ref = ll.MPCC_LL(mf)
    ref.con_tol, ref.max_its = 1e-6, 50
    ref.t1 = np.zeros((o, v))
    ref.t2 = np.zeros((o, o, v, v))
    ad = diis.DIIS()
    err, count, tr_de, tr_res, secs = np.inf, 0, [], [], []
    out = {}
    while err > ref.con_tol and count < ref.max_its:        # the loop of kernel(), dfrmpcc_lowlevel.py:82-95
        t0 = time.perf_counter()
        res, dE, t1n, t2n = ref.updated_amps()
        t_upd = time.perf_counter() - t0
        if count == 0:
            t2c = np.ascontiguousarray(t2n)
            out["it1_t1_new"] = t1n
            out["it1_t2_norm"] = np.array(np.linalg.norm(t2c))
            out["it1_t2_samples"] = t2c.reshape(-1)[::SAMPLE_STRIDE].copy()
            vec = ref.amplitudes_to_vector(t1n, t2n)
            out["it1_vec_norm"] = np.array(np.linalg.norm(vec))
            out["it1_vec_samples"] = vec[::SAMPLE_STRIDE].copy()
            del t2c, vec
        t0 = time.perf_counter()
        ref.t1, ref.t2 = ref.run_diis(t1n, t2n, ad)
        t_diis = time.perf_counter() - t0
        count += 1
        err = res
        tr_de.append(dE)
        tr_res.append(res)
        secs.append((t_upd, t_diis))
        print(f"It {count} dE={dE!r} res={res!r}  updated_amps {t_upd:.1f} s, run_diis {t_diis:.1f} s", flush=True)
    out["shape"] = np.array(SHAPE)
    out["scale"] = np.array(scale)
    out["iterations"] = np.array(count)
    out["trace_dE"] = np.array(tr_de)
    out["trace_res"] = np.array(tr_res)
    out["final_t1"] = ref.t1
    out["final_t2_norm"] = np.array(np.linalg.norm(ref.t2))
    out["final_t2_sum"] = np.array(ref.t2.sum())
    out["final_t2_samples"] = ref.t2.reshape(-1)[::SAMPLE_STRIDE].copy()
    out["sample_stride"] = np.array(SAMPLE_STRIDE)
    out["reference_seconds"] = np.array(secs)        # (updated_amps, run_diis) per iteration in the build container
    np.savez_compressed(os.path.join(HERE, f"config2_mpcc_{o}_{v}_{L}.npz"), **out)
    print(f"{SHAPE} -> {count} iterations, dE = {tr_de[-1]!r}")


if __name__ == "__main__":
    main()
