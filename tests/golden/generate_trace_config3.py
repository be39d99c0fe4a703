"""Iteration trace at BASELINE.json configs[2] — nocc=64, nvir=512, naux=1536, the shape the metric is quoted on.

Run in the build container (about 50 GB of RAM, a few minutes per iteration on 8 cores):
    python tests/golden/generate_trace_config3.py [--shape 64 512 1536]
The reference as written would need more than an hour per iteration of single-threaded ``np.einsum`` at this shape
(SURVEY.md §8d), so the trace comes from the GEMM-shaped oracle ``LowLevelOracle.updated_amps_packed`` +
``PulayDIIS`` — pinned to the reference itself at 1e-13 on the small shapes (``tests/test_oracle_vs_reference.py``)
and bitwise-equivalent in structure to ``kernel()``'s loop (dfrmpcc_lowlevel.py:82-95). Inputs:
``pyscf_mpcc_b200.synthetic.host_blocks(64, 512, 1536, seed=0)`` — the generator of every other fixture, drawn in
chunks; ``bench.py`` and ``tests/test_gpu_fullsize.py`` build the same tensors on the GPU box.

Recorded per iteration: dE, res, |t1_new|, |t2|, |DIIS vector|, the DIIS coefficients; plus t1_new of the first
iteration and the final (extrapolated) t1.
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle.mpcc_oracle import LowLevelOracle, PulayDIIS                # noqa: E402
from pyscf_mpcc_b200.synthetic import default_scale, host_blocks        # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shape", type=int, nargs=3, default=[64, 512, 1536])
    ap.add_argument("--max-its", type=int, default=6, help="memory bound: (2n) vectors of N doubles are alive at iteration n")
    ap.add_argument("--slab", type=int, default=32)
    args = ap.parse_args()
    o, v, L = args.shape
    scale = default_scale(o, v)
    t0 = time.perf_counter()
    Loo, Lov, Lvv, eo, ev = host_blocks(o, v, L, seed=0, scale=scale)
    print(f"inputs in {time.perf_counter() - t0:.1f} s", flush=True)
    ll = LowLevelOracle(blocks=(Loo, Lov, Lvv, np.diag(eo), np.diag(ev), np.concatenate([eo, ev])), mode="blas")
    t1 = np.zeros((o, v))
    ad = PulayDIIS()
    out = {"shape": np.array([o, v, L]), "scale": np.array(scale), "seed": np.array(0)}
    tr = {k: [] for k in ("dE", "res", "t1_new_norm", "t2_norm", "vec_norm", "t1_next_norm")}
    coeffs = np.zeros((args.max_its, 7))
    err, count = np.inf, 0
    while err > 1e-6 and count < args.max_its:                # kernel()'s loop, con_tol = 1e-6 (mpcc.py:14)
        t0 = time.perf_counter()
        res, dE, t1n, vec, t2ss = ll.updated_amps_packed(t1, slab=args.slab)
        t_upd = time.perf_counter() - t0
        if count == 0:
            out["it1_t1_new"] = t1n.copy()
        tr["dE"].append(dE); tr["res"].append(res)
        tr["t1_new_norm"].append(float(np.linalg.norm(t1n)))
        tr["t2_norm"].append(float(np.sqrt(t2ss)))
        tr["vec_norm"].append(float(np.sqrt(np.vdot(vec, vec))))
        t0 = time.perf_counter()
        xnew = ad.update(vec)
        del vec
        t1 = np.array(xnew[:o * v]).reshape(o, v)
        if ad.last_coeff is not None:
            coeffs[count, :len(ad.last_coeff)] = ad.last_coeff
        tr["t1_next_norm"].append(float(np.linalg.norm(t1)))
        count += 1
        err = res
        print(f"It {count} dE={dE!r} res={res!r} ({t_upd:.0f} s + DIIS {time.perf_counter() - t0:.0f} s)", flush=True)
    out["iterations"] = np.array(count)
    out["converged"] = np.array(err <= 1e-6)
    for k, val in tr.items():
        out["trace_" + k] = np.array(val)
    out["diis_coeff"] = coeffs[:count]
    out["final_t1"] = t1
    np.savez_compressed(os.path.join(HERE, f"trace_{o}_{v}_{L}.npz"), **out)
    print(f"({o},{v},{L}) -> {count} iterations, converged={err <= 1e-6}, dE = {tr['dE'][-1]!r}")


if __name__ == "__main__":
    main()
