"""Regenerate tests/golden/*.npz by running the REFERENCE's own code.

Run in the build container only (needs /root/reference):
    python tests/golden/generate_golden.py
Loads ``pyscf/mpcc/dfrmpcc_lowlevel.py`` and ``pyscf/lib/diis.py`` unmodified
through ``oracle/ref_shim.py`` and records, for seeded synthetic problems
(``pyscf_mpcc_b200.synthetic.SyntheticMF``):
  * the full ``kernel()`` trace (iteration count, dE and res per iteration),
    final t1, final t2 (o,o,v,v) — stored whole for the small cases, as
    norm / strided samples for the larger ones;
  * known-answer ``updated_amps()`` calls from t1 = 0 and from a random t1
    (res, dE, t1_new, t2 (v,o,v,o) whole or sampled);
  * a DIIS known-answer sequence (8 random vectors through ``DIIS.update``).
"""
import contextlib
import io
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle import ref_shim                                   # noqa: E402
from pyscf_mpcc_b200.synthetic import SyntheticMF             # noqa: E402

CASES = [  # (nocc, nvir, naux, scale, store_t2_whole)
    (4, 10, 30, 0.3, True),
    (5, 19, 116, 0.1, True),
    (3, 7, 17, 0.2, True),
    (8, 40, 120, 0.1, False),
    (12, 64, 192, 0.1, False),
    (21, 30, 70, 0.1, False),
]
SAMPLE_STRIDE = 97


def parse_trace(text):
    its, de, res = [], [], []
    for line in text.splitlines():
        if line.startswith("It "):
            w = line.split()
            its.append(int(w[1])); de.append(float(w[4])); res.append(float(w[7]))
    return its, de, res


def main():
    ll, diis = ref_shim.load_reference()
    for (o, v, L, scale, whole) in CASES:
        mf = SyntheticMF(o, v, L, seed=0, scale=scale)
        out = {}
        ref = ll.MPCC_LL(mf)
        ref.con_tol, ref.max_its = 1e-6, 50
        # exact per-iteration values: re-run the loop by hand with the reference's methods
        ref.t1 = np.zeros((o, v)); ref.t2 = np.zeros((o, o, v, v))
        ad = diis.DIIS()
        err, count, tr_de, tr_res = np.inf, 0, [], []
        while err > ref.con_tol and count < ref.max_its:
            res, dE, t1n, t2n = ref.updated_amps()
            ref.t1, ref.t2 = ref.run_diis(t1n, t2n, ad)
            count += 1; err = res
            tr_de.append(dE); tr_res.append(res)
        # and confirm the reference's own kernel() prints the same trace
        chk = ll.MPCC_LL(mf); chk.con_tol, chk.max_its = 1e-6, 50
        buf = io.StringIO()
        with contextlib.redirect_stdout(buf):
            E = chk.kernel()
        its, pde, pres = parse_trace(buf.getvalue())
        assert len(its) == count and E == tr_de[-1]
        assert np.array_equal(chk.t1, ref.t1) and np.array_equal(chk.t2, ref.t2)
        out["shape"] = np.array([o, v, L]); out["scale"] = np.array(scale)
        out["iterations"] = np.array(count)
        out["trace_dE"] = np.array(tr_de); out["trace_res"] = np.array(tr_res)
        out["final_t1"] = ref.t1
        out["final_t2_norm"] = np.array(np.linalg.norm(ref.t2))
        out["final_t2_sum"] = np.array(ref.t2.sum())
        out["final_t2_samples"] = ref.t2.reshape(-1)[::SAMPLE_STRIDE].copy()
        if whole:
            out["final_t2"] = ref.t2
        # single-call KATs
        rng = np.random.default_rng(1234)
        for tag, t1 in (("zero", np.zeros((o, v))), ("rand", 0.05 * rng.standard_normal((o, v)))):
            ref.t1 = t1
            res, dE, t1n, t2 = ref.updated_amps()
            out[f"kat_{tag}_t1_in"] = t1
            out[f"kat_{tag}_res"] = np.array(res); out[f"kat_{tag}_dE"] = np.array(dE)
            out[f"kat_{tag}_t1_new"] = t1n
            out[f"kat_{tag}_t2_norm"] = np.array(np.linalg.norm(t2))
            out[f"kat_{tag}_t2_samples"] = np.ascontiguousarray(t2).reshape(-1)[::SAMPLE_STRIDE].copy()
            vec = ref.amplitudes_to_vector(t1n, t2)
            out[f"kat_{tag}_vec_norm"] = np.array(np.linalg.norm(vec))
            out[f"kat_{tag}_vec_samples"] = vec[::SAMPLE_STRIDE].copy()
            if whole:
                out[f"kat_{tag}_t2"] = np.ascontiguousarray(t2)
        np.savez_compressed(os.path.join(HERE, f"mpcc_{o}_{v}_{L}.npz"), **out)
        print(f"({o},{v},{L},{scale}) -> {count} its, dE={tr_de[-1]!r}")

    # DIIS known-answer sequence (reference diis.py, xerr=None path, ring wrap included)
    rng = np.random.default_rng(7)
    n = 257
    base = rng.standard_normal(n)
    ad = diis.DIIS()
    ins, outs = [], []
    for k in range(10):
        x = base + (0.5 ** k) * rng.standard_normal(n)
        ins.append(x.copy())
        outs.append(np.array(ad.update(x.copy())))
    np.savez_compressed(os.path.join(HERE, "diis_kat.npz"), inputs=np.array(ins), outputs=np.array(outs),
                        H=ad._H)
    print("diis_kat written")


if __name__ == "__main__":
    main()
