"""Throughput of the CCSD-type residual (rows f1/f2) on synthetic DF tensors: one rmpccsd.update_amps,
CUDA-event timed; FP64 TFLOP/s from the GEMM flops the engine issued (2*M*N*K per contraction).

    python tools/cc_bench.py [o,v,L ...]
"""
import json
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from pyscf_mpcc_b200 import rmpccsd
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.synthetic import device_blocks, default_scale


def run(o, v, L):
    Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
    ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])))
    t0 = torch.cuda.Event(enable_timing=True); t1e = torch.cuda.Event(enable_timing=True)
    t0.record()
    eris = ll.high_level_eris()
    t1e.record(); torch.cuda.synchronize()
    ms_eris = t0.elapsed_time(t1e)
    eng = eris.eng
    fl_eris = eng.flops
    cc = types.SimpleNamespace(cc2=False)
    emp2, t1, t2 = rmpccsd.init_amps(cc, eris)
    act_h, act_p = list(range(o // 2, o)), list(range(v // 4))
    out = {}
    reps_ms = []
    for rep in range(4):
        eng.flops = 0.0; eng.gemms = 0
        launches0 = ll._ctx.launch_count()
        t0.record()
        t1n, t2n = rmpccsd.update_amps(cc, t1, t2, eris, act_h, act_p, [0, 1, 2], list(range(15)))
        t1e.record(); torch.cuda.synchronize()
        ms = t0.elapsed_time(t1e)
        reps_ms.append(ms)
        out = {"nocc": o, "nvir": v, "naux": L, "ms_update_amps": ms, "gemm_flops": eng.flops, "gemms": eng.gemms,
               "launches": ll._ctx.launch_count() - launches0, "tflops": eng.flops / (ms * 1e-3) * 1e-12,
               "pp_ladder_flops": 2.0 * v ** 4 * (L + o * o + o), "ms_eris_build": ms_eris, "eris_flops": fl_eris,
               "emp2": emp2, "e_corr_after": rmpccsd.energy(cc, t1n, t2n, eris)}
    # per-contraction profile (one more call, synchronising after every einsum)
    eng.profile = {}
    t0.record()
    rmpccsd.update_amps(cc, t1, t2, eris, act_h, act_p, [0, 1, 2], list(range(15)))
    t1e.record(); torch.cuda.synchronize()
    tot = t0.elapsed_time(t1e)
    top = sorted(eng.profile.items(), key=lambda kv: -kv[1][0])[:45]
    out["profiled_ms_total"] = tot
    out["einsum_ms_total"] = sum(v[0] for v in eng.profile.values())
    out["top_einsums"] = [{"spec": k, "ms": round(v[0], 3), "tflops": round(v[1] / max(v[0], 1e-9) * 1e-9, 2), "calls": v[2]}
                          for k, v in top]
    eng.profile = None
    out["reps_ms"] = reps_ms          # first call builds the cached copies of the integral blocks
    out["perm_cache_bytes"] = getattr(eng, "perm_cache_bytes", 0)
    return out


if __name__ == "__main__":
    shapes = [tuple(int(x) for x in s.split(",")) for s in sys.argv[1:]] or [(8, 40, 120), (16, 64, 192), (32, 128, 384), (32, 192, 576)]
    rows = []
    for s in shapes:
        rows.append(run(*s))
        print(json.dumps(rows[-1]), flush=True)
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/cc_bench.json", "w"), indent=1)
