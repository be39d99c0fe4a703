#!/usr/bin/env python
"""DF-MPCC iteration benchmark (BASELINE.json metric: s/iteration and FP64 TFLOP/s).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One *step* = one trip of ``MPCC_LL.kernel()``'s loop = ``updated_amps()`` +
``run_diis()`` (reference ``pyscf/mpcc/dfrmpcc_lowlevel.py:82-95``) on the
synthetic target shape nocc=64, nvir=512, naux=1536 (BASELINE.json configs[2],
the configuration the metric is quoted on; it fits one GPU). FP64 throughout.

* ``value``  — TFLOP/s (algorithmic flops of SURVEY.md §8d per iteration) with the
  amplitudes resident in HBM (``MPCC_LL.step_device``); CUDA-event timed.
* ``e2e``    — same metric through the public API the reference user calls
  (``updated_amps()`` / ``run_diis()`` with host numpy t1 in and out every step).
* ``roofline`` — the dominant kernel (fused eris/t2 DMMA GEMM), timed live with
  CUDA events on its stream inside the timed region, against the FP64 GEMM rate
  measured on this pool's B200 (cuBLAS DGEMM yardstick, profiles/r1_fp64_microbench.txt).
* ``cpu_baseline`` — the oracle port of the reference's einsum path on a bounded
  sample (rank 0, N=1 only).

With N > 1 (torchrun, one rank per GPU, NCCL) the virtual index is sharded over
ranks (strong scaling: the problem is fixed).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# FP64 GEMM rate measured on this pool's B200 with cuBLAS DGEMM on the eris-shaped product
# (32768 x 8192 x 1536; tools/fp64_yardstick.py -> profiles/r1_fp64_microbench.txt). MEASURED_PEAKS.json
# holds no FP64 figure. Nominal: 148 SMs x 64 DFMA/clk x 1.965 GHz = 37.2 TFLOP/s.
FP64_PEAK_TFLOPS_MEASURED = 36.07
FP64_PEAK_TFLOPS_NOMINAL = 37.2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nocc", type=int, default=64)
    ap.add_argument("--nvir", type=int, default=512)
    ap.add_argument("--naux", type=int, default=1536)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tma", action="store_true")
    ap.add_argument("--diis-layout", default=None, choices=["columns", "rows"],
                    help="t2 part of the DIIS vector over ranks: balanced column ownership (default for N>1) or row segments")
    ap.add_argument("--slabs", default="equal", choices=["balanced", "equal"],
                    help="split of the virtual index over ranks (sharding.balanced_bounds / equal widths)")
    return ap.parse_args()


def ncu_traffic(o, v, L, world):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel from the committed
    `ncu --set full` capture (profiles/ncu_traffic.json, written from the .ncu-rep); None for shapes not captured."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            tab = json.load(f)
        return tab.get(f"{o},{v},{L},{world}", {})
    except Exception:
        return None


class ClockSampler:
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self.reasons = set()
        self._stop = threading.Event()
        self._th = None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self.samples.append((float(f[0]), float(f[1])))
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                    if val.lower().startswith("active"):
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._th = threading.Thread(target=self._run, daemon=True)
        self._th.start()

    def stop(self):
        self._stop.set()
        if self._th:
            self._th.join(timeout=6)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": sorted(self.reasons)}
        sm = sorted(s[0] for s in self.samples)
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": self.samples[0][1], "reasons": sorted(self.reasons),
                "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU legs (oracle port of the reference's einsum path). Only these functions touch oracle/.
# ---------------------------------------------------------------------------------------------
def cpu_sample_inputs(o, v, L, seed):
    rng = np.random.default_rng(seed)
    amp = 0.1 / np.sqrt(L)
    Jvo = rng.standard_normal((L, v, o)) * amp
    Lov = rng.standard_normal((L, o, v)) * amp
    eo = np.sort(rng.uniform(-2.0, -0.5, o))
    ev = np.sort(rng.uniform(0.3, 3.0, v))
    return Jvo, Lov, eo, ev


def cpu_step(inputs, na, mode):
    """One bounded sample: the two O(L nov^2) contractions + t2 build on `na` virtual rows."""
    from oracle.mpcc_oracle import doubles_sample
    Jvo, Lov, eo, ev = inputs
    t0 = time.perf_counter()
    _, _, flops = doubles_sample(Jvo, Lov, eo, ev, 0, na, mode=mode)
    return time.perf_counter() - t0, flops


def cpu_baseline(o, v, L, seed):
    from pyscf_mpcc_b200.flops import updated_amps_flops
    inputs = cpu_sample_inputs(o, v, L, seed)
    na_e = 1
    dt_e, fl_e = cpu_step(inputs, na_e, "einsum")
    na_b = max(1, min(v, 64))
    cpu_step(inputs, min(4, na_b), "blas")          # thread-pool warm-up
    dt_b, fl_b = cpu_step(inputs, na_b, "blas")
    total = updated_amps_flops(o, v, L)
    try:
        import threadpoolctl
        blas_threads = max((p.get("num_threads", 1) for p in threadpoolctl.threadpool_info()), default=1)
    except Exception:
        blas_threads = os.cpu_count()
    return {
        "value": fl_e / dt_e * 1e-12, "unit": "TFLOP/s", "cores": 1, "kind": "port",
        "host_cores": os.cpu_count(),
        "sample": (f"oracle einsum port of eris+t2+Yvo (dfrmpcc_lowlevel.py:135-138) on {na_e} of {v} virtual rows "
                   f"({fl_e:.3g} flop, {dt_e:.2f} s); np.einsum is single-threaded as in the reference"),
        "s_per_iteration_extrapolated": total / (fl_e / dt_e),
        "blas_fair": {"value": fl_b / dt_b * 1e-12, "unit": "TFLOP/s", "cores": int(blas_threads),
                      "sample": f"same contractions as GEMMs (OpenBLAS) on {na_b} of {v} virtual rows, {dt_b:.2f} s",
                      "s_per_iteration_extrapolated": total / (fl_b / dt_b),
                      "note": "not the reference: what a BLAS-backed CPU implementation would reach"},
    }


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port) on bounded samples of the same workload."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from pyscf_mpcc_b200.flops import updated_amps_flops
    o, v, L = args.nocc, args.nvir, args.naux
    inputs = cpu_sample_inputs(o, v, L, args.seed)
    na = 1
    for _ in range(args.warmup):
        cpu_step(inputs, na, "einsum")
    t, fl = 0.0, 0.0
    for _ in range(args.steps):
        dt, f = cpu_step(inputs, na, "einsum")
        t += dt
        fl += f
    rate = fl / t * 1e-12
    total = updated_amps_flops(o, v, L)
    line = {
        "impl": "reference", "metric": "dfmpcc_iteration_fp64_tflops", "value": rate, "unit": "TFLOP/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": total / (rate * 1e12) * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"synthetic DF-MPCC nocc={o} nvir={v} naux={L}",
                   "ms_per_step_is": "extrapolated from the sample rate to one full iteration"},
        "cpu_baseline": {"value": rate, "unit": "TFLOP/s", "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
                         "sample": f"per step: oracle einsum port of eris+t2+Yvo on {na} of {v} virtual rows "
                                   f"({fl / args.steps:.3g} flop, {t / args.steps:.2f} s)"},
        "e2e": {"value": rate, "unit": "TFLOP/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from pyscf_mpcc_b200 import flops as F
    from pyscf_mpcc_b200.diis import DIIS
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import device_blocks, default_scale

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    o, v, L = args.nocc, args.nvir, args.naux
    K, W = args.steps, args.warmup

    # synthetic DF tensors generated on the device (same seed on every rank -> replicated inputs)
    Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=args.seed, scale=default_scale(o, v), device=dev)
    foo = torch.diag(eo)
    fvv = torch.diag(ev)
    ll = MPCC_LL(blocks=(Loo, Lov, Lvv, foo, fvv, torch.cat([eo, ev])), device=dev, use_tma=not args.no_tma,
                 slabs=args.slabs, diis_layout=args.diis_layout)
    ll.verbose_print = False
    ctx = ll._ctx
    nov = o * v

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- warm-up: W iterations of a first solve (allocations, module load, DIIS buffers) ----
    t1 = torch.zeros(nov, dtype=torch.float64, device=dev)
    adiis = DIIS()
    for _ in range(max(W, 3)):
        _, _, t1 = ll.step_device(t1, adiis)
    del adiis
    barrier()

    # ---- timed: the first K iterations of a fresh solve, amplitudes resident in HBM ----
    ctx.call("mpccb_set_option", b"profile", 1)
    ctx.profile_read()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    t1 = torch.zeros(nov, dtype=torch.float64, device=dev)
    adiis = DIIS().reserve(ll)          # history buffers allocated before the timed region
    trace = []
    launches0 = ctx.launch_count()
    e0 = torch.cuda.Event(enable_timing=True)
    e1 = torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(K):
        res, dE, t1 = ll.step_device(t1, adiis)
        trace.append((dE, res))
    e1.record()
    barrier()
    ms_dev = max_over_ranks(e0.elapsed_time(e1))
    launches = ctx.launch_count() - launches0
    stages = ctx.profile_read()
    ctx.call("mpccb_set_option", b"profile", 0)
    del adiis

    # ---- e2e: the public API with host numpy amplitudes in and out each step ----
    ll.h2d_bytes = ll.d2h_bytes = 0
    ll.t1 = np.zeros((o, v))
    adiis = DIIS().reserve(ll)
    barrier()
    e0.record()
    for _ in range(K):
        res, dE, t1n, t2 = ll.updated_amps()
        ll.t1, ll.t2 = ll.run_diis(t1n, t2, adiis)
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if rank == 0 else None
    h2d, d2h = ll.h2d_bytes / K, ll.d2h_bytes / K + 7 * 8

    flop_it = F.updated_amps_flops(o, v, L)
    tflops = flop_it * K / (ms_dev * 1e-3) * 1e-12
    tflops_e2e = flop_it * K / (ms_e2e * 1e-3) * 1e-12

    # roofline of the dominant kernel on this rank (slab of na rows), live CUDA-event timing
    na = ll._na
    er_ms, er_n = stages["eris_t2"]
    yv_ms, yv_n = stages["yvo"]
    launched, total_tiles = ctx.t2_tiles()
    er_alg = F.eris_flops(o, v, L, na)                       # algorithmic: every (a,b) block of the slab
    er_exec = er_alg * launched / max(total_tiles, 1)        # executed: the mirror scheme computes one block per pair
    er_tf = er_exec / (er_ms / max(er_n, 1) * 1e-3) * 1e-12
    er_tf_alg = er_alg / (er_ms / max(er_n, 1) * 1e-3) * 1e-12
    yv_tf = F.yvo_flops(o, v, L, na) / (yv_ms / max(yv_n, 1) * 1e-3) * 1e-12
    traffic = ncu_traffic(o, v, L, world) or {}
    eris_info = {"kernel": ("gemm_tma_kernel<1,WN>" if not args.no_tma else "gemm_t2_kernel")
                           + " (fused eris -> t2 with mirror scheme, TMA-fed DMMA.8x8x4)",
                 "achieved": er_tf, "frac": er_tf / FP64_PEAK_TFLOPS_MEASURED,
                 "frac_of_nominal": er_tf / FP64_PEAK_TFLOPS_NOMINAL, "ms_per_launch": er_ms / max(er_n, 1),
                 "flops_counted": "executed: eris is a Gram matrix, the kernel computes one tile of every mirror pair "
                                  f"({launched} of {total_tiles} tiles) and writes both t2 blocks",
                 "algorithmic_tflops": er_tf_alg, "traffic": traffic.get("eris_t2_bytes")}
    yvo_info = {"kernel": "gemm_tma_kernel<0,4> (Yvo = t2 . Lov, TMA-fed DMMA.8x8x4, 4-D tensor map over t2)",
                "achieved": yv_tf, "frac": yv_tf / FP64_PEAK_TFLOPS_MEASURED,
                "frac_of_nominal": yv_tf / FP64_PEAK_TFLOPS_NOMINAL, "ms_per_launch": yv_ms / max(yv_n, 1),
                "traffic": traffic.get("yvo_bytes")}
    dom, other = (yvo_info, eris_info) if yv_ms >= er_ms else (eris_info, yvo_info)
    roofline = {
        "bound": "tensor", "kernel": dom["kernel"], "achieved": dom["achieved"], "peak": FP64_PEAK_TFLOPS_MEASURED,
        "unit": "TFLOP/s", "frac": dom["frac"], "traffic": dom["traffic"],
        "peak_source": "cuBLAS DGEMM 32768x8192x1536 measured on this pool's B200 (FP64; MEASURED_PEAKS.json has "
                       "only bf16); nominal 37.2",
        "frac_of_nominal": dom["frac_of_nominal"], "ms_per_launch": dom["ms_per_launch"],
        "second_kernel": other,
        "stages_ms_per_step": {k: s[0] / K for k, s in stages.items()},
    }
    if "flops_counted" in dom:
        roofline["flops_counted"] = dom["flops_counted"]
        roofline["algorithmic_tflops"] = dom["algorithmic_tflops"]

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    line = {
        "metric": "dfmpcc_iteration_fp64_tflops", "value": tflops, "unit": "TFLOP/s",
        "n_gpus": world, "steps": K, "warmup": max(W, 3), "ms_per_step": ms_dev / K,
        "s_per_iteration": ms_dev / K * 1e-3,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"synthetic DF-MPCC nocc={o} nvir={v} naux={L} (BASELINE.json configs[2])",
                   "step": "updated_amps + run_diis (first K iterations of a solve from t1=0, DIIS on)",
                   "l2": "inputs_exceed_l2 (t2 buffer 8.6 GB, DIIS vectors 4.3 GB each at the default shape)",
                   "parallelism": f"a-slab x{world}", "slab_bounds": list(ll._bounds) if world > 1 else None,
                   "diis_vector": "column-owned" if ll._diis_columns else "row-owned", "jk_exchange": ll._jk_mode, "t2_peer_mirror": bool(ll._t2_peers),
                   "flop_per_iteration": flop_it,
                   "flop_note": "value counts the ALGORITHMIC flops of the reference iteration (SURVEY.md 8d: 4 L nov^2 + ...); "
                                "the eris GEMM executes about half of its share by symmetry, so value can exceed the FP64 peak"},
        "e2e": {"value": tflops_e2e, "unit": "TFLOP/s", "ms_per_step": ms_e2e / K,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "note": "integrals are uploaded once per solve at construction, as the reference builds them once"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roofline,
        "trace_dE_res": trace[:K],
    }
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(o, v, L, args.seed)
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
