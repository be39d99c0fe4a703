"""Per-kernel timeline of a few iterations (CUPTI through torch.profiler; every kernel of the process is seen, the
library's own included): where a step's time goes BETWEEN the kernels — launch gaps, barriers, collectives, host
solves — which the per-stage CUDA events of bench.py do not show.

    python tools/timeline.py [--steps 3]                                   (one GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/timeline.py

Rank 0 writes gpurun_out/timeline_n<N>.json: the kernels of the LAST profiled step in start order
(name, stream, start_us relative to the first kernel of the step, dur_us, gap_us to the end of the previous kernel on
any stream), per-name totals, busy time (union over streams) and idle time of the step.
A profiler run is diagnostic only: no bench number comes from it.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from torch.profiler import ProfilerActivity, profile

from pyscf_mpcc_b200.diis import DIIS
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.synthetic import default_scale, host_blocks

ap = argparse.ArgumentParser()
ap.add_argument("--nocc", type=int, default=64)
ap.add_argument("--nvir", type=int, default=512)
ap.add_argument("--naux", type=int, default=1536)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--warm", type=int, default=8)
args = ap.parse_args()

world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
o, v, L = args.nocc, args.nvir, args.naux
Loo, Lov, Lvv, eo, ev = (torch.from_numpy(x).to(dev) for x in host_blocks(o, v, L, seed=0, scale=default_scale(o, v)))
ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])), device=dev)
ll.verbose_print = False
t1 = torch.zeros(o * v, dtype=torch.float64, device=dev)
adiis = DIIS().reserve(ll)
for _ in range(args.warm):                       # DIIS space full (6 vectors) before anything is recorded
    _, _, t1 = ll.step_device(t1, adiis)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
marks = []
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for s in range(args.steps):
        with torch.profiler.record_function(f"mpcc_step_{s}"):
            _, _, t1 = ll.step_device(t1, adiis)
    torch.cuda.synchronize()
if world > 1:
    dist.barrier()

if rank == 0:
    evs = prof.events()
    cuda_t = torch.autograd.DeviceType.CUDA
    kern = sorted([e for e in evs if e.device_type == cuda_t and not e.name.startswith("mpcc_step_")],
                  key=lambda e: e.time_range.start)
    marks = sorted([(e.time_range.start, e.time_range.end, e.name) for e in evs
                    if e.name.startswith("mpcc_step_") and e.device_type != cuda_t])
    t0 = kern[0].time_range.start
    raw = [[k.name[:100], round(k.time_range.start - t0, 2), round(k.time_range.end - k.time_range.start, 2)]
           for k in kern]
    # steps are cut at the first kernel of each T1 dressing (the t1 transpose is the first launch of mpccb_dress)
    starts = [i for i, r in enumerate(raw) if "k_transpose_batched" in r[0]]
    lo = starts[-2] if len(starts) >= 2 else 0
    hi = starts[-1] if len(starts) >= 2 else len(raw)
    ks = raw[lo:hi]                                   # the last complete step
    rows, totals = [], {}
    s0 = ks[0][1]
    end_prev, busy, cur_lo, cur_hi = s0, 0.0, None, None
    for name, a, d in ks:
        b = a + d
        rows.append({"name": name, "start_us": round(a - s0, 2), "dur_us": d, "gap_us": round(a - end_prev, 2)})
        end_prev = max(end_prev, b)
        t = totals.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += d
        if cur_lo is None:
            cur_lo, cur_hi = a, b
        elif a <= cur_hi:
            cur_hi = max(cur_hi, b)
        else:
            busy += cur_hi - cur_lo
            cur_lo, cur_hi = a, b
    busy += cur_hi - cur_lo
    span = raw[hi][1] - s0 if hi < len(raw) else end_prev - s0       # start of this dressing to start of the next
    out = {"n_gpus": world, "shape": [o, v, L], "step_span_us": round(span, 1), "busy_us": round(busy, 1),
           "idle_us": round(span - busy, 1), "cpu_steps_us": [[round(a - t0, 1), round(b - t0, 1), n] for a, b, n in marks],
           "totals_us": {n: [c, round(d, 1)] for n, (c, d) in sorted(totals.items(), key=lambda x: -x[1][1])},
           "kernels": rows, "all_kernels": raw}
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open(f"gpurun_out/timeline_n{world}.json", "w"), indent=1)
    print(json.dumps({k: out[k] for k in ("n_gpus", "step_span_us", "busy_us", "idle_us")}))
    for r in rows:
        if r["gap_us"] > 15:
            print(f"gap {r['gap_us']:9.1f} us before {r['name'][:70]} at {r['start_us']:.0f}")
if world > 1:
    dist.destroy_process_group()
