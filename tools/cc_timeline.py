"""Kernel-time breakdown of one steady-state rmpccsd.update_amps (torch.profiler / CUPTI): totals per kernel name,
busy and idle time. Diagnostic only.

    python tools/cc_timeline.py [o,v,L]
"""
import json
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

from pyscf_mpcc_b200 import rmpccsd
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.synthetic import default_scale, device_blocks

o, v, L = (int(x) for x in sys.argv[1].split(",")) if len(sys.argv) > 1 else (32, 192, 576)
Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])))
eris = ll.high_level_eris()
cc = types.SimpleNamespace(cc2=False)
emp2, t1, t2 = rmpccsd.init_amps(cc, eris)
args = (list(range(o // 2, o)), list(range(v // 4)), [0, 1, 2], list(range(15)))
for _ in range(2):
    rmpccsd.update_amps(cc, t1, t2, eris, *args)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    rmpccsd.update_amps(cc, t1, t2, eris, *args)
    torch.cuda.synchronize()
ks = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA], key=lambda e: e.time_range.start)
tot, busy, hi = {}, 0.0, None
for k in ks:
    a, b = k.time_range.start, k.time_range.end
    t = tot.setdefault(k.name[:110], [0, 0.0])
    t[0] += 1
    t[1] += b - a
    if hi is None or a > hi:
        busy += b - a
        hi = b
    elif b > hi:
        busy += b - hi
        hi = b
span = ks[-1].time_range.end - ks[0].time_range.start
out = {"shape": [o, v, L], "span_us": span, "busy_us": busy, "idle_us": span - busy, "launches": len(ks),
       "totals_us": {n: [c, round(d, 1)] for n, (c, d) in sorted(tot.items(), key=lambda x: -x[1][1])}}
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open("gpurun_out/cc_timeline.json", "w"), indent=1)
print(json.dumps({k: out[k] for k in ("span_us", "busy_us", "idle_us", "launches")}))
for n, (c, d) in list(out["totals_us"].items())[:16]:
    print(f"{d / 1e3:9.2f} ms  n={c:4d}  {n[:100]}")
