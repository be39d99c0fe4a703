"""GPU timeline of one iteration (kernels, copies, gaps) per rank, via torch.profiler (CUPTI) — the image has no nsys.
Run under torchrun; writes gpurun_out/timeline_rank{r}.txt. Diagnostic only: numbers under the profiler are not bench values.
usage: torchrun ... tools/timeline.py [nocc nvir naux]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from torch.profiler import profile, ProfilerActivity
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.diis import DIIS
from pyscf_mpcc_b200.synthetic import device_blocks, default_scale

o, v, L = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (64, 512, 1536)))
rank = int(os.environ.get("RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
group = None
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    group = dist.group.WORLD
Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])), group=group)
ad = DIIS(); ad.reserve(ll)
t1 = torch.zeros(o * v, dtype=torch.float64, device="cuda")
for _ in range(4):
    _, _, t1 = ll.step_device(t1, ad)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        _, _, t1 = ll.step_device(t1, ad)
    torch.cuda.synchronize()
ev_gpu = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev_gpu.sort(key=lambda e: e.time_range.start)
os.makedirs("gpurun_out", exist_ok=True)
with open(f"gpurun_out/timeline_rank{rank}.txt", "w") as f:
    t0 = ev_gpu[0].time_range.start if ev_gpu else 0
    prev_end = t0
    for e in ev_gpu:
        s, en = e.time_range.start, e.time_range.end
        f.write(f"{(s - t0) / 1e3:10.3f} ms  dur {(en - s):9.1f} us  gap {(s - prev_end):8.1f} us  {e.name[:110]}\n")
        prev_end = max(prev_end, en)
    f.write(f"total span {(prev_end - t0) / 1e3:.3f} ms over 3 iterations\n")
if world > 1:
    dist.destroy_process_group()
