"""Residual throughput sweep (BASELINE.json configs[1], [3], [4]): one `updated_amps` (device-resident,
no DIIS) per shape, CUDA-event timed, FP64 TFLOP/s from the SURVEY.md §8d flop formula.

    python tools/sweep.py [--out gpurun_out/sweep.json] [--shapes o,v,L ...]

One GPU, or several under torchrun (a-slab sharding; `ms` is the max over ranks); shapes whose t2 slab + integrals
do not fit the free memory are skipped.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/sweep.py --shapes 128,1024,3072
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from pyscf_mpcc_b200 import flops as F
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.synthetic import device_blocks, default_scale

NOMINAL = 37.2      # 148 SMs x 64 DFMA/clk x 1.965 GHz

DEFAULT = [(21, 243, 654), (32, 256, 768), (32, 512, 1536), (32, 1024, 3072), (64, 256, 768), (64, 512, 1536),
           (64, 1024, 3072), (80, 848, 2224), (128, 256, 768), (128, 512, 1536)]


WORLD = int(os.environ.get("WORLD_SIZE", 1))
RANK = int(os.environ.get("RANK", 0))


def need_bytes(o, v, L):
    nov = o * v
    return 8 * (nov * nov // WORLD + 2 * L * v * v + 9 * L * nov + 600 * nov) + (2 << 30)


def run(o, v, L, steps, warmup):
    Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
    group = torch.distributed.group.WORLD if WORLD > 1 else None
    ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])), group=group)
    ll._ctx.call("mpccb_set_option", b"profile", 1)
    t1 = torch.zeros(o * v, dtype=torch.float64, device="cuda")
    for _ in range(warmup):
        res, dE, t1 = ll.step_device(t1, None)
    ll._ctx.profile_read()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    if WORLD > 1:
        torch.distributed.barrier()
    e0.record()
    for _ in range(steps):
        res, dE, t1 = ll.step_device(t1, None)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if WORLD > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t.item())
    st = ll._ctx.profile_read()
    fl = F.updated_amps_flops(o, v, L)
    er = F.eris_flops(o, v, L, ll._na) / (st["eris_t2"][0] / steps * 1e-3) * 1e-12      # this rank's slab
    yv = F.yvo_flops(o, v, L, ll._na) / (st["yvo"][0] / steps * 1e-3) * 1e-12
    # EXECUTED flops: the mirror scheme computes `launched` of `total` eris tiles (about half); a roofline fraction
    # divides executed flops by the peak, the algorithmic rate is reported separately and may exceed the peak
    launched, total = ll._ctx.t2_tiles()
    skipped = F.eris_flops(o, v, L, ll._na) * (1.0 - launched / max(total, 1)) * WORLD
    fl_exec = fl - skipped
    return {"nocc": o, "nvir": v, "naux": L, "n_gpus": WORLD, "ms_per_updated_amps": ms,
            "tflops_algorithmic": fl / (ms * 1e-3) * 1e-12, "tflops_executed": fl_exec / (ms * 1e-3) * 1e-12,
            "frac_of_nominal_executed": fl_exec / (ms * 1e-3) * 1e-12 / (NOMINAL * WORLD),
            "eris_t2_tflops_algorithmic": er, "eris_t2_tflops_executed": er * launched / max(total, 1), "yvo_tflops": yv,
            "stages_ms": {k: s[0] / steps for k, s in st.items() if s[1]}, "res": res, "dE": dE}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--shapes", nargs="*")
    a = ap.parse_args()
    shapes = [tuple(int(x) for x in s.split(",")) for s in a.shapes] if a.shapes else DEFAULT
    if WORLD > 1:
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        torch.distributed.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    free = torch.cuda.mem_get_info()[0]
    rows = []
    for o, v, L in shapes:
        if need_bytes(o, v, L) > free:
            rows.append({"nocc": o, "nvir": v, "naux": L, "skipped": f"needs ~{need_bytes(o, v, L) / 2**30:.0f} GiB"})
            print(rows[-1], flush=True)
            continue
        try:
            rows.append(run(o, v, L, a.steps, a.warmup))
        except Exception as e:                                   # keep the sweep going; record the failure
            rows.append({"nocc": o, "nvir": v, "naux": L, "error": repr(e)[:300]})
        if RANK == 0:
            print(json.dumps(rows[-1]), flush=True)
        torch.cuda.empty_cache()
    if RANK == 0:
        with open(a.out, "w") as f:
            json.dump(rows, f, indent=1)
    if WORLD > 1:
        torch.distributed.destroy_process_group()
