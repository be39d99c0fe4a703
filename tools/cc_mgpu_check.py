"""Several ranks (torchrun, NCCL): the CCSD-type residual with the particle-particle ladder and the ring / exchange
terms sharded over the ranks (slabs of the first virtual index, one all-reduce of the [a,b,i,j] accumulator) against
the fixtures the reference's own update_amps produced, then a timing of one update at a larger shape.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/cc_mgpu_check.py

The integrals of the high-level stage are replicated: every process owns a one-rank low-level solver (its own
singleton process group), only ``cc.group`` (the WORLD group) shards the ladder.
"""
import glob
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import torch
import torch.distributed as dist

from cc_problem import make_problem
from pyscf_mpcc_b200 import rmpccsd
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.synthetic import device_blocks, default_scale

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
singletons = [dist.new_group([r]) for r in range(world)]      # collective: every rank creates every group
mine = singletons[rank]

ok_all = True
paths = sorted(glob.glob(os.path.join(ROOT, "tests", "golden", "rmpccsd_[0-9]*.npz"))
               + glob.glob(os.path.join(ROOT, "tests", "golden", "rmpccsd_cc2_*.npz")))
for path in paths:
    g = np.load(path)
    o, v, L = (int(x) for x in g["shape"])
    p = make_problem(o, v, L, seed=int(g["seed"]))
    ll = MPCC_LL(blocks=(p["Loo"], p["Lov"], p["Lvv"], p["fock"][:o, :o].copy(), p["fock"][o:, o:].copy(),
                         np.concatenate([p["eo"], p["ev"]])), device=dev, group=mine)
    eris = ll.high_level_eris(fock=p["fock"])
    cc = types.SimpleNamespace(cc2="cc2" in os.path.basename(path), group=dist.group.WORLD)
    t1n, t2n = rmpccsd.update_amps(cc, p["t1"], p["t2"], eris, list(g["act_hole"]), list(g["act_particle"]),
                                   list(g["idx_singles"]), list(g["idx_doubles"]))
    scale = max(1.0, float(np.abs(g["t2new_samples"]).max()))
    ok = (np.abs(t1n - g["t1new"]).max() <= 1e-10 * max(1.0, float(np.abs(g["t1new"]).max()))
          and np.abs(t2n.reshape(-1)[::int(g["stride"])] - g["t2new_samples"]).max() <= 1e-10 * scale)
    flag = torch.tensor([int(ok)], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok_all &= bool(flag.item())
    if rank == 0:
        print(("OK      " if flag.item() else "MISMATCH"), os.path.basename(path), f"{world} ranks", flush=True)
    del ll, eris

o, v, L = 32, 192, 576
Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device=dev)
ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])), device=dev, group=mine)
eris = ll.high_level_eris()
res = {}
for label, group in (("replicated", None), ("sharded", dist.group.WORLD)):
    cc = types.SimpleNamespace(cc2=False, group=group)
    emp2, t1, t2 = rmpccsd.init_amps(cc, eris)
    for rep in range(3):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        t1n, t2n = rmpccsd.update_amps(cc, t1, t2, eris, list(range(o // 2, o)), list(range(v // 4)), [0, 1, 2],
                                       list(range(15)))
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    res[label] = {"ms": float(ms.item()), "e_corr_after": rmpccsd.energy(cc, t1n, t2n, eris)}
if rank == 0:
    res.update(n_gpus=world, shape=[o, v, L], parity_ok=ok_all)
    print(json.dumps(res), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(res, open(f"gpurun_out/cc_mgpu_n{world}.json", "w"), indent=1)
dist.destroy_process_group()
