"""Short driver for ncu: N iterations (updated_amps + DIIS) at the target shape, device-resident."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.diis import DIIS
from pyscf_mpcc_b200.synthetic import device_blocks, default_scale

o, v, L = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (64, 512, 1536)))
its = int(sys.argv[4]) if len(sys.argv) > 4 else 3
use_tma = (os.environ.get("MPCC_TMA", "1") != "0")
Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
ll = MPCC_LL(blocks=(Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev])), use_tma=use_tma)
t1 = torch.zeros(o * v, dtype=torch.float64, device="cuda")
ad = DIIS()
for k in range(its):
    res, dE, t1 = ll.step_device(t1, ad)
    print(k, res, dE)
torch.cuda.synchronize()
