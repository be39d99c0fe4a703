"""Small end-to-end runs for compute-sanitizer (memcheck): ragged shapes through every kernel variant, the device
AO->MO transform, the CCSD-type stage and the unrestricted residual.   compute-sanitizer --tool memcheck python tools/sanitize_small.py"""
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np

from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.synthetic import SyntheticMF
from pyscf_mpcc_b200 import rmpccsd, umpcc_lowlevel as U
from pyscf_mpcc_b200.tensor_engine import make_cuda_engine
from cc_problem import make_problem
from u_problem import make_u_problem

for (o, v, L, kw) in [(5, 19, 33, {}), (16, 21, 40, {}), (16, 21, 40, {"t2_wide_tiles": True}), (7, 9, 11, {"t2_unfused": True}),
                      (16, 18, 35, {"yvo_split": 3}), (3, 4, 5, {"use_tma": False}), (33, 35, 20, {"t2_sym": False}), (70, 72, 24, {})]:
    ll = MPCC_LL(SyntheticMF(o, v, L, seed=1, scale=0.1), **kw)
    ll.verbose_print = False
    ll.max_its = 3
    e = ll.kernel()
    t2 = np.asarray(ll.t2)
    print("lowlevel", o, v, L, kw, e, float(np.abs(t2).max()), flush=True)
p = make_problem(4, 9, 14, seed=3)
ll = MPCC_LL(blocks=(p["Loo"], p["Lov"], p["Lvv"], p["fock"][:4, :4].copy(), p["fock"][4:, 4:].copy(), np.concatenate([p["eo"], p["ev"]])))
eris = ll.high_level_eris(fock=p["fock"])
t1n, t2n = rmpccsd.update_amps(types.SimpleNamespace(cc2=False), p["t1"], p["t2"], eris, [0, 2], [1, 3, 4], [0], [1, 2, 9])
print("rmpccsd", float(np.abs(t2n).max()), flush=True)
pu = make_u_problem(3, 2, 9, 12, seed=2)
eng, keep = make_cuda_engine(3, 7, 12)
st = U.UMPCCState(eng, pu["a"], pu["b"])
st.set_t1(pu["a"]["t1"], pu["b"]["t1"])
print("umpcc", U.updated_amp(st)[:2], flush=True)
