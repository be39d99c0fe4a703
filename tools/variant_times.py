"""Stage times of one iteration under the t2 kernel variants (fused 128x64 / fused 128x128 / unfused), one GPU.
usage: python tools/variant_times.py [nocc nvir naux] [iterations]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from pyscf_mpcc_b200.diis import DIIS
from pyscf_mpcc_b200.synthetic import device_blocks, default_scale

o, v, L = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (64, 512, 1536)))
its = int(sys.argv[4]) if len(sys.argv) > 4 else 4
Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device="cuda")
blocks = (Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev]))
out = {}
for name, kw in (("fused_128x64", {}), ("fused_128x128", {"t2_wide_tiles": True}), ("unfused", {"t2_unfused": True}),
                 ("fused_128x64_no_sym", {"t2_sym": False})):
    ll = MPCC_LL(blocks=blocks, **kw)
    ad = DIIS()
    ad.reserve(ll)
    t1 = torch.zeros(o * v, dtype=torch.float64, device="cuda")
    for _ in range(2):
        _, _, t1 = ll.step_device(t1, ad)
    ll._ctx.call("mpccb_set_option", b"profile", 1)
    ll._ctx.profile_read()
    for _ in range(its):
        _, _, t1 = ll.step_device(t1, ad)
    torch.cuda.synchronize()
    st = ll._ctx.profile_read()
    out[name] = {k: round(s[0] / its, 3) for k, s in st.items()}
    print(name, out[name], flush=True)
    del ll, ad, t1
    torch.cuda.empty_cache()
print(json.dumps(out))
