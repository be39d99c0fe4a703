"""Throughput of the unrestricted low-level residual (row f4): one umpcc_lowlevel.updated_amp on synthetic DF
tensors, CUDA-event timed; FP64 TFLOP/s from the GEMM flops the engine issued.   python tools/u_bench.py [oa,ob,nmo,L ...]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch

from pyscf_mpcc_b200 import umpcc_lowlevel as U
from pyscf_mpcc_b200.tensor_engine import make_cuda_engine
from u_problem import make_u_problem


def run(oa, ob, n, L):
    p = make_u_problem(oa, ob, n, L, seed=1, scale=0.05)
    eng, keep = make_cuda_engine(max(oa, ob), n - min(oa, ob), L)
    st = U.UMPCCState(eng, p["a"], p["b"])
    st.set_t1(p["a"]["t1"] * 0.1, p["b"]["t1"] * 0.1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    row = {}
    for rep in range(3):
        eng.flops = 0.0; eng.gemms = 0
        e0.record()
        res, dE, t1, t2 = U.updated_amp(st)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        row = {"nocc_a": oa, "nocc_b": ob, "nmo": n, "naux": L, "ms_updated_amp": ms, "gemm_flops": eng.flops,
               "gemms": eng.gemms, "tflops": eng.flops / (ms * 1e-3) * 1e-12, "res": res, "dE": dE}
    return row


if __name__ == "__main__":
    shapes = [tuple(int(x) for x in s.split(",")) for s in sys.argv[1:]] or [(16, 15, 144, 384), (32, 31, 288, 768), (48, 47, 432, 1152)]
    rows = []
    for s in shapes:
        rows.append(run(*s))
        print(json.dumps(rows[-1]), flush=True)
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/u_bench.json", "w"), indent=1)
