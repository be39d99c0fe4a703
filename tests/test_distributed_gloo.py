"""world_size-2 check (gloo, CPU) of the a-slab sharding arithmetic the multi-GPU path relies on:
slab-local doubles + all-reduce(partial Omega, dE, DIIS dots) reproduces the unsharded iteration.
The per-rank arithmetic is the oracle's (no GPU here); the partition/segment logic is the package's."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.mpcc_oracle import LowLevelOracle, amps_to_vector
    from pyscf_mpcc_b200 import sharding
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    o, v, L = 4, 11, 30
    mf = SyntheticMF(o, v, L, seed=0, scale=0.2)
    orc = LowLevelOracle(mf, mode="blas")
    t1 = 0.05 * np.random.default_rng(3).standard_normal((o, v))
    a0, a1 = sharding.slab_bounds(v, world, rank)
    # partial singles residual of this slab: total with slab minus the replicated part (empty slab)
    _, dE_s, t1n_s, t2_s = orc.updated_amps(t1, a_slab=(a0, a1))
    _, dE_0, t1n_0, _ = orc.updated_amps(t1, a_slab=(0, 0))
    part = torch.from_numpy(np.concatenate([(-(t1n_s - t1n_0) * orc.eia).T.reshape(-1), [dE_s - dE_0]]))
    dist.all_reduce(part)
    om_rep = (-t1n_0 * orc.eia).T
    om = om_rep + part[:-1].numpy().reshape(v, o)
    t1n = -om.T / orc.eia
    res_f, dE_f, t1n_f, t2_f = orc.updated_amps(t1)
    ok = np.abs(t1n - t1n_f).max() < 1e-12 and abs(dE_0 + part[-1].item() - dE_f) < 1e-12
    # sharded DIIS dot: slab rows of the packed vector are a contiguous segment
    vec = amps_to_vector(t1n_f, t2_f, o, v)
    nov = o * v
    s0, s1 = sharding.packed_segment(o, a0, a1)
    m_local = np.ascontiguousarray(t2_s.transpose(0, 2, 1, 3)).reshape((a1 - a0) * o, nov)
    rows = np.arange(a0 * o, a1 * o)
    seg = np.concatenate([m_local[k, :r + 1] for k, r in enumerate(rows)])
    ok = ok and np.allclose(seg, vec[nov + s0:nov + s1], rtol=0, atol=1e-14)
    ok = ok and len(seg) + nov == sharding.local_vector_length(o, v, a0, a1)
    d = torch.tensor([float(seg @ seg) + (float(vec[:nov] @ vec[:nov]) if rank == 0 else 0.0)], dtype=torch.float64)
    dist.all_reduce(d)
    ok = ok and abs(d.item() - float(vec @ vec)) < 1e-10 * float(vec @ vec)
    # column-owned DIIS vector (the multi-GPU default): elements whose second virtual index is in the slab, read
    # from the local [a,b,i,j] slab through t2[A,B,I,J] = t2[B,A,J,I]; same dot after the all-reduce
    loc, ref = sharding.column_owned_positions(o, v, a0, a1)
    segc = m_local.reshape(-1)[loc]
    ok = ok and np.allclose(segc, vec[nov + ref], rtol=0, atol=1e-14)
    ok = ok and len(segc) == int(sharding.column_owned_prefix(o, v, a0, a1)[-1])
    d = torch.tensor([float(segc @ segc) + (float(vec[:nov] @ vec[:nov]) if rank == 0 else 0.0), float(len(segc))],
                     dtype=torch.float64)
    dist.all_reduce(d)
    ok = ok and abs(d[0].item() - float(vec @ vec)) < 1e-10 * float(vec @ vec)
    ok = ok and int(d[1].item()) == nov * (nov + 1) // 2
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_two_rank_slab_sharding_reproduces_full_iteration():
    world = 2
    port = 29500 + (os.getpid() % 2000)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}
