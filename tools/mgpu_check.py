"""Multi-rank parity check (run under torchrun, one rank per GPU, NCCL): the a-slab sharded solver must
reproduce the reference-generated fixtures of tests/golden — iteration counts, per-iteration (dE, res),
final t1 and the (all-gathered) final t2 — and, at the full target shape, the single-rank iteration.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
        tools/mgpu_check.py [--full] [--shape=o,v,L]
"""
import glob
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

E_TOL, AMP_TOL, STRIDE = 1e-8, 1e-10, 97


def main():
    rank, world, local = (int(os.environ[k]) for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import SyntheticMF
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    bad = 0
    fixtures = [] if "--only-shape" in sys.argv else sorted(glob.glob(os.path.join(root, "tests", "golden", "mpcc_*.npz")))
    for path in fixtures:
        g = np.load(path)
        o, v, L = (int(x) for x in g["shape"])
        if v < world:
            continue
        for slabs, unfused in ((None, None), (uneven_bounds(v, world), None), (None, True)):
            if slabs is not None and slabs == even_bounds(v, world):
                continue
            ll = MPCC_LL(SyntheticMF(o, v, L, seed=0, scale=float(g["scale"])), device=dev, slabs=slabs,
                         t2_unfused=unfused)
            ll.verbose_print = False
            e = ll.kernel()
            tr = np.array(ll.history)
            t2 = np.asarray(ll.t2)                                  # collective: all-gather + symmetric unpack
            ok = (ll.iterations == int(g["iterations"])
                  and np.abs(tr[:, 1] - g["trace_dE"]).max() <= E_TOL
                  and np.allclose(tr[:, 2], g["trace_res"], rtol=1e-6, atol=1e-12)
                  and np.abs(ll.t1 - g["final_t1"]).max() <= AMP_TOL
                  and t2.shape == (o, o, v, v)
                  and np.abs(t2.reshape(-1)[::STRIDE] - g["final_t2_samples"]).max() <= AMP_TOL
                  and abs(np.linalg.norm(t2) - g["final_t2_norm"]) <= 1e-9)
            # one residual evaluation: gathered (v,o,v,o) t2 and the flat DIIS vector
            ll.t1 = g["kat_rand_t1_in"]
            res, dE, t1n, t2h = ll.updated_amps()
            t2a = np.asarray(t2h)
            vec = ll.amplitudes_to_vector(t1n, t2h).numpy()
            ok = (ok and abs(dE - g["kat_rand_dE"]) <= E_TOL and np.abs(t1n - g["kat_rand_t1_new"]).max() <= AMP_TOL
                  and np.abs(t2a.reshape(-1)[::STRIDE] - g["kat_rand_t2_samples"]).max() <= AMP_TOL
                  and np.abs(vec[::STRIDE] - g["kat_rand_vec_samples"]).max() <= AMP_TOL
                  and abs(np.linalg.norm(vec) - g["kat_rand_vec_norm"]) <= 1e-9)
            flag = torch.tensor([0 if ok else 1], device=dev)
            dist.all_reduce(flag)
            if rank == 0:
                print(f"{os.path.basename(path)} world={world} slabs={ll._bounds} t2={'unfused' if unfused else 'auto'} peers={ll._t2_peers} jk={ll._jk_mode} its={ll.iterations} E={e:.12e} "
                      f"{'OK' if flag.item() == 0 else 'MISMATCH'}", flush=True)
            bad += int(flag.item())
            del ll
    if "--full" in sys.argv:
        bad += full_shape(rank, world, dev)
    for arg in sys.argv:
        # --shape=o,v,L : the same sharded-vs-single comparison at another shape, DIIS off (13 history vectors of a
        # large shape do not fit one GPU), two iterations
        if arg.startswith("--shape="):
            bad += full_shape(rank, world, dev, tuple(int(x) for x in arg.split("=")[1].split(",")), diis=False, its=2)
    dist.destroy_process_group()
    if bad:
        sys.exit(1)


def even_bounds(v, world):
    from pyscf_mpcc_b200 import sharding
    return [sharding.slab_bounds(v, world, r)[0] for r in range(world)] + [v]


def uneven_bounds(v, world):
    """Equal split with one index moved from each of the last ranks to the first ones (exercises the
    unequal-slab paths that sharding.balanced_bounds produces at large shapes)."""
    b = even_bounds(v, world)
    w = [hi - lo for lo, hi in zip(b[:-1], b[1:])]
    for r in range(world // 2):
        if w[-1 - r] > 1:
            w[r] += 1
            w[-1 - r] -= 1
    out = [0]
    for x in w:
        out.append(out[-1] + x)
    return out


def full_shape(rank, world, dev, shape=(64, 512, 1536), diis=True, its=4):
    """(64,512,1536) by default: sharded vs single-rank (group of one) on the same device inputs."""
    from pyscf_mpcc_b200.lowlevel import MPCC_LL
    from pyscf_mpcc_b200.synthetic import device_blocks, default_scale
    o, v, L = shape
    Loo, Lov, Lvv, eo, ev = device_blocks(o, v, L, seed=0, scale=default_scale(o, v), device=dev)
    blocks = (Loo, Lov, Lvv, torch.diag(eo), torch.diag(ev), torch.cat([eo, ev]))
    ll = MPCC_LL(blocks=blocks, device=dev)
    ll.verbose_print = False
    ll.max_its = its
    ll.diis = diis
    ll.kernel()
    hist_sharded = np.array(ll.history)
    t1_sharded = ll.t1.copy()
    del ll
    torch.cuda.empty_cache()
    # single-rank run on every GPU: a one-rank process group per rank
    groups = [dist.new_group([r]) for r in range(world)]
    ll1 = MPCC_LL(blocks=blocks, device=dev, group=groups[rank])
    ll1.verbose_print = False
    ll1.max_its = its
    ll1.diis = diis
    ll1.kernel()
    h1 = np.array(ll1.history)
    ok = (np.abs(h1[:, 1] - hist_sharded[:, 1]).max() <= E_TOL and np.allclose(h1[:, 2], hist_sharded[:, 2], rtol=1e-6)
          and np.abs(ll1.t1 - t1_sharded).max() <= AMP_TOL)
    flag = torch.tensor([0 if ok else 1], device=dev)
    dist.all_reduce(flag)
    if rank == 0:
        print(f"full shape {shape} world={world}: sharded vs single  dE diff {np.abs(h1[:, 1] - hist_sharded[:, 1]).max():.2e} "
              f"t1 diff {np.abs(ll1.t1 - t1_sharded).max():.2e} {'OK' if flag.item() == 0 else 'MISMATCH'}", flush=True)
    return int(flag.item())


if __name__ == "__main__":
    main()
