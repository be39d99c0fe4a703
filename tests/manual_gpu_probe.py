"""Development probe (test infrastructure: it uses the oracle, so it lives under tests/; not collected by pytest): every intermediate of one updated_amps vs the oracle, then full kernel() traces."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from pyscf_mpcc_b200.synthetic import SyntheticMF
from pyscf_mpcc_b200.lowlevel import MPCC_LL
from oracle.mpcc_oracle import LowLevelOracle, amps_to_vector

def rel(a, b):
    return float(np.abs(a - b).max() / (np.abs(b).max() + 1e-300))

def check(o, v, L, scale, seed=0):
    mf = SyntheticMF(o, v, L, seed=seed, scale=scale)
    orc = LowLevelOracle(mf, mode="blas")
    ll = MPCC_LL(mf); ll.verbose_print = False
    rng = np.random.default_rng(5)
    t1 = 0.05 * rng.standard_normal((o, v))
    # oracle intermediates
    Loo, Lov, Lvv = orc.Loo, orc.Lov, orc.Lvv
    Xvo = np.einsum("Lab,ib->Lai", Lvv, t1, optimize=True)
    Xoo = np.einsum("Lia,ja->Lij", Lov, t1, optimize=True)
    X = np.einsum("Lia,ia->L", Lov, t1, optimize=True)
    Joo = Xoo + Loo
    Jvo = Xvo + Lov.transpose(0, 2, 1) - np.einsum("Lij,ja->Lai", Joo, t1, optimize=True)
    res_o, dE_o, t1n_o, t2_o = orc.updated_amps(t1)
    ll.t1 = t1
    res, dE, t1n, t2 = ll.updated_amps()
    print(f"--- ({o},{v},{L}) res {res:.12e} vs {res_o:.12e}  dE {dE:.12e} vs {dE_o:.12e}")
    B = ll.debug_buffer
    ldk = (L + 1) // 2 * 2
    print("  Lvo  ", rel(B("Lvo").reshape(L, v, o), Lov.transpose(0, 2, 1)))
    print("  Xov  ", rel(B("Xov").reshape(L, o, v), Xvo.transpose(0, 2, 1)))
    print("  Xoo  ", rel(B("Xoo").reshape(L, o, o), Xoo))
    print("  Joo  ", rel(B("Joo").reshape(L, o, o), Joo))
    print("  X    ", rel(B("X"), X))
    print("  JK   ", rel(B("JK").reshape(v * o, ldk)[:, :L], Jvo.transpose(1, 2, 0).reshape(v * o, L)))
    T = B("T").reshape(v, v, o, o)
    print("  T    ", rel(T, t2_o.transpose(0, 2, 1, 3)))
    Y_o = np.einsum("aibj,Ljb->Lai", t2_o, Lov, optimize=True)
    print("  Y    ", rel(B("Y").reshape(v * o, L), Y_o.transpose(1, 2, 0).reshape(v * o, L)))
    print("  t1n  ", rel(t1n, t1n_o), " t2 handle", rel(np.asarray(t2), t2_o))
    vec = ll.amplitudes_to_vector(t1n, t2).numpy()
    print("  vec  ", rel(vec, amps_to_vector(t1n_o, t2_o, o, v)))
    # full kernel
    ll2 = MPCC_LL(mf); ll2.verbose_print = False
    t0 = time.time(); E = ll2.kernel(); dt = time.time() - t0
    Eo = orc.kernel()
    print(f"  kernel: its {ll2.iterations} vs {orc.iterations}; E {E:.15e} vs {Eo:.15e}; dE {abs(E-Eo):.2e}; "
          f"t1 {np.abs(ll2.t1-orc.t1).max():.2e}; t2 {np.abs(np.asarray(ll2.t2)-orc.t2).max():.2e}; {dt:.2f}s")
    for (k, de, r), (k2, de2, r2) in zip(ll2.history, orc.history):
        print(f"     it {k}: dE {de:.12e} ({abs(de-de2):.1e})  res {r:.12e} ({abs(r-r2):.1e})")

if __name__ == "__main__":
    print(torch.cuda.get_device_name(0))
    for case in [(4, 10, 30, 0.3), (5, 19, 116, 0.1), (3, 7, 17, 0.2), (8, 40, 120, 0.1), (21, 30, 70, 0.1), (12, 64, 192, 0.1), (32, 72, 200, 0.05)]:
        try:
            check(*case)
        except Exception as e:
            import traceback; traceback.print_exc()
            print("FAILED", case)
            break
