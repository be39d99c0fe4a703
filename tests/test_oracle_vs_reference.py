"""Pins the oracle to the reference's own code, run unmodified through oracle/ref_shim.py.
Only runs where /root/reference exists (the build container); the committed golden
fixtures carry the same pin to the GPU box."""
import contextlib
import io

import numpy as np
import pytest

from oracle import ref_shim
from oracle.mpcc_oracle import LowLevelOracle, PulayDIIS, amps_to_vector, vector_to_amps
from pyscf_mpcc_b200.synthetic import SyntheticMF

pytestmark = pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree not mounted")


@pytest.mark.parametrize("shape", [(4, 10, 30, 0.3), (3, 7, 17, 0.2), (6, 11, 40, 0.15)])
def test_kernel_bitwise_in_einsum_mode(shape):
    o, v, L, s = shape
    ll, _ = ref_shim.load_reference()
    mf = SyntheticMF(o, v, L, seed=0, scale=s)
    with contextlib.redirect_stdout(io.StringIO()):
        ref = ll.MPCC_LL(mf)
        ref.con_tol, ref.max_its = 1e-6, 50
        e_ref = ref.kernel()
    orc = LowLevelOracle(mf, mode="einsum")
    e = orc.kernel()
    assert e == e_ref
    np.testing.assert_array_equal(orc.t1, ref.t1)
    np.testing.assert_array_equal(orc.t2, ref.t2)
    blas = LowLevelOracle(mf, mode="blas")
    assert abs(blas.kernel() - e_ref) < 1e-13 and blas.iterations == orc.iterations
    np.testing.assert_allclose(blas.t2, ref.t2, atol=1e-13)


def test_setup_tensors_match():
    ll, _ = ref_shim.load_reference()
    mf = SyntheticMF(5, 9, 23, seed=2, scale=0.2)
    mf.mo_coeff = np.linalg.qr(np.random.default_rng(0).standard_normal((14, 14)))[0]   # non-trivial AO->MO
    ref = ll.MPCC_LL(mf)
    orc = LowLevelOracle(mf)
    for name in ("Loo", "Lov", "Lvv", "foo", "fvv", "eia"):
        np.testing.assert_allclose(getattr(orc, name), getattr(ref, name), atol=1e-13, err_msg=name)
    np.testing.assert_allclose(orc.denominators(), ref.D, atol=1e-14)


def test_vector_layout_and_diis_match():
    ll, diis = ref_shim.load_reference()
    mf = SyntheticMF(4, 6, 15, seed=1, scale=0.2)
    ref = ll.MPCC_LL(mf)
    rng = np.random.default_rng(0)
    t1 = rng.standard_normal((4, 6)); t2 = rng.standard_normal((6, 4, 6, 4))
    v_ref = ref.amplitudes_to_vector(t1, t2)
    np.testing.assert_array_equal(amps_to_vector(t1, t2, 4, 6), v_ref)
    a, b = vector_to_amps(v_ref, 4, 6)
    a2, b2 = ref.vector_to_amplitudes(v_ref)
    np.testing.assert_array_equal(a, a2); np.testing.assert_array_equal(b, b2)
    d_ref, d = diis.DIIS(), PulayDIIS()
    base = rng.standard_normal(50)
    for k in range(9):
        x = base + 0.3 ** k * rng.standard_normal(50)
        np.testing.assert_allclose(d.update(x.copy()), d_ref.update(x.copy()), rtol=1e-12, atol=1e-13)


@pytest.mark.skipif(not ref_shim.full_tree_available(), reason="pyscf/cc/ccsd.py not present")
def test_s4_vector_helpers_match_the_text_of_ccsd_py():
    """oracle/cc_oracle.py restates ccsd._unpack_4fold / amplitudes_to_vector_s4 / vector_to_amplitudes_s4
    (pyscf/cc/ccsd.py:705-757); ccsd.py cannot be imported (libcc), so the SOURCE TEXT of the three functions is taken
    from the reference file at test time (nothing is copied into the repository) and executed with NumPy forms of
    lib.take_2d / lib.takebak_2d (numpy_helper.py:478-541: a[idx[:,None],idy] and out[idx[:,None],idy] += a)."""
    import ast
    import os
    import types
    from oracle import cc_oracle
    path = os.path.join(ref_shim.REFERENCE_ROOT, "pyscf", "cc", "ccsd.py")
    src = open(path).read()
    wanted = {"_unpack_4fold", "amplitudes_to_vector_s4", "vector_to_amplitudes_s4"}
    tree = ast.parse(src)
    funcs = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in wanted]
    assert {f.name for f in funcs} == wanted
    ns = {"numpy": np, "lib": types.SimpleNamespace(
        take_2d=lambda a, idx, idy, out=None: cc_oracle.take_2d(a, idx, idy, out=out),
        takebak_2d=lambda out, a, idx, idy, thread_safe=True: cc_oracle.takebak_2d(out, a, idx, idy))}
    exec(compile(ast.Module(body=funcs, type_ignores=[]), path, "exec"), ns)
    rng = np.random.default_rng(0)
    for nocc, nvir in ((1, 3), (2, 2), (3, 5), (4, 4), (5, 2), (3, 1)):
        t1 = rng.standard_normal((nocc, nvir))
        t2 = rng.standard_normal((nvir, nocc, nvir, nocc))       # the prototype hands in [a,i,b,j] blocks
        v_ref = ns["amplitudes_to_vector_s4"](t1, t2)
        v = cc_oracle.amplitudes_to_vector_s4(t1, t2)
        np.testing.assert_array_equal(v, v_ref)
        a1, a2 = ns["vector_to_amplitudes_s4"](v_ref, nocc + nvir, nocc)
        b1, b2 = cc_oracle.vector_to_amplitudes_s4(v, nocc + nvir, nocc)
        np.testing.assert_array_equal(a1, b1)
        np.testing.assert_array_equal(a2, b2)
        assert b2.shape == (nocc, nocc, nvir, nvir)
        np.testing.assert_array_equal(b2, -b2.transpose(1, 0, 2, 3))
