"""Allocation-free entry point ``mpccb_dot_static`` (include/mpcc_b200.h): the dot product of ``mpccb_dot`` with its
per-block partial sums in a file-scope device array of the library instead of the caller-allocated workspace.

Collected last on purpose (file name): a failure here must not hide other tests.
Passes alone in a fresh process and after the rest of the suite: it creates its own contexts and touches no shared
state but the library's static scratch.

Reference for the values: the existing entry point ``mpccb_dot`` (same kernels, same block count -> identical bits;
the tolerance written below is the one of the other GPU tests, 1e-10 relative) and NumPy.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL_TOL = 1e-10


def _contexts():
    """(ctx with a bound workspace for mpccb_dot, ctx WITHOUT any workspace for mpccb_dot_static, keep-alive)."""
    import torch
    from pyscf_mpcc_b200 import _cabi
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    ref = _cabi.Context(4, 8, 16, 0, 8, 0)
    ws = torch.empty(ref.workspace_bytes() + 256, dtype=torch.uint8, device=dev)
    ref.call("mpccb_bind_workspace", (ws.data_ptr() + 255) // 256 * 256, ref.workspace_bytes())
    bare = _cabi.Context(4, 8, 16, 0, 8, 0)
    stream = torch.cuda.current_stream(dev).cuda_stream
    ref.call("mpccb_set_stream", stream)
    bare.call("mpccb_set_stream", stream)
    return ref, bare, ws


@pytest.mark.parametrize("n", [1, 255, 1024, 1000003, 592 * 1024 + 17, 20_000_000])
def test_dot_static_matches_dot_on_two_calls_in_a_row(n):
    import torch
    dev = torch.device("cuda", 0)
    ref, bare, ws = _contexts()
    rng = np.random.default_rng(n)
    # everything the calls need is allocated by the CALLER, before them
    xs = [torch.from_numpy(rng.standard_normal(n)).to(dev) for _ in range(2)]
    ys = [torch.from_numpy(rng.standard_normal(n)).to(dev) for _ in range(2)]
    out_ref = torch.zeros(2, dtype=torch.float64, device=dev)
    out_new = torch.full((2,), np.nan, dtype=torch.float64, device=dev)
    for k in range(2):                                  # second call finds the first call's partials in the scratch
        ref.call("mpccb_dot", xs[k].data_ptr(), ys[k].data_ptr(), n, out_ref[k:].data_ptr())
        bare.call("mpccb_dot_static", xs[k].data_ptr(), ys[k].data_ptr(), n, out_new[k:].data_ptr())
    torch.cuda.synchronize()
    got, want = out_new.cpu().numpy(), out_ref.cpu().numpy()
    for k in range(2):
        exact = float(np.dot(xs[k].cpu().numpy(), ys[k].cpu().numpy()))
        scale = max(1.0, float(np.sqrt(n)))
        assert abs(got[k] - want[k]) <= REL_TOL * scale, (k, got[k], want[k])
        assert got[k] == want[k]                        # same kernels, same block count: the same bits
        assert abs(got[k] - exact) <= REL_TOL * scale
    assert got[0] != got[1]                             # the two calls really were different problems


def test_dot_static_allocates_nothing_and_rejects_oversize():
    import torch
    from pyscf_mpcc_b200 import _cabi
    dev = torch.device("cuda", 0)
    ref, bare, ws = _contexts()
    n = 3_000_000
    x = torch.arange(n, dtype=torch.float64, device=dev) / n
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    bare.call("mpccb_dot_static", x.data_ptr(), x.data_ptr(), n, out.data_ptr())     # loads the module, looks up the symbol
    torch.cuda.synchronize()
    first = float(out.item())
    free0 = torch.cuda.mem_get_info(dev)[0]
    for _ in range(5):
        bare.call("mpccb_dot_static", x.data_ptr(), x.data_ptr(), n, out.data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info(dev)[0] == free0     # device memory in use did not move
    assert float(out.item()) == first
    assert abs(first - float(np.dot(np.arange(n) / n, np.arange(n) / n))) <= REL_TOL * n
    # sizes beyond the stated maximum (2^32 elements) and non-positive sizes: a clear error, nothing is launched
    launches = bare.launch_count()
    for bad in ((1 << 32) + 1, 0, -5):
        with pytest.raises(_cabi.MpccError, match="mpccb_dot_static"):
            bare.call("mpccb_dot_static", x.data_ptr(), x.data_ptr(), bad, out.data_ptr())
    with pytest.raises(_cabi.MpccError, match="null"):
        bare.call("mpccb_dot_static", None, x.data_ptr(), n, out.data_ptr())
    assert bare.launch_count() == launches
    # the existing entry point still wants its workspace: that requirement is what the static form removes
    with pytest.raises(_cabi.MpccError, match="workspace"):
        bare.call("mpccb_dot", x.data_ptr(), x.data_ptr(), n, out.data_ptr())
