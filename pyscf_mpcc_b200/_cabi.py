"""ctypes binding of ``libmpcc_b200.so`` (C ABI declared in ``include/mpcc_b200.h``).

The reference reaches its native helpers the same way — ``lib.load_library`` +
ctypes, raw pointers and ``c_int`` sizes, caller-owned buffers
(``pyscf/lib/misc.py:94-106``, ``pyscf/lib/numpy_helper.py:286-289``). There is
no CPU fallback: if the library is missing, importing the solver fails loudly.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libmpcc_b200.so")

c_double_p = ctypes.POINTER(ctypes.c_double)
c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_ll = ctypes.c_longlong

# name -> (restype, argtypes); every symbol include/mpcc_b200.h declares
SIGNATURES = {
    "mpccb_create": (c_int, [ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_int, c_int, c_int]),
    "mpccb_destroy": (None, [c_void_p]),
    "mpccb_last_error": (ctypes.c_char_p, [c_void_p]),
    "mpccb_version": (ctypes.c_char_p, []),
    "mpccb_set_stream": (c_int, [c_void_p, c_void_p]),
    "mpccb_synchronize": (c_int, [c_void_p]),
    "mpccb_workspace_bytes": (ctypes.c_size_t, [c_void_p]),
    "mpccb_bind_workspace": (c_int, [c_void_p, c_void_p, ctypes.c_size_t]),
    "mpccb_bind_integrals": (c_int, [c_void_p] + [c_void_p] * 7),
    "mpccb_ao2mo_scratch_bytes": (ctypes.c_size_t, [c_void_p, c_int]),
    "mpccb_ao2mo_block": (c_int, [c_void_p, c_void_p, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, ctypes.c_size_t]),
    "mpccb_bind_t2": (c_int, [c_void_p, c_void_p]),
    "mpccb_bind_jk_exchange": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int, c_void_p]),
    "mpccb_bind_jk_subset": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int, c_int, c_void_p]),
    "mpccb_dress": (c_int, [c_void_p, c_void_p]),
    "mpccb_dress_post": (c_int, [c_void_p]),
    "mpccb_dress_rest": (c_int, [c_void_p]),
    "mpccb_bind_xoo_exchange": (c_int, [c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int, c_int, c_void_p]),
    "mpccb_doubles": (c_int, [c_void_p, c_void_p]),
    "mpccb_t2_tiles": (c_int, [c_void_p, ctypes.POINTER(c_ll), ctypes.POINTER(c_ll)]),
    "mpccb_sym_query": (c_int, [c_int, c_int, ctypes.POINTER(c_int), c_int, c_int, c_int]),
    "mpccb_jk_subset_needs": (c_int, [c_int, c_int, c_int]),
    "mpccb_doubles_begin": (c_int, [c_void_p, c_void_p]),
    "mpccb_doubles_end": (c_int, [c_void_p, c_void_p]),
    "mpccb_bind_t2_peers": (c_int, [c_void_p, c_int, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_void_p)]),
    "mpccb_finish": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p]),
    "mpccb_updated_amps_host": (c_int, [c_void_p, c_void_p, c_void_p, c_double_p, c_double_p]),
    "mpccb_t2_to_aibj": (c_int, [c_void_p, c_void_p]),
    "mpccb_pack_tril_rows": (c_int, [c_void_p, c_void_p]),
    "mpccb_bind_diis_columns": (c_int, [c_void_p, c_void_p]),
    "mpccb_diis_scatter_tril": (c_int, [c_void_p, c_void_p, c_void_p]),
    "mpccb_unpack_oovv": (c_int, [c_void_p, c_void_p, c_void_p]),
    "mpccb_diis_push": (c_int, [c_void_p, c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_double_p, c_int,
                                c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_int, c_void_p]),
    "mpccb_vec_diis_push": (c_int, [c_void_p, c_void_p, c_ll, c_void_p, ctypes.POINTER(c_void_p), c_double_p, c_int,
                                    c_void_p, c_void_p, ctypes.POINTER(c_void_p), c_int, c_int, c_int, c_void_p]),
    "mpccb_pack_tril": (c_int, [c_void_p, c_void_p, c_ll, c_void_p]),
    "mpccb_unpack_tril": (c_int, [c_void_p, c_void_p, c_ll, c_void_p]),
    "mpccb_diis_combine": (c_int, [c_void_p, ctypes.POINTER(c_void_p), c_double_p, c_int, c_void_p, c_ll]),
    "mpccb_get_buffer": (c_int, [c_void_p, ctypes.c_char_p, ctypes.POINTER(c_void_p), ctypes.POINTER(c_ll)]),
    "mpccb_copy_to_host": (c_int, [c_void_p, c_void_p, c_void_p, c_ll]),
    "mpccb_copy_to_device": (c_int, [c_void_p, c_void_p, c_void_p, c_ll]),
    "mpccb_launch_count": (c_ll, [c_void_p]),
    "mpccb_profile_stages": (c_int, []),
    "mpccb_profile_stage_name": (ctypes.c_char_p, [c_int]),
    "mpccb_profile_read": (c_int, [c_void_p, c_double_p, ctypes.POINTER(c_ll)]),
    "mpccb_set_option": (c_int, [c_void_p, ctypes.c_char_p, c_ll]),
    "mpccb_permute_axpby": (c_int, [c_void_p, c_void_p, c_void_p, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int),
                                    ctypes.c_double, ctypes.c_double]),
    "mpccb_div_denominators": (c_int, [c_void_p, c_void_p, c_int]),
    "mpccb_div_general": (c_int, [c_void_p, c_void_p, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_void_p),
                                  c_double_p]),
    "mpccb_class_overwrite": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_void_p, c_void_p, ctypes.c_uint]),
    "mpccb_take_2d": (c_int, [c_void_p, c_void_p, c_ll, c_void_p, c_int, c_void_p, c_int, c_void_p]),
    "mpccb_takebak_2d": (c_int, [c_void_p, c_void_p, c_ll, c_void_p, c_void_p, c_int, c_void_p, c_int, ctypes.c_double]),
    "mpccb_dot": (c_int, [c_void_p, c_void_p, c_void_p, c_ll, c_void_p]),
    "mpccb_dot_static": (c_int, [c_void_p, c_void_p, c_void_p, c_ll, c_void_p]),
    "mpccb_test_fast_div": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_ll]),
    "mpccb_gemv": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_ll, c_ll, c_int, ctypes.c_double, ctypes.c_double]),
    "mpccb_gemm_nt_mapped": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_ll, c_ll,
                                     c_int, c_ll, c_ll, ctypes.c_double, ctypes.c_double]),
    "mpccb_gemm_nt": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int]),
}

_lib = None


def build(verbose: bool = False) -> str:
    """Compile the CUDA sources for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    out = subprocess.run(["make", "-C", CSRC], capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("building libmpcc_b200.so failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stdout)
    return LIB_PATH


def load():
    """Load the shared library and attach the prototypes. No fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C pyscf_mpcc_b200/csrc`). The DF-MPCC GPU path has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class MpccError(RuntimeError):
    pass


class Context:
    """Thin RAII wrapper of ``mpccb_ctx``; every call raises on a non-zero status."""

    def __init__(self, nocc, nvir, naux, a_begin, a_end, device_index):
        self.lib = load()
        h = c_void_p()
        rc = self.lib.mpccb_create(ctypes.byref(h), nocc, nvir, naux, a_begin, a_end, device_index)
        self.h = h
        if rc != 0:
            msg = self.lib.mpccb_last_error(h).decode() if h else "allocation failed"
            if h:
                self.lib.mpccb_destroy(h)
            self.h = None
            raise MpccError(f"mpccb_create: {msg}")

    def call(self, name, *args):
        rc = getattr(self.lib, name)(self.h, *args)
        if rc != 0:
            raise MpccError(f"{name}: {self.lib.mpccb_last_error(self.h).decode()}")

    def workspace_bytes(self) -> int:
        return int(self.lib.mpccb_workspace_bytes(self.h))

    def launch_count(self) -> int:
        return int(self.lib.mpccb_launch_count(self.h))

    def t2_tiles(self):
        """(launched, total) tiles of the fused eris -> t2 kernel."""
        a, b = c_ll(), c_ll()
        self.call("mpccb_t2_tiles", ctypes.byref(a), ctypes.byref(b))
        return int(a.value), int(b.value)

    def buffer(self, name: str):
        p = c_void_p()
        n = c_ll()
        self.call("mpccb_get_buffer", name.encode(), ctypes.byref(p), ctypes.byref(n))
        return p.value, int(n.value)

    def profile_read(self):
        """{stage: (total_ms, count)} since the last read; synchronises the stream."""
        n = self.lib.mpccb_profile_stages()
        ms = (ctypes.c_double * n)()
        cnt = (c_ll * n)()
        self.call("mpccb_profile_read", ms, cnt)
        return {self.lib.mpccb_profile_stage_name(i).decode(): (float(ms[i]), int(cnt[i])) for i in range(n)}

    def close(self):
        if getattr(self, "h", None):
            self.lib.mpccb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
