"""B200 low-level DF-MPCC solver — drop-in for the reference's plug-in module.

The reference solver object takes a *module* and instantiates
``lowlevel.MPCC_LL(mf)`` (``pyscf/mpcc/mpcc.py:5,13``; ``pyscf/mpcc/dfrmpcc.py:6``
passes ``dfrmpcc_lowlevel``). This module exports ``MPCC_LL`` with the same
constructor, attributes and methods (``pyscf/mpcc/dfrmpcc_lowlevel.py:6-176``):

    kernel() -> dE                           (:72-100)
    updated_amps() -> (res, dE, t1, t2)      (:103-152)   aliases: update_amps, energy
    run_diis(t1, t2, adiis) -> (t1, t2)      (:155-160)
    amplitudes_to_vector / vector_to_amplitudes (:162-176)
    compute_three_center_ints()              (:46-69)

Everything numerical runs in hand-written sm_100a kernels behind the C ABI of
``libmpcc_b200.so`` (``include/mpcc_b200.h``); PyTorch only owns the device
buffers and the NCCL process group. There is no CPU fallback: constructing the
solver without a CUDA device or without the built library raises.

Bug-compatible details kept on purpose: the ``+`` sign in Jvv,
the dressed ``Jvo`` in ``Fov``, t2 returned as (v,o,v,o) from ``updated_amps``
and as (o,o,v,v) after DIIS, the non-symmetric lower-triangle DIIS vector, the
energy lagging the amplitudes by one update, convergence tested on T1 only.
"""
from __future__ import annotations

import ctypes
import os
import sys

import numpy as np
import torch

from . import _cabi, sharding
from .diis import DIIS, AmplitudeVector


class DeviceT2:
    """Lazy host view of doubles amplitudes that live on the GPU.

    ``layout == "aibj"``: the (v,o,v,o) tensor ``updated_amps`` returns (:137,:152),
    backed by the solver's [a,b,i,j] buffer. ``layout == "oovv"``: the (o,o,v,v)
    tensor ``vector_to_amplitudes`` returns (:170-176), backed by a packed vector.
    ``numpy()`` / ``np.asarray`` materialise it; nothing on the hot path does.
    """

    def __init__(self, solver, layout, packed=None, serial=0, packed_is_full=False, owner=None, source=None):
        self.solver = solver
        self.layout = layout
        self.packed = packed
        self.source = source        # "oovv": the AmplitudeVector to take the packed t2 part from, on first use
        self.serial = serial
        # "oovv": `packed` is either this rank's local part of a DIIS vector (row segment or column-owned order) or,
        # when it came from a host vector, the whole reference-order packed triangle (no gather needed).
        self.packed_is_full = packed_is_full
        # the DIIS object whose extrapolation buffer `packed` aliases (None: an own buffer, never stale)
        self.owner = owner
        o, v = solver.nocc, solver.nvir
        self.shape = (v, o, v, o) if layout == "aibj" else (o, o, v, v)
        self.dtype = np.dtype(np.float64)
        self.ndim = 4

    def numpy(self):
        """Full host tensor. With several ranks this is a collective call: every rank
        contributes its slab (all-gather over the solver's process group)."""
        ll = self.solver
        if ll.world_size == 1:
            return self.local_numpy()
        o, v = ll.nocc, ll.nvir
        if self.layout == "aibj":
            slab = self._local_device()                      # (na, o, v, o)
            return ll._gather_rows(slab.reshape(ll._na, -1), v).reshape(v, o, v, o).cpu().numpy()
        self._check_fresh()
        seg = self._packed()                                 # local packed rows of the lower triangle
        full = seg if self.packed_is_full else ll._gather_packed(seg)
        out = torch.empty((o, o, v, v), dtype=torch.float64, device=ll.device)
        ll._ctx.call("mpccb_unpack_oovv", full.data_ptr(), out.data_ptr())
        ll._sync()
        return out.cpu().numpy()

    def _packed(self):
        """The packed t2 part. An extrapolated vector is materialised only here — nothing on the hot path reads the
        t2 part of the DIIS extrapolation (the next iteration rebuilds t2 from t1, :135-137)."""
        if self.packed is None:
            self.packed = self.source.packed_device()[self.solver.nov:]
        return self.packed

    def _check_fresh(self):
        """An "oovv" handle from ``run_diis`` aliases the DIIS object's extrapolation buffer, which the next
        ``update`` overwrites in place (the reference returns a fresh array each time, diis.py:268-275)."""
        if self.owner is not None and self.serial != self.owner.generation:
            raise RuntimeError("this t2 handle is stale: a later DIIS update overwrote the extrapolated vector; "
                               "materialise it (np.asarray) before the next run_diis()")

    def _local_device(self):
        ll = self.solver
        o, v = ll.nocc, ll.nvir
        if self.serial != ll._t2_serial:
            raise RuntimeError("this t2 handle is stale: the device buffer was overwritten by a later updated_amps()")
        out = torch.empty((ll._na, o, v, o), dtype=torch.float64, device=ll.device)
        ll._ctx.call("mpccb_t2_to_aibj", out.data_ptr())
        ll._sync()
        return out

    def local_numpy(self):
        """This rank's part: the (na,o,v,o) slab for "aibj"; the whole (o,o,v,v) tensor on one rank."""
        ll = self.solver
        o, v = ll.nocc, ll.nvir
        if self.layout == "aibj":
            return self._local_device().cpu().numpy()
        self._check_fresh()
        if ll.world_size > 1 and not self.packed_is_full:
            raise RuntimeError("the (o,o,v,v) form of a sharded vector needs the collective numpy()")
        packed = ll._gather_packed(self._packed()) if (ll._diis_columns and not self.packed_is_full) else self._packed()
        out = torch.empty((o, o, v, v), dtype=torch.float64, device=ll.device)
        ll._ctx.call("mpccb_unpack_oovv", packed.data_ptr(), out.data_ptr())
        ll._sync()
        return out.cpu().numpy()

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a if dtype is None else a.astype(dtype)


class MPCC_LL:
    """GPU-resident equivalent of ``pyscf.mpcc.dfrmpcc_lowlevel.MPCC_LL``."""

    def __init__(self, mf=None, *, blocks=None, device=None, group=None, use_tma=True, t2_wide_tiles=None, t2_unfused=None, yvo_split=None,
                 jk_exchange=None, t2_sym=None, slabs=None, diis_layout=None):
        """``MPCC_LL(mf)`` as in the reference (:7-43). Extensions (keyword only):

        blocks : (Loo, Lov, Lvv, foo, fvv, mo_energy) numpy arrays or CUDA tensors
                 already in the MO basis — skips ``compute_three_center_ints``.
        device : torch device (default ``cuda:<LOCAL_RANK or 0>``).
        group  : torch.distributed process group to shard the doubles over
                 (default: the WORLD group if torch.distributed is initialised).
        slabs  : how the virtual index is split over the ranks: "equal" (default), "balanced" (widths from the
                 cost model ``sharding.balanced_bounds``; measured slower) or an explicit list
                 of world_size + 1 bounds.
        diis_layout : t2 part of the DIIS vector. "columns" (default with several ranks: each rank keeps the packed elements whose SECOND
                 virtual index is in its slab — 1/P of the vector each) or "rows" (contiguous reference-order
                 segments, (2P-1):1 unbalanced; the only form on one rank).
        """
        if not torch.cuda.is_available():
            raise RuntimeError("pyscf_mpcc_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        self.mf = mf
        self.with_df = getattr(mf, "with_df", None) if mf is not None else None
        if mf is not None and self.with_df is None:
            # the reference's no-with_df branch raises NameError (df is not imported, :13); be explicit
            raise RuntimeError("MPCC_LL needs a density-fitted mean field (mf.density_fit())")
        self.t1 = None
        self.t2 = None
        self.diis = True
        self.con_tol = 1e-6
        self.max_its = 50
        self.verbose_print = True
        self.stdout = sys.stdout

        if torch.distributed.is_available() and torch.distributed.is_initialized():
            self.group = group if group is not None else torch.distributed.group.WORLD
            self.world_size = torch.distributed.get_world_size(self.group)
            self.rank = torch.distributed.get_rank(self.group)
        else:
            self.group, self.world_size, self.rank = None, 1, 0
        if device is None:
            # one process per GPU under torchrun: cuda:<LOCAL_RANK>; otherwise the current device
            local = os.environ.get("LOCAL_RANK")
            if local is not None and self.world_size > 1 and int(local) < torch.cuda.device_count():
                device = torch.device("cuda", int(local))
            else:
                device = torch.device("cuda", torch.cuda.current_device())
        self.device = torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        torch.cuda.set_device(self.device)

        if mf is not None:
            self.nao = int(mf.mol.nao)
            self.nocc = int(mf.mol.nelec[0])
            self.nvir = self.nao - self.nocc
            self.naux = int(self.with_df.get_naoaux())
            self.mo_energy = np.asarray(mf.mo_energy, dtype=np.float64)
            self.mo_coeff = np.asarray(mf.mo_coeff, dtype=np.float64)
        else:
            Loo, Lov, Lvv, foo, fvv, mo_energy = blocks
            self.naux, self.nocc, self.nvir = (int(s) for s in Lov.shape)
            self.nao = self.nocc + self.nvir
            self.mo_energy = self._to_host(mo_energy)
            self.mo_coeff = None
        self.nvir_pair = self.nvir * (self.nvir + 1) // 2
        self.nov = self.nocc * self.nvir
        # pitch of the innermost index of the device t2 buffer T[a][b][i][j] (and of Lvo[L][b][j]): nocc rounded up to
        # even, so the TMA tensor maps of the Yvo GEMM are legal for odd nocc as well (16-byte strides)
        self._jp = self.nocc + (self.nocc & 1)
        o, v, L = self.nocc, self.nvir, self.naux

        # --- native context, workspace, device buffers (torch owns the memory) ---
        if slabs is None:
            slabs = os.environ.get("MPCC_SLABS", "equal")
        if isinstance(slabs, str):
            if slabs not in ("balanced", "equal"):
                raise ValueError(f"slabs={slabs!r}")
            self._bounds = (sharding.balanced_bounds(o, v, L, self.world_size) if slabs == "balanced" else
                            [sharding.slab_bounds(v, self.world_size, r)[0] for r in range(self.world_size)] + [v])
        else:
            self._bounds = [int(b) for b in slabs]
            if (len(self._bounds) != self.world_size + 1 or self._bounds[0] != 0 or self._bounds[-1] != v
                    or any(hi <= lo for lo, hi in zip(self._bounds[:-1], self._bounds[1:]))):
                raise ValueError(f"slabs={slabs!r}: need {self.world_size + 1} increasing bounds from 0 to nvir={v}")
        self._a_begin, self._a_end = self._bounds[self.rank], self._bounds[self.rank + 1]
        self._na = self._a_end - self._a_begin
        self._count_t1 = (self.rank == 0)
        self._vec_len_local = sharding.local_vector_length(o, v, self._a_begin, self._a_end)
        if diis_layout is None:
            diis_layout = os.environ.get("MPCC_DIIS_LAYOUT", "columns" if self.world_size > 1 else "rows")
        if diis_layout not in ("columns", "rows"):
            raise ValueError(f"diis_layout={diis_layout!r}")
        self._diis_columns = diis_layout == "columns"
        # xprev folded into the push (row-owned kernel only) or stored and combined next to the next dressing
        imp = os.environ.get("MPCC_DIIS_IMPLICIT", "auto")
        if imp not in ("auto", "0", "1"):
            raise ValueError(f"MPCC_DIIS_IMPLICIT={imp!r}")
        self._diis_implicit = (not self._diis_columns) if imp == "auto" else (imp == "1" and not self._diis_columns)
        self._diis_prefix = None
        if self._diis_columns:
            prefix = sharding.column_owned_prefix(o, v, self._a_begin, self._a_end)
            self._vec_len_local = self.nov + int(prefix[-1])
            self._diis_prefix = torch.from_numpy(prefix).to(self.device)
        self._ctx = _cabi.Context(o, v, L, self._a_begin, self._a_end, self.device.index)
        self._ctx.call("mpccb_set_option", b"tma", int(bool(use_tma)))
        if os.environ.get("MPCC_YVO_FLAT"):                     # A/B switch: 1 / -1 force the flat-K Yvo form on / off
            self._ctx.call("mpccb_set_option", b"yvo_flat", int(os.environ["MPCC_YVO_FLAT"]))
        if os.environ.get("MPCC_MAPPED_NARROW_K"):              # A/B switch: K up to which mapped GEMMs use 128 x 64 tiles
            self._ctx.call("mpccb_set_option", b"mapped_narrow_k", int(os.environ["MPCC_MAPPED_NARROW_K"]))
        for key in ("raster_yvo", "raster_eris"):               # A/B switch: column tiles per raster group (0 = default)
            if os.environ.get("MPCC_" + key.upper()):
                self._ctx.call("mpccb_set_option", key.encode(), int(os.environ["MPCC_" + key.upper()]))
        if os.environ.get("MPCC_XVO_GENERIC", "0") != "0":      # A/B switch: Xvo GEMM (:119) on the generic kernel
            self._ctx.call("mpccb_set_option", b"xvo_generic", 1)
        if self._diis_columns:
            self._ctx.call("mpccb_bind_diis_columns", self._diis_prefix.data_ptr())
        if t2_wide_tiles is None and os.environ.get("MPCC_T2_WIDE_TILES"):
            t2_wide_tiles = os.environ["MPCC_T2_WIDE_TILES"] != "0"
        if t2_sym is None and os.environ.get("MPCC_T2_SYM"):
            t2_sym = os.environ["MPCC_T2_SYM"] != "0"
        if t2_sym is not None:             # mirror scheme of the eris -> t2 step (default on)
            self._ctx.call("mpccb_set_option", b"t2_sym", int(bool(t2_sym)))
        self._t2_sym = True if t2_sym is None else bool(t2_sym)
        if yvo_split is not None:
            self._ctx.call("mpccb_set_option", b"yvo_split", int(yvo_split))
        if t2_unfused is not None:
            self._ctx.call("mpccb_set_option", b"t2_unfused", int(bool(t2_unfused)))
        if t2_wide_tiles is not None:      # tuning knob: 128x128 tiles (1 CTA/SM) vs 128x64 (2 CTAs/SM) for eris->t2
            self._ctx.call("mpccb_set_option", b"t2_wide_tiles", int(bool(t2_wide_tiles)))
        self._bind_stream()
        f64 = dict(dtype=torch.float64, device=self.device)
        self._ws = torch.empty(self._ctx.workspace_bytes() + 256, dtype=torch.uint8, device=self.device)
        ws_ptr = (self._ws.data_ptr() + 255) // 256 * 256
        self._ctx.call("mpccb_bind_workspace", ws_ptr, self._ctx.workspace_bytes())
        self._jk_mode = "local"
        self._xoo_hdl, self._xoo_mode = None, "replicated"
        if self.world_size > 1:
            self._setup_jk_exchange(jk_exchange or os.environ.get("MPCC_JK_EXCHANGE", "auto"))
        self._t2_peers = False
        if self.world_size > 1 and self._t2_sym and self._jk_mode in ("multicast", "p2p", "subset"):
            self._setup_t2_peers()
        if self._jk_mode == "subset" and not self._t2_peers:
            self._bind_jk(subset=False)                  # without the cross-rank mirror every rank needs every JK row
        if not self._t2_peers:
            # odd nocc: the pad column is read by TMA boxes (and multiplied by Lvo's zero column) — it must stay finite
            self._T = (torch.zeros if self._jp != o else torch.empty)(self._na * v * o * self._jp, **f64)
            self._ctx.call("mpccb_bind_t2", self._T.data_ptr())
        self._t2_serial = 0
        self._red = torch.zeros(self.nov + 1, **f64)
        self._t1_dev = torch.zeros(self.nov, **f64)
        self._t1n_dev = torch.zeros(self.nov, **f64)
        self._scal_dev = torch.zeros(2, **f64)
        self._dots_dev = torch.zeros(7, **f64)
        self._t1_host = torch.zeros(self.nov, dtype=torch.float64).pin_memory()
        self._t1n_host = torch.zeros(self.nov, dtype=torch.float64).pin_memory()
        self._scal_host = torch.zeros(2, dtype=torch.float64).pin_memory()
        self._dots_host = torch.zeros(7, dtype=torch.float64).pin_memory()

        if mf is not None:
            self.compute_three_center_ints()
            fock_mo = self.mo_coeff.T @ np.asarray(mf.get_fock()) @ self.mo_coeff
            self._fock_mo = fock_mo
            foo = np.ascontiguousarray(fock_mo[:o, :o])
            fvv = np.ascontiguousarray(fock_mo[o:, o:])
        else:
            self._d = {"Loo": self._to_dev(Loo), "Lov": self._to_dev(Lov), "Lvv": self._to_dev(Lvv)}
        self._d["foo"] = self._to_dev(foo)
        self._d["fvv"] = self._to_dev(fvv)
        self._d["eo"] = self._to_dev(self.mo_energy[:o])
        self._d["ev"] = self._to_dev(self.mo_energy[o:])
        d = self._d
        self._ctx.call("mpccb_bind_integrals", d["Loo"].data_ptr(), d["Lov"].data_ptr(), d["Lvv"].data_ptr(),
                       d["foo"].data_ptr(), d["fvv"].data_ptr(), d["eo"].data_ptr(), d["ev"].data_ptr())
        # eia[i,a] = e_a - e_i  (:40-42)
        self.eia = self.mo_energy[o:][None, :] - self.mo_energy[:o][:, None]
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    # ------------------------------------------------------------------ plumbing
    def _bind_stream(self):
        self._stream = torch.cuda.current_stream(self.device)
        self._ctx.call("mpccb_set_stream", self._stream.cuda_stream)

    def _sync(self):
        self._stream.synchronize()

    def _to_dev(self, x):
        if isinstance(x, torch.Tensor):
            return x.to(device=self.device, dtype=torch.float64).contiguous()
        return torch.from_numpy(np.ascontiguousarray(x, dtype=np.float64)).to(self.device)

    @staticmethod
    def _to_host(x):
        if isinstance(x, torch.Tensor):
            return x.detach().cpu().numpy().astype(np.float64)
        return np.asarray(x, dtype=np.float64)

    def _host_attr(self, name):
        self._sync()
        return self._d[name].cpu().numpy()

    def _gather_flat(self, seg, lengths):
        """All-gather of per-rank 1-D device segments of unequal `lengths` (padded to the longest)."""
        n = max(lengths)
        pad = torch.zeros(n, dtype=torch.float64, device=self.device)
        pad[:seg.numel()].copy_(seg.reshape(-1))
        parts = [torch.empty_like(pad) for _ in range(self.world_size)]
        self._sync()
        torch.distributed.all_gather(parts, pad, group=self.group)
        return torch.cat([p[:k] for p, k in zip(parts, lengths)])

    def _gather_rows(self, slab, nvir):
        """All-gather of the per-rank (na_r, width) row slabs into (nvir, width)."""
        width = slab.shape[1]
        lens = []
        for a0, a1 in zip(self._bounds[:-1], self._bounds[1:]):
            lens.append((a1 - a0) * width)
        return self._gather_flat(slab.reshape(-1), lens).reshape(nvir, width)

    def _gather_packed(self, seg):
        """The reference-order packed lower triangle (nov(nov+1)/2 doubles, on every rank) from the per-rank t2
        parts of a DIIS vector. Collective; host-facing paths only (``numpy()`` of a vector / an (o,o,v,v) t2)."""
        if not self._diis_columns:
            lens = [sharding.local_vector_length(self.nocc, self.nvir, a0, a1) - self.nov
                    for a0, a1 in zip(self._bounds[:-1], self._bounds[1:])]
            return self._gather_flat(seg, lens)
        full = torch.zeros(self.nov * (self.nov + 1) // 2, dtype=torch.float64, device=self.device)
        self._bind_stream()
        self._ctx.call("mpccb_diis_scatter_tril", seg.data_ptr(), full.data_ptr())
        self._sync()
        if self.world_size > 1:
            torch.distributed.all_reduce(full, group=self.group)     # disjoint supports: a sum of one value and zeros
        return full

    # host-readable views of the device-resident tensors (reference attribute names, :66-69,:36-37)
    Loo = property(lambda self: self._host_attr("Loo"))
    Lov = property(lambda self: self._host_attr("Lov"))
    Lvv = property(lambda self: self._host_attr("Lvv"))
    foo = property(lambda self: self._host_attr("foo"))
    fvv = property(lambda self: self._host_attr("fvv"))

    @property
    def Lvo(self):
        """(L,v,o) transpose of Lov — the reference builds it (:50,:63) and never uses it; here it
        is the K-major operand of the Yvo GEMM."""
        p, n = self._ctx.buffer("Lvo")
        return np.ascontiguousarray(self._read_device(p, n).reshape(self.naux, self.nvir, self._jp)[:, :, :self.nocc])

    @property
    def D(self):
        """D[a,i,b,j] = -eia[i,a] - eia[j,b] (:43). Never materialised on the device."""
        e = self.eia.T
        return -(e[:, :, None, None] + e[None, None, :, :])

    def _read_device(self, ptr, count):
        out = np.empty(count, dtype=np.float64)
        self._ctx.call("mpccb_copy_to_host", ptr, out.ctypes.data, count)
        return out

    def debug_buffer(self, name):
        """Host copy of a named scratch tensor (tests / profiling only)."""
        p, n = self._ctx.buffer(name)
        out = self._read_device(p, n)
        if name in ("T", "Lvo") and self._jp != self.nocc:        # strip the pad column of the j index
            out = np.ascontiguousarray(out.reshape(-1, self._jp)[:, :self.nocc]).reshape(-1)
        return out

    # ------------------------------------------------------------------ :46-69
    def compute_three_center_ints(self):
        """AO->MO transform of the Cholesky tensor (:46-69), block by block as ``with_df.loop()``
        yields it: the packed rows are uploaded and unpacked, transformed (two DMMA GEMM passes) and
        sliced into Loo/Lov/Lvv on the device (``mpccb_ao2mo_block``). T1-independent, once per solve."""
        nao, o, L = self.nao, self.nocc, self.naux
        f64 = dict(dtype=torch.float64, device=self.device)
        Loo = torch.empty((L, o, o), **f64)
        Lov = torch.empty((L, o, nao - o), **f64)
        Lvv = torch.empty((L, nao - o, nao - o), **f64)
        C = self._to_dev(self.mo_coeff)
        if tuple(C.shape) != (nao, nao):
            raise ValueError("mo_coeff must be (nao, nao): no frozen orbitals on this path (:23-25)")
        scratch, p1 = None, 0
        for packed in self.with_df.loop():
            nb = int(packed.shape[0])
            blk = self._to_dev(packed)
            need = int(self._ctx.lib.mpccb_ao2mo_scratch_bytes(self._ctx.h, nb))
            if scratch is None or scratch.numel() < need:
                scratch = torch.empty(need, dtype=torch.uint8, device=self.device)
            self._ctx.call("mpccb_ao2mo_block", blk.data_ptr(), nb, p1, C.data_ptr(), Loo.data_ptr(),
                           Lov.data_ptr(), Lvv.data_ptr(), scratch.data_ptr(), need)
            self.h2d_bytes = getattr(self, "h2d_bytes", 0) + blk.numel() * 8
            p1 += nb
        if p1 != L:
            raise ValueError(f"with_df.loop() yielded {p1} auxiliary rows, get_naoaux() says {L}")
        self._sync()
        self._d = {"Loo": Loo, "Lov": Lov, "Lvv": Lvv}

    # ------------------------------------------------------------------ :103-152
    def _setup_jk_exchange(self, mode):
        """How the slab-dressed integrals Jvo reach every rank.

        "multicast" / "p2p": the JK buffers live in symmetric (peer-mapped) memory and the dressing kernel
        stores its rows into every rank's buffer itself — through the NVSwitch multicast address when the
        allocation has one, else one NVLink store per peer; only a stream barrier follows.
        "nccl": slab-local store + ``all_gather_into_tensor``. "auto" tries the fused forms first; all ranks
        agree on the outcome (a failed rendezvous on any rank sends everyone to "nccl")."""
        dist = torch.distributed
        o, v = self.nocc, self.nvir
        ptr, count = self._ctx.buffer("JK")
        ldk = count // self.nov
        if mode not in ("auto", "multicast", "p2p", "subset", "nccl"):
            raise ValueError(f"jk_exchange={mode!r}")
        ok, hdl, buf = 0, None, None
        if mode != "nccl" and self._symm_agreed():
            try:
                import torch.distributed._symmetric_memory as symm
                buf = symm.empty(count + 2 * self.nov, dtype=torch.float64, device=self.device)   # JK rows + w + Fvo
                hdl = symm.rendezvous(buf, self.group)
                ok = 1
            except Exception as e:                      # no peer mapping on this system: use NCCL
                self._jk_exchange_error = repr(e)
        flag = torch.tensor([ok], device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) == 1:
            buf.zero_()                                  # K padding columns must read as zero
            self._sync()
            torch.cuda.current_stream(self.device).synchronize()
            hdl.barrier(channel=0)                       # nobody stores into a buffer that is still being cleared
            mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
            self._jk_buf, self._jk_hdl, self._jk_mc = buf, hdl, mc
            # "subset": one store per rank that needs the slab (world/2 + 1 of them under the mirror scheme) instead of
            # a multicast to everybody, and w / Fvo broadcast with it (F2 becomes a GEMM over the slab's rows); needs
            # the cross-rank mirror (peer-mapped t2). Measured faster than the multicast form from 2 ranks on.
            want_subset = mode in ("subset", "auto")
            if want_subset and self._t2_sym:
                self._bind_jk(subset=True)
            else:
                self._bind_jk(subset=False, multicast=mode in ("auto", "multicast", "subset"))
            self._setup_xoo_exchange()
            return
        off = ptr - self._ws.data_ptr()
        self._jk_view = self._ws[off:off + count * 8].view(torch.float64)
        self._jk_rows = []
        for b0, b1 in zip(self._bounds[:-1], self._bounds[1:]):
            self._jk_rows.append((b0 * o * ldk, b1 * o * ldk))
        self._equal_slabs = len({hi - lo for lo, hi in self._jk_rows}) == 1
        self._jk_mode = "nccl"

    def _setup_xoo_exchange(self):
        """Xoo / Joo (:120, :123) sharded over the auxiliary index and exchanged through symmetric memory
        (include/mpcc_b200.h, mpccb_bind_xoo_exchange) instead of the same GEMM on every rank. MPCC_XOO_EXCHANGE=0
        keeps the replicated form. All ranks agree on the outcome."""
        want = os.environ.get("MPCC_XOO_EXCHANGE", "1") != "0" and self.naux >= self.world_size
        ok, hdl, buf = 0, None, None
        if want:
            try:
                import torch.distributed._symmetric_memory as symm
                buf = symm.empty(2 * self.naux * self.nocc * self.nocc, dtype=torch.float64, device=self.device)
                hdl = symm.rendezvous(buf, self.group)
                ok = 1
            except Exception as e:
                self._xoo_exchange_error = repr(e)
        flag = torch.tensor([ok], device=self.device)
        torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN, group=self.group)
        if int(flag.item()) != 1:
            return
        peers = (ctypes.c_void_p * self.world_size)(*[int(p) for p in hdl.buffer_ptrs])
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        use_mc = mc != 0 and os.environ.get("MPCC_XOO_MULTICAST", "1") != "0"
        self._ctx.call("mpccb_bind_xoo_exchange", buf.data_ptr(), peers, self.world_size, self.rank,
                       mc if use_mc else None)
        self._xoo_buf, self._xoo_hdl = buf, hdl
        self._xoo_mode = "multicast" if use_mc else "p2p"

    def _bind_jk(self, subset, multicast=True):
        hdl, buf, mc = self._jk_hdl, self._jk_buf, self._jk_mc
        peers = (ctypes.c_void_p * self.world_size)(*[int(p) for p in hdl.buffer_ptrs])
        if subset:
            self._ctx.call("mpccb_bind_jk_subset", buf.data_ptr(), peers, self.world_size, self.rank, mc if mc else None)
            self._jk_mode = "subset"
            return
        use_mc = mc != 0 and multicast
        self._ctx.call("mpccb_bind_jk_exchange", buf.data_ptr(), peers, self.world_size, mc if use_mc else None)
        self._jk_mode = "multicast" if use_mc else "p2p"

    def _symm_agreed(self):
        """All ranks agree, BEFORE anyone enters the collective rendezvous, that symmetric memory can be tried: a
        local probe (module importable, allocation works) reduced with MIN over the group. A rank that cannot even
        allocate would otherwise leave the others blocked inside ``rendezvous``."""
        if getattr(self, "_symm_ok", None) is None:
            ok = 1
            try:
                import torch.distributed._symmetric_memory as symm
                probe = symm.empty(16, dtype=torch.float64, device=self.device)
                del probe
            except Exception as e:
                ok = 0
                self._symm_probe_error = repr(e)
            flag = torch.tensor([ok], device=self.device)
            torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN, group=self.group)
            self._symm_ok = int(flag.item()) == 1
        return self._symm_ok

    def _setup_t2_peers(self):
        """Cross-rank half of the mirror scheme (include/mpcc_b200.h, mpccb_bind_t2_peers): the t2 slabs live in
        symmetric memory so that the rank computing block (a,b) can store the mirror block t2[b,a] straight into
        the buffer of the rank that owns b. All ranks agree; on failure everyone keeps private buffers."""
        dist = torch.distributed
        o, v = self.nocc, self.nvir
        bounds = self._bounds
        n_max = max(bounds[r + 1] - bounds[r] for r in range(self.world_size)) * v * o * self._jp
        ok, buf, hdl = 0, None, None
        if not self._symm_agreed():
            return
        try:
            import torch.distributed._symmetric_memory as symm
            buf = symm.empty(n_max, dtype=torch.float64, device=self.device)
            hdl = symm.rendezvous(buf, self.group)
            ok = 1
        except Exception as e:
            self._t2_peers_error = repr(e)
        flag = torch.tensor([ok], device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if int(flag.item()) != 1:
            return
        peers = (ctypes.c_void_p * self.world_size)(*[int(p) for p in hdl.buffer_ptrs])
        self._ctx.call("mpccb_bind_t2_peers", self.world_size, self.rank, (ctypes.c_int * (self.world_size + 1))(*bounds), peers)
        if self._jp != o:
            buf.zero_()
            self._sync()
            hdl.barrier(channel=1)
        self._T = buf[:self._na * v * o * self._jp]
        self._t2_buf, self._t2_hdl = buf, hdl
        self._t2_peers = True

    def _doubles(self):
        """Phase 2 (:135-147). With peer-mapped t2 buffers other ranks store mirror blocks into this rank's slab
        during the eris/t2 kernel, so a cross-rank barrier separates it from the Yvo GEMM that reads the slab."""
        ctx = self._ctx
        if not self._t2_peers:
            ctx.call("mpccb_doubles", self._red.data_ptr())
            return
        ctx.call("mpccb_doubles_begin", self._red.data_ptr())
        self._t2_hdl.barrier(channel=1)
        ctx.call("mpccb_doubles_end", self._red.data_ptr())

    def _dress(self, t1_dev):
        """T1 dressing (:119-133). One rank: a single native call. Several ranks: each dresses its own
        virtual slab, the dressed integrals Jvo (rows of the K-major JK buffer; a slab is a contiguous
        row range) are all-gathered over NVLink, then the terms that need every row are formed."""
        ctx = self._ctx
        ctx.call("mpccb_dress", t1_dev.data_ptr())
        if self.world_size == 1:
            return
        if self._xoo_hdl is not None:
            self._xoo_hdl.barrier(channel=0)             # every rank's L slab of Xoo / Joo has landed everywhere
            ctx.call("mpccb_dress_rest")
        if self._jk_mode in ("multicast", "p2p", "subset"):
            self._jk_hdl.barrier(channel=0)              # every rank's rows have landed in every buffer
        else:
            jk, rows = self._jk_view, self._jk_rows
            if self._equal_slabs:
                torch.distributed.all_gather_into_tensor(jk, jk[rows[self.rank][0]:rows[self.rank][1]],
                                                         group=self.group)
            else:
                for r, (lo, hi) in enumerate(rows):
                    torch.distributed.broadcast(jk[lo:hi], src=torch.distributed.get_global_rank(self.group, r),
                                                group=self.group)
        ctx.call("mpccb_dress_post")

    # ------------------------------------------------------------------ :103-152
    def updated_amps(self):
        """One residual evaluation from ``self.t1``; returns (res, dE, t1_new, t2) with
        t1_new a host (o,v) array and t2 a lazy (v,o,v,o) handle on the device buffer."""
        self._bind_stream()
        t1 = self.t1
        if isinstance(t1, torch.Tensor) and t1.is_cuda:
            self._t1_dev.copy_(t1.reshape(-1))
        else:
            self._t1_host.copy_(torch.from_numpy(np.ascontiguousarray(t1, dtype=np.float64).reshape(-1)))
            self._t1_dev.copy_(self._t1_host, non_blocking=True)
            self.h2d_bytes += self.nov * 8
        ctx = self._ctx
        self._dress(self._t1_dev)
        self._doubles()
        self._t2_serial += 1
        if self.world_size > 1:
            torch.distributed.all_reduce(self._red, group=self.group)
        ctx.call("mpccb_finish", self._red.data_ptr(), self._t1n_dev.data_ptr(), self._scal_dev.data_ptr())
        self._t1n_host.copy_(self._t1n_dev, non_blocking=True)
        self._scal_host.copy_(self._scal_dev, non_blocking=True)
        self._sync()
        self.d2h_bytes += self.nov * 8 + 16
        res = float(np.sqrt(self._scal_host[0].item()))
        dE = float(self._scal_host[1].item())
        t1_new = self._t1n_host.numpy().reshape(self.nocc, self.nvir).copy()
        self._last_t1_new = t1_new          # lets run_diis reuse the device copy instead of re-uploading
        return res, dE, t1_new, DeviceT2(self, "aibj", serial=self._t2_serial)

    update_amps = updated_amps       # north_star / PySCF-CC spelling

    def step_device(self, t1_dev, adiis=None):
        """One trip of the ``kernel()`` loop (:84-86) with the amplitudes resident in HBM:
        ``t1_dev`` (nov doubles, CUDA) in, next t1 (CUDA view) out; only the two scalars and
        the <= 6 DIIS dot products cross PCIe. Returns (res, dE, t1_next_dev)."""
        self._bind_stream()
        ctx = self._ctx
        self._dress(t1_dev)
        self._doubles()
        self._t2_serial += 1
        if self.world_size > 1:
            torch.distributed.all_reduce(self._red, group=self.group)
        ctx.call("mpccb_finish", self._red.data_ptr(), self._t1n_dev.data_ptr(), self._scal_dev.data_ptr())
        self._scal_host.copy_(self._scal_dev, non_blocking=True)
        if adiis is not None:
            # update() waits for the DIIS dot products, which follow the two scalars on the stream: both are on the
            # host when it returns, and the kernels it launches afterwards need not be waited for
            nxt = adiis.update(AmplitudeVector(self, vec=None, t1_dev=self._t1n_dev)).t1_dev
            if not getattr(adiis, "synced", False):
                self._sync()
        else:
            nxt = self._t1n_dev
            self._sync()
        return float(np.sqrt(self._scal_host[0].item())), float(self._scal_host[1].item()), nxt

    def energy(self, t1=None):
        """Correlation energy dE of ``t1`` (default ``self.t1``): the second return value
        of ``updated_amps`` (:146-147; standalone form ``compute_CC2_energy``,
        ``pyscf/cc/rmpcc_lowlevel.py:82-109``)."""
        keep = self.t1
        if t1 is not None:
            self.t1 = t1
        try:
            return self.updated_amps()[1]
        finally:
            self.t1 = keep

    # ------------------------------------------------------------------ PySCF's to_gpu convention
    @classmethod
    def from_cpu(cls, other, **kwargs):
        """GPU solver from an already constructed CPU ``MPCC_LL`` of the reference (or anything with its
        attributes ``Loo, Lov, Lvv, foo, fvv, mo_energy``, ``dfrmpcc_lowlevel.py:29-43,66-69``): the MO-basis
        blocks are uploaded as they are — no second AO->MO transform — and the outside pokes (``con_tol``,
        ``max_its``, ``diis``), the mean field and the current ``t1`` are carried over. This is what
        ``obj.to_gpu()`` means elsewhere in PySCF (``lib.to_gpu``, ``pyscf/lib/misc.py:1508-1556``); the MPCC
        classes of the reference do not define it, so a maintainer adds
        ``MPCC_LL.to_gpu = lambda self: pyscf_mpcc_b200.lowlevel.to_gpu(self)``."""
        blocks = tuple(np.ascontiguousarray(getattr(other, k), dtype=np.float64)
                       for k in ("Loo", "Lov", "Lvv", "foo", "fvv", "mo_energy"))
        new = cls(blocks=blocks, **kwargs)
        new.mf = getattr(other, "mf", None)
        new.with_df = getattr(other, "with_df", None)
        if getattr(other, "mo_coeff", None) is not None:
            new.mo_coeff = np.asarray(other.mo_coeff, dtype=np.float64)
        for k in ("con_tol", "max_its", "diis"):
            if hasattr(other, k):
                setattr(new, k, getattr(other, k))
        t1 = getattr(other, "t1", None)
        if t1 is not None:
            new.t1 = np.array(t1, dtype=np.float64).reshape(new.nocc, new.nvir)
        return new

    # ------------------------------------------------------------------ high-level stage (rows f1/f2)
    def tensor_engine(self):
        """Device contraction engine bound to this solver's native context (stream, workspace, orbital energies)."""
        from .tensor_engine import CudaBackend, Engine
        self._bind_stream()
        return Engine(CudaBackend(self._ctx, self.device))

    def high_level_eris(self, fock=None):
        """Four-index integrals for the CCSD-type stage (``pyscf_mpcc_b200.rmpccsd``), built on the device from
        this solver's DF tensors. ``fock``: full MO Fock matrix (default: the one of the mean field, or
        blockdiag(foo, fvv) when the solver was built from MO blocks)."""
        from . import rmpccsd
        from .tensor_engine import DTensor
        if self.world_size > 1:
            raise NotImplementedError(
                "the CCSD-type stage keeps its integrals replicated: build this solver on a one-rank group "
                "(group=torch.distributed.new_group([rank])) and pass the sharding group to rmpccsd.RMPCCSD(group=...) "
                "/ cc.group, which shards the particle-particle ladder and the ring terms (tools/cc_mgpu_check.py)")
        o, v, L = self.nocc, self.nvir, self.naux
        if fock is None:
            fock = getattr(self, "_fock_mo", None)
        if fock is None:
            fock = np.zeros((o + v, o + v))
            fock[:o, :o] = self.foo
            fock[o:, o:] = self.fvv
        eng = self.tensor_engine()
        d = self._d
        return rmpccsd.DFERIs(eng, DTensor(d["Loo"].reshape(-1), (L, o, o)), DTensor(d["Lov"].reshape(-1), (L, o, v)),
                              DTensor(d["Lvv"].reshape(-1), (L, v, v)), fock)

    # ------------------------------------------------------------------ :155-176
    def amplitudes_to_vector(self, t1, t2, out=None):
        """[t1 | tril(M)] (:162-168) as a device-side handle; ``.numpy()`` gives the flat host vector."""
        if isinstance(t2, DeviceT2):
            if t2.layout != "aibj" or t2.serial != self._t2_serial:
                raise RuntimeError("amplitudes_to_vector needs the t2 of the latest updated_amps()")
        else:
            self._upload_t2(np.asarray(t2))
        if t1 is getattr(self, "_last_t1_new", None):
            t1_dev = self._t1n_dev
        elif isinstance(t1, torch.Tensor):
            t1_dev = t1.to(self.device).reshape(-1)
        else:
            t1_dev = self._to_dev(np.ascontiguousarray(t1).reshape(-1))
            self.h2d_bytes += self.nov * 8
        return AmplitudeVector(self, vec=None, t1_dev=t1_dev)

    def _upload_t2(self, t2):
        """Accept a host 4-index array the way the reference would (:166): M = t2.transpose(0,2,1,3)
        read as (nov,nov). Stored into the device buffer in that order (single-rank, small cases)."""
        if self.world_size > 1:
            raise NotImplementedError("host t2 upload with several ranks")
        m = np.ascontiguousarray(t2.transpose(0, 2, 1, 3)).reshape(-1)
        o, jp = self.nocc, self._jp
        if m.size * jp != self._T.numel() * o:
            raise ValueError("t2 has the wrong size")
        if jp != o:                                    # rows of nocc elements onto the even pitch of the device buffer
            padded = np.zeros((m.size // o, jp))
            padded[:, :o] = m.reshape(-1, o)
            m = padded.reshape(-1)
        self._T.copy_(torch.from_numpy(m))
        self._t2_serial += 1

    def vector_to_amplitudes(self, vector):
        """(:170-176) t1 (o,v) host copy; t2 (o,o,v,v) lazy handle."""
        if isinstance(vector, AmplitudeVector) and vector.t1_dev is not None:
            # device-side vector: t1 from its first nov elements; the t2 part stays where it is (for an extrapolated
            # vector: un-materialised) until somebody reads the (o,o,v,v) tensor
            self._t1n_host.copy_(vector.t1_dev.reshape(-1), non_blocking=True)
            self._sync()
            self.d2h_bytes += self.nov * 8
            t1 = self._t1n_host.numpy().reshape(self.nocc, self.nvir).copy()
            return t1, DeviceT2(self, "oovv", source=vector, owner=vector.owner, serial=vector.serial)
        full, owner, serial = False, None, 0
        if isinstance(vector, AmplitudeVector):
            vec = vector.packed_device()
            owner, serial = vector.owner, vector.serial
        else:
            # a host vector is the reference-order flat vector (what AmplitudeVector.numpy() returns): the whole
            # packed triangle, whatever the device-side sharding / ownership of DIIS vectors is
            host = np.asarray(vector, dtype=np.float64).reshape(-1)
            if host.size != self.nov + self.nov * (self.nov + 1) // 2:
                raise ValueError(f"vector has {host.size} elements, expected nov + nov(nov+1)/2 = "
                                 f"{self.nov + self.nov * (self.nov + 1) // 2}")
            vec = self._to_dev(host)
            self.h2d_bytes += host.size * 8
            full = True
        self._t1n_host.copy_(vec[:self.nov], non_blocking=True)
        self._sync()
        self.d2h_bytes += self.nov * 8
        t1 = self._t1n_host.numpy().reshape(self.nocc, self.nvir).copy()
        return t1, DeviceT2(self, "oovv", packed=vec[self.nov:], packed_is_full=full, owner=owner, serial=serial)

    def run_diis(self, t1, t2, adiis):
        vec = self.amplitudes_to_vector(t1, t2)
        t1, t2 = self.vector_to_amplitudes(adiis.update(vec))
        return t1, t2

    # ------------------------------------------------------------------ :72-100
    def kernel(self):
        self.t1 = np.zeros((self.nocc, self.nvir))
        self.t2 = np.zeros((self.nocc, self.nocc, self.nvir, self.nvir)) if self.nov ** 2 < 2 ** 24 else None
        err = np.inf
        count = 0
        adiis = DIIS()
        self.history = []
        DE = None
        while err > self.con_tol and count < self.max_its:
            res, DE, t1_new, t2_new = self.updated_amps()
            if self.diis:
                t1, t2 = self.run_diis(t1_new, t2_new, adiis)
            else:
                t1, t2 = t1_new, t2_new
            self.t1 = t1
            self.t2 = t2
            count += 1
            err = res
            self.history.append((count, DE, res))
            if self.verbose_print and self.rank == 0:
                print(f"It {count} Energy progress {DE:.6e} residual progress {res:.6e}", file=self.stdout)
        if self.verbose_print and self.rank == 0:
            print("In Kernel!", file=self.stdout)
        self.iterations = count
        self._adiis = adiis
        return DE


def to_gpu(cpu_lowlevel, **kwargs):
    """``lib.to_gpu``-style conversion (``pyscf/lib/misc.py:1508-1556``) of a constructed reference
    ``dfrmpcc_lowlevel.MPCC_LL``: see :meth:`MPCC_LL.from_cpu`."""
    return MPCC_LL.from_cpu(cpu_lowlevel, **kwargs)
