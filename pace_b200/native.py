"""NativeCore: owns the device buffers (PyTorch tensors) of all subdomains resident on one GPU and
drives libpace_b200 through the C ABI.

Layout (see csrc/common.cuh): 3-D fields are torch tensors of shape (nt, npz+1, ny+7, sj), fp64,
i fastest, i.e. the transpose of the reference's numpy-backend (i, j, k) storage
(gt4py storage/cartesian/layout.py:143-149) padded to the library's row stride.
"""
import ctypes

import numpy as np
import torch

from . import _lib

DEFAULTS = dict(
    n_split=1, k_split=1, hord_dp=6, hord_mt=6, hord_tm=6, hord_tr=8, hord_vt=6, kord_mt=9, kord_tm=-9,
    kord_tr=9, kord_wz=9, nord=3, n_sponge=1, nf_omega=1, grid_type=0, do_sat_adj=1, do_vort_damp=1, rf_fast=1,
    hydrostatic=0, convert_ke=0, fill=1, adiabatic=0, do_qa=1, use_old_omega=1, do_f3d=0, inline_q=0, nwat=6,
    dt_atmos=225.0, a_imp=1.0, beta=0.0, p_fac=0.05, d_con=1.0, d2_bg=0.0, d2_bg_k1=0.2, d2_bg_k2=0.1,
    d4_bg=0.15, dddmp=0.5, delt_max=0.002, ke_bg=0.0, rf_cutoff=3000.0, tau=10.0, vtdm4=0.06, d_ext=0.0,
    consv_te=0.0, ptop=300.0, da_min=0.0, da_min_c=0.0, ks=18,
)


class NativeError(RuntimeError):
    pass


class NativeCore:
    def __init__(self, nx, ny, npz, n_tiles=1, edge_flags=None, device="cuda:0", _lib_handle=None, **cfg):
        # _lib_handle: dependency injection for tests/hostemu.py only (kernel sources compiled for the
        # host to debug logic without a GPU); the product always loads the CUDA library.
        self.lib = _lib_handle if _lib_handle is not None else _lib.load()
        self.device = torch.device(device)
        if _lib_handle is None:
            if not torch.cuda.is_available() or self.device.type != "cuda":
                raise NativeError("pace_b200 needs a CUDA device (there is no CPU fallback)")
            torch.cuda.set_device(self.device)
        self.nx, self.ny, self.npz, self.nt = nx, ny, npz, n_tiles
        self.sj = self.lib.pace_b200_row_stride(nx)
        self.nj = ny + 7
        c = _lib.Config()
        vals = dict(DEFAULTS)
        vals.update(cfg)
        vals.update(nx=nx, ny=ny, npz=npz, n_tiles_local=n_tiles)
        vals.setdefault("npx", nx + 1)
        vals.setdefault("npy", ny + 1)
        for name, ctype in _lib.Config._fields_:
            v = vals[name]
            setattr(c, name, int(v) if ctype is ctypes.c_int else float(v))
        self.cfg = vals
        flags = bytes(edge_flags if edge_flags is not None else [15] * n_tiles)
        self.handle = ctypes.c_void_p()
        rc = self.lib.pace_b200_create(ctypes.byref(c), flags, ctypes.byref(self.handle))
        if rc != 0:
            raise NativeError(self.lib.pace_b200_last_error(None).decode())
        self.t = {}  # name -> torch tensor
        if self.device.type == "cuda":
            self.lib.pace_b200_set_stream(self.handle, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))

    def __del__(self):
        h = getattr(self, "handle", None)
        if h:
            self.lib.pace_b200_destroy(h)
            self.handle = None

    # -- buffers ---------------------------------------------------------------------------------
    def _register(self, name, tensor):
        self.t[name] = tensor
        self.lib.pace_b200_set_field(self.handle, name.encode(), ctypes.c_void_p(tensor.data_ptr()))
        return tensor

    def alloc3(self, name, fill=0.0):
        return self._register(
            name, torch.full((self.nt, self.npz + 1, self.nj, self.sj), fill, dtype=torch.float64, device=self.device)
        )

    def alloc2(self, name):
        return self._register(name, torch.zeros((self.nt, self.nj, self.sj), dtype=torch.float64, device=self.device))

    def allock(self, name, n=None):
        return self._register(name, torch.zeros((n or self.npz + 1,), dtype=torch.float64, device=self.device))

    def put(self, name, arrays):
        """arrays: per-subdomain list of reference-layout numpy arrays (i,j,k) / (i,j) / (k,)."""
        a0 = np.asarray(arrays[0])
        if a0.ndim == 1:
            t = self.t.get(name)
            if t is None:
                t = self.allock(name, a0.shape[0])
            t.copy_(torch.from_numpy(np.ascontiguousarray(a0, dtype=np.float64)))
            return t
        t = self.t.get(name)
        if t is None:
            t = self.alloc3(name) if a0.ndim == 3 else self.alloc2(name)
        host = np.zeros(tuple(t.shape), dtype=np.float64)
        for n, a in enumerate(arrays):
            a = np.asarray(a, dtype=np.float64)
            if a.ndim == 3:
                ni, nj, nk = a.shape
                host[n, :nk, :nj, :ni] = a.transpose(2, 1, 0)
            else:
                ni, nj = a.shape
                host[n, :nj, :ni] = a.T
        t.copy_(torch.from_numpy(host))
        return t

    def get(self, name):
        """Back to the reference layout: list over subdomains of (i,j,k) / (i,j) arrays."""
        t = self.t[name].cpu().numpy()
        ni = self.nx + 7
        if t.ndim == 4:
            return [np.ascontiguousarray(t[n, :, :, :ni].transpose(2, 1, 0)) for n in range(self.nt)]
        if t.ndim == 3:
            return [np.ascontiguousarray(t[n, :, :ni].T) for n in range(self.nt)]
        return [t]

    # -- single-precision fields (the reference's PACE_FLOAT_PRECISION=32 mode; MapSingle only) ---------------
    def put32(self, arrays):
        """float32 device tensor in the library layout from per-subdomain reference-layout (i,j,k) / (i,j) arrays."""
        a0 = np.asarray(arrays[0])
        shape = (self.nt, self.npz + 1, self.nj, self.sj) if a0.ndim == 3 else (self.nt, self.nj, self.sj)
        host = np.zeros(shape, dtype=np.float32)
        for n, a in enumerate(arrays):
            a = np.asarray(a, dtype=np.float32)
            if a.ndim == 3:
                ni, nj, nk = a.shape
                host[n, :nk, :nj, :ni] = a.transpose(2, 1, 0)
            else:
                ni, nj = a.shape
                host[n, :nj, :ni] = a.T
        return torch.from_numpy(host).to(self.device)

    def get32(self, t):
        t = t.cpu().numpy()
        ni = self.nx + 7
        if t.ndim == 4:
            return [np.ascontiguousarray(t[n, :, :, :ni].transpose(2, 1, 0)) for n in range(self.nt)]
        return [np.ascontiguousarray(t[n, :, :ni].T) for n in range(self.nt)]

    def map_single_f32(self, q, pe1, pe2, qs=None, iv=1, kord=9, x_interface=False, y_interface=False, qmin=0.0):
        """MapSingle.__call__ on float32 tensors in the library layout (q is remapped in place)."""
        for t in (q, pe1, pe2) + ((qs,) if qs is not None else ()):
            if t.dtype != torch.float32 or not t.is_contiguous():
                raise NativeError("map_single_f32 needs contiguous float32 tensors")
        rc = self.lib.pace_b200_map_single_f32(
            self.handle, ctypes.c_void_p(q.data_ptr()), ctypes.c_void_p(pe1.data_ptr()), ctypes.c_void_p(pe2.data_ptr()),
            ctypes.c_void_p(qs.data_ptr()) if qs is not None else None, int(iv), int(kord), int(x_interface),
            int(y_interface), float(qmin))
        if rc != 0:
            raise NativeError("map_single_f32: %s" % self.lib.pace_b200_last_error(self.handle).decode())

    def set_column(self, name, values):
        """Small per-level column (host values; the library keeps host + device copies)."""
        v = np.ascontiguousarray(values, dtype=np.float64)
        rc = self.lib.pace_b200_set_column(self.handle, name.encode(), v.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), v.size)
        if rc != 0:
            raise NativeError("set_column %s: %s" % (name, self.lib.pace_b200_last_error(self.handle).decode()))

    def set_columns_from_config(self):
        from .hostprep import all_columns

        for name, v in all_columns(self.cfg, self.npz, self.cfg["da_min"], self.cfg["da_min_c"]).items():
            self.set_column(name, v)

    def put_grid(self, grids):
        """grids: list over subdomains of dict name -> reference-layout array ("grid.*", "damp.*")."""
        from .hostprep import cubic_spline_constants, derived_grid_fields

        for name, v in cubic_spline_constants(grids[0]["grid.dp_ref"], self.npz).items():
            self.set_column(name, v)

        g0 = grids[0]
        for name in g0:
            if np.ndim(g0[name]) == 0 or name in ("grid.edge_s", "grid.edge_n"):
                continue
            if np.ndim(g0[name]) == 1:
                self.set_column(name, g0[name])
            elif np.ndim(g0[name]) == 2:
                self.put(name, [g[name] for g in grids])
        for name, arrs in derived_grid_fields(grids, self.nx, self.ny).items():
            if name == "grid.a2b_corner_coef":
                self.set_column(name, arrs[0])
            else:
                self.put(name, arrs)

    # -- halo exchange ---------------------------------------------------------------------------
    def set_halo_plan(self, plan_id, ops, rank_to_local):
        """ops: list of pace_b200.halo.HaloOp whose dst and src subdomains are both resident here."""
        arr = (_lib.HaloOp * max(1, len(ops)))()
        for n, o in enumerate(ops):
            arr[n] = _lib.HaloOp(rank_to_local[o.dst_rank], rank_to_local[o.src_rank], o.comp, o.src_comp, o.sign,
                                 o.i0, o.j0, o.ni, o.nj, o.aii, o.aij, o.bi, o.aji, o.ajj, o.bj)
        rc = self.lib.pace_b200_halo_plan_set(self.handle, plan_id, ctypes.cast(arr, ctypes.c_void_p), len(ops))
        if rc != 0:
            raise NativeError("halo_plan_set: %s" % self.lib.pace_b200_last_error(self.handle).decode())

    def set_halo_plan_raw(self, plan_id, rows):
        """rows: (dst_t, src_t, HaloOp) with the two subdomain slots given explicitly — for pack / unpack plans one
        of them is the op's cell offset in the message (include/pace_b200.h, pace_b200_halo_pack)."""
        arr = (_lib.HaloOp * max(1, len(rows)))()
        for n, (dst_t, src_t, o) in enumerate(rows):
            arr[n] = _lib.HaloOp(dst_t, src_t, o.comp, o.src_comp, o.sign, o.i0, o.j0, o.ni, o.nj, o.aii, o.aij, o.bi,
                                 o.aji, o.ajj, o.bj)
        rc = self.lib.pace_b200_halo_plan_set(self.handle, plan_id, ctypes.cast(arr, ctypes.c_void_p), len(rows))
        if rc != 0:
            raise NativeError("halo_plan_set: %s" % self.lib.pace_b200_last_error(self.handle).decode())

    def _halo_msg(self, fn, what, plan_id, x_names, y_names, nk, buf):
        n = len(x_names)
        xs = (ctypes.c_char_p * n)(*[s.encode() for s in x_names])
        ys = (ctypes.c_char_p * n)(*[s.encode() for s in y_names]) if y_names else None
        ptr = buf if isinstance(buf, int) else buf.data_ptr()  # a raw device address (peer memory) or a tensor
        rc = fn(self.handle, plan_id, xs, ys, n, nk, ctypes.c_void_p(ptr))
        if rc != 0:
            raise NativeError("%s: %s" % (what, self.lib.pace_b200_last_error(self.handle).decode()))

    def halo_pack(self, plan_id, x_names, y_names, nk, buf):
        self._halo_msg(self.lib.pace_b200_halo_pack, "halo_pack", plan_id, x_names, y_names, nk, buf)

    def halo_unpack(self, plan_id, x_names, y_names, nk, buf):
        self._halo_msg(self.lib.pace_b200_halo_unpack, "halo_unpack", plan_id, x_names, y_names, nk, buf)

    def halo_update(self, plan_id, x_names, y_names=None, nk=None):
        n = len(x_names)
        xs = (ctypes.c_char_p * n)(*[s.encode() for s in x_names])
        ys = (ctypes.c_char_p * n)(*[s.encode() for s in y_names]) if y_names else None
        rc = self.lib.pace_b200_halo_update(self.handle, plan_id, xs, ys, n, self.npz if nk is None else nk)
        if rc != 0:
            raise NativeError("halo_update: %s" % self.lib.pace_b200_last_error(self.handle).decode())

    def sat_adjust(self, mdt, last_step, kmp, **over):
        cfg = _lib.SatAdjConfig.make(**over)
        rc = self.lib.pace_b200_sat_adjust(self.handle, ctypes.byref(cfg), float(mdt), int(last_step), int(kmp))
        if rc != 0:
            raise NativeError("sat_adjust: %s" % self.lib.pace_b200_last_error(self.handle).decode())

    # -- calls -----------------------------------------------------------------------------------
    def call(self, name, *scalars):
        fn = getattr(self.lib, "pace_b200_" + name)
        scalars = [a.encode() if isinstance(a, str) else a for a in scalars]
        rc = fn(self.handle, *scalars)
        if rc != 0:
            raise NativeError("%s: %s" % (name, self.lib.pace_b200_last_error(self.handle).decode()))

    def sync(self):
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
            if getattr(self, "peer_exchange", False):
                self.check_peers()

    def check_peers(self):
        """Raises if a flag barrier of the library-driven halo exchange gave up waiting for a peer GPU (a rank that
        failed or fell out of step): the halos of that update are then not valid."""
        done, bad = ctypes.c_ulonglong(0), ctypes.c_ulonglong(0)
        rc = self.lib.pace_b200_halo_peer_status(self.handle, ctypes.byref(done), ctypes.byref(bad))
        if rc != 0:
            raise NativeError("halo_peer_status: %s" % self.lib.pace_b200_last_error(self.handle).decode())
        if bad.value:
            raise NativeError("halo exchange: barrier %d of %d timed out waiting for a peer GPU" % (bad.value, done.value))
