"""Host-side mirror of the reference's function surface (LB.c:85-123) over the C ABI.

``Lattice`` owns one ``lbm_ctx`` (one slab on one GPU).  Method names follow the reference:
``InitializeArrays``, ``DoTimeStep``, ``ApplySource``, ``StreamCollide``, ``ComputeReynolds``,
``PrintLattice``; host arrays use the reference layout ``f[s][i][j]`` (LB.c:71-72).  Everything
here is plumbing -- all arithmetic happens in the CUDA kernels behind ``liblbm_b200_*.so``.
"""
import ctypes

import numpy as np

from . import _capi


def _ptr(a):
    return None if a is None else a.ctypes.data_as(ctypes.c_void_p)


class Lattice:
    def __init__(self, nx, ny, prec="dp", tau=0.7, accel=0.005, vecwidth=0, device=0, row_lo=0, row_hi=0, inplace=False,
                 outflow=False, experimental=False):
        """``inplace=True`` keeps ONE buffer and steps with the in-place (AA) kernels: buffer A after every
        even step count is bit-identical, buffer B does not exist (lbm_config.inplace).  ``outflow=True``: open y ends
        (lbm_config.boundary_y = LBM_Y_OUTFLOW).  ``experimental=True`` loads the experiment build (exp/)."""
        self.lib = _capi.load(prec, experimental)
        self.inplace = bool(inplace)
        self.prec, self.dtype = prec, self.lib.dtype
        self.nx, self.ny = int(nx), int(ny)
        cfg = self.lib.Config(self.nx, self.ny, tau, accel, vecwidth, device, row_lo, row_hi, int(self.inplace), int(bool(outflow)))
        self._ctx = ctypes.c_void_p()
        self.lib.check(self.lib.dll.lbm_create(ctypes.byref(self._ctx), ctypes.byref(cfg)))
        self.tau = cfg.tau
        info = self.slab_info()
        self.row_lo, self.row_hi, self.nx_local = info.row_lo, info.row_hi, info.nx_local

    # ---- lifetime -------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) and self._ctx.value:
            self.lib.dll.lbm_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _call(self, name, *args):
        self.lib.check(getattr(self.lib.dll, name)(self._ctx, *args))

    # ---- inputs -----------------------------------------------------------------------------
    def scene_reference(self):
        return self.lib.dll.lbm_scene_reference()

    def InitializeArrays(self, scene=None, **kw):
        """InitializeArrays (LB.c:551-606) on the device; ``kw`` overrides scene fields
        (rho0, channel_walls, bar_i0.., disc_ci, disc_cj, disc_r)."""
        sc = scene if scene is not None else self.scene_reference()
        for k, v in kw.items():
            setattr(sc, k, v)
        self._call("lbm_initialize_arrays", ctypes.byref(sc))

    def upload(self, fA, walls, fB=None, ldy=None):
        """Global-lattice host arrays: fA/fB (9,nx,ldy) or flat, walls (nx,ldy) int8 0/1."""
        ldy = self.ny if ldy is None else ldy
        fA = np.ascontiguousarray(fA, dtype=self.dtype)
        walls = np.ascontiguousarray(walls, dtype=np.int8)
        fB = None if fB is None else np.ascontiguousarray(fB, dtype=self.dtype)
        assert fA.size == 9 * self.nx * ldy and walls.size == self.nx * ldy
        self._call("lbm_upload", _ptr(fA), _ptr(fB), _ptr(walls), ldy)

    def set_walls(self, walls, ldy=None):
        """The wall mask alone (global layout): 0 fluid, 1 bounce-back, 2 / 3 reflect with normal x / y (lbm_set_walls)."""
        ldy = self.ny if ldy is None else ldy
        walls = np.ascontiguousarray(walls, dtype=np.int8)
        assert walls.size == self.nx * ldy
        self._call("lbm_set_walls", _ptr(walls), ldy)

    def load_walls(self, path):
        """Geometry from a text file (lbm_load_walls): one lattice row per line, entries 0..3."""
        self._call("lbm_load_walls", str(path).encode())

    def download(self, which="A", out=None, ldy=None):
        """Returns (9,nx,ldy); rows outside this slab are left as they were in ``out`` (zeros if new)."""
        ldy = self.ny if ldy is None else ldy
        if out is None:
            out = np.zeros((9, self.nx, ldy), dtype=self.dtype)
        a, b = (_ptr(out), None) if which == "A" else (None, _ptr(out))
        self._call("lbm_download", a, b, None, ldy)
        return out

    def upload_slab(self, fA, walls, fB=None):
        """Host arrays holding only this slab's rows: fA (9,nx_local,ny), walls (nx_local,ny)."""
        assert fA.flags.c_contiguous and fA.dtype == self.dtype and fA.size == 9 * self.nx_local * self.ny
        walls = np.ascontiguousarray(walls, dtype=np.int8)
        self._call("lbm_upload_slab", _ptr(fA), _ptr(fB), _ptr(walls), self.ny)

    def download_slab(self, out=None, which="A", walls=False):
        if walls:
            w = np.zeros((self.nx_local, self.ny), dtype=np.int8)
            self._call("lbm_download_slab", None, None, _ptr(w), self.ny)
            return w
        if out is None:
            out = np.zeros((9, self.nx_local, self.ny), dtype=self.dtype)
        a, b = (_ptr(out), None) if which == "A" else (None, _ptr(out))
        self._call("lbm_download_slab", a, b, None, self.ny)
        return out

    def download_window(self, r0, r1, j0, j1, which="A"):
        """(9, r1-r0, j1-j0): local rows [r0,r1) x columns [j0,j1) of this slab (lbm_download_window)."""
        out = np.zeros((9, r1 - r0, j1 - j0), dtype=self.dtype)
        self._call("lbm_download_window", 0 if which == "A" else 1, int(r0), int(r1), int(j0), int(j1), _ptr(out))
        return out

    def download_walls(self, ldy=None):
        ldy = self.ny if ldy is None else ldy
        out = np.zeros((self.nx, ldy), dtype=np.int8)
        self._call("lbm_download", None, None, _ptr(out), ldy)
        return out

    # ---- the timestep path --------------------------------------------------------------------
    def DoTimeStep(self):
        self._call("lbm_do_timestep")

    def run(self, nsteps):
        self._call("lbm_run", int(nsteps))

    def run_timed(self, nsteps):
        ms = ctypes.c_float()
        self._call("lbm_run_timed", int(nsteps), ctypes.byref(ms))
        return ms.value

    def ApplySource(self, which="A"):
        self._call("lbm_apply_source", 0 if which == "A" else 1)

    def StreamCollide(self, src="A"):
        self._call("lbm_stream_collide", 0 if src == "A" else 1)

    def sync(self):
        self._call("lbm_sync")

    # ---- observers ------------------------------------------------------------------------------
    def ComputeReynolds(self):
        re = self.lib.real()
        self._call("lbm_compute_reynolds", ctypes.byref(re))
        return float(re.value)

    def reynolds_column(self):
        uy = np.zeros(self.nx_local, dtype=self.dtype)
        fluid = np.zeros(self.nx_local, dtype=np.int8)
        self._call("lbm_reynolds_column", _ptr(uy), _ptr(fluid))
        return uy, fluid

    def reynolds_column_static(self, uy=None, fluid=None):
        """lbm_reynolds_column_static: the same column without any allocation on the library's side (static device scratch,
        at most LBM_STATIC_MAX_ROWS local rows); pass ``uy`` / ``fluid`` to reuse the caller's buffers."""
        uy = np.zeros(self.nx_local, dtype=self.dtype) if uy is None else uy
        fluid = np.zeros(self.nx_local, dtype=np.int8) if fluid is None else fluid
        self._call("lbm_reynolds_column_static", _ptr(uy), _ptr(fluid))
        return uy, fluid

    def reynolds_from_column(self, uy, fluid):
        uy = np.ascontiguousarray(uy, dtype=self.dtype)
        fluid = np.ascontiguousarray(fluid, dtype=np.int8)
        return float(self.lib.dll.lbm_reynolds_from_column(_ptr(uy), _ptr(fluid), len(uy), self.tau))

    def macroscopic(self):
        shape = (self.nx_local, self.ny)
        rho, ux, uy = (np.zeros(shape, dtype=self.dtype) for _ in range(3))
        self._call("lbm_macroscopic", _ptr(rho), _ptr(ux), _ptr(uy))
        return rho, ux, uy

    def PrintLattice(self, path):
        self._call("lbm_print_lattice", str(path).encode())

    def checksum(self, which="A"):
        """Decomposition-invariant 64-bit checksum of this slab's part of a buffer (lbm_checksum); the slabs of a ring
        add up (mod 2**64) to the single-GPU value.  geometry.checksum is the numpy statement of the same sum."""
        out = ctypes.c_ulonglong()
        self._call("lbm_checksum", 0 if which == "A" else 1, ctypes.byref(out))
        return int(out.value)

    def check_finite(self):
        """Raises LbmError(ERR_FPE) if buffer A holds NaN/Inf (the reference traps instead, LB.c:129)."""
        n = ctypes.c_ulonglong()
        self._call("lbm_check_finite", ctypes.byref(n))

    # ---- slabs / tuning ----------------------------------------------------------------------------
    def slab_info(self):
        info = self.lib.SlabInfo()
        self._call("lbm_slab_info_get", ctypes.byref(info))
        return info

    def ipc_export(self):
        blob = np.zeros(_capi.BLOB_BYTES, dtype=np.uint8)
        self._call("lbm_ipc_export", _ptr(blob))
        return blob

    def ipc_connect(self, up_blob, dn_blob):
        up = np.ascontiguousarray(up_blob, dtype=np.uint8)
        dn = np.ascontiguousarray(dn_blob, dtype=np.uint8)
        self._call("lbm_ipc_connect", _ptr(up), _ptr(dn))

    def connect_local(self, up, dn):
        self._call("lbm_connect_local", up._ctx, dn._ctx)

    def set_tuning(self, **kw):
        t = self.lib.Tuning()
        self._call("lbm_get_tuning", ctypes.byref(t))
        for k, v in kw.items():
            setattr(t, k, v)
        self._call("lbm_set_tuning", ctypes.byref(t))

    def set_experimental(self, **kw):
        """Experiment build only (include/lbm_b200_experimental.h): stage_smem, rows_per_thread, block, smem_pad."""
        x = self.lib.Experimental(**kw)
        self._call("lbm_set_experimental", ctypes.byref(x))

    def get_tuning(self):
        t = self.lib.Tuning()
        self._call("lbm_get_tuning", ctypes.byref(t))
        return {k: getattr(t, k) for k, _ in t._fields_}

    @property
    def launches(self):
        return int(self.lib.dll.lbm_kernel_launches(self._ctx))


def group_run(slabs, nsteps):
    """Step a ring of in-process slabs together (lbm_group_run)."""
    lib = slabs[0].lib
    arr = (ctypes.c_void_p * len(slabs))(*[s._ctx.value for s in slabs])
    lib.check(lib.dll.lbm_group_run(arr, len(slabs), int(nsteps)))


def print_lattice_group(slabs, path):
    """PrintLattice (LB.c:610-639) of a lattice spread over in-process slabs: one CSV, rows in order."""
    lib = slabs[0].lib
    arr = (ctypes.c_void_p * len(slabs))(*[s._ctx.value for s in slabs])
    lib.check(lib.dll.lbm_print_lattice_group(arr, len(slabs), str(path).encode()))


def make_ring(nx, ny, nslabs, prec="dp", device=0, bounds=None, **kw):
    """Split rows [0,nx) into ``nslabs`` contiguous slabs on one device and wire them as a ring --
    decomposition-invariance tests without a cluster."""
    from .slab import partition_rows
    bounds = bounds or partition_rows(nx, nslabs)
    slabs = [Lattice(nx, ny, prec, device=device, row_lo=lo, row_hi=hi, **kw) for lo, hi in bounds]
    n = len(slabs)
    if n > 1:
        for k, s in enumerate(slabs):
            s.connect_local(slabs[(k + 1) % n], slabs[(k - 1) % n])
    return slabs


def tb_schedule(nx, ny, tb_depth, tb_rows=0, tb_cols=0, prec="dp", multi=False):
    """Launch list of one temporal-blocking macro-step (lbm_tb_schedule; host logic, no GPU):
    returns (depth_used, [(level, r0, r1, j0, j1), ...]); ranges past nx / ny wrap periodically,
    j0 == j1 == 0 means whole rows.  ``multi=True``: the list of one slab of a multi-slab ring, each entry
    with a sixth field ``seam`` (bit 0: wait for the neighbours' previous event first, bit 1: announce after)."""
    lib = _capi.load(prec)
    k = ctypes.c_int()
    mode = int(bool(multi))
    n = lib.dll.lbm_tb_schedule(nx, ny, mode, tb_depth, tb_rows, tb_cols, ctypes.byref(k), None, 0)
    if n < 0:
        raise _capi.LbmError(_capi.ERR_ARG, lib.dll.lbm_last_error().decode())
    buf = (ctypes.c_int * (6 * max(n, 1)))()
    lib.dll.lbm_tb_schedule(nx, ny, mode, tb_depth, tb_rows, tb_cols, ctypes.byref(k), buf, n)
    return k.value, [tuple(buf[6 * i + m] for m in range(6 if mode == 1 else 5)) for i in range(n)]


def tb_schedule_chains(nx, ny, chains, tb_depth, tb_rows=0, tb_cols=0, prec="dp", multi=False):
    """Launch list of one macro-step with concurrent tile chains (lbm_tb_schedule_chains; host logic, no GPU):
    returns (depth_used, chains_used, [(level, r0, r1, j0, j1, stream, outer), ...]) in enqueue order; streams 0..C-1 are the
    lanes, C..2C-1 the seam chains of the boundaries after lane (stream - C).  chains_used < 2: no lanes (see tb_schedule)."""
    lib = _capi.load(prec)
    k, cu = ctypes.c_int(), ctypes.c_int()
    n = lib.dll.lbm_tb_schedule_chains(nx, ny, int(bool(multi)), chains, tb_depth, tb_rows, tb_cols, ctypes.byref(k), ctypes.byref(cu), None, 0)
    if n < 0:
        raise _capi.LbmError(_capi.ERR_ARG, lib.dll.lbm_last_error().decode())
    buf = (ctypes.c_int * (7 * max(n, 1)))()
    lib.dll.lbm_tb_schedule_chains(nx, ny, int(bool(multi)), chains, tb_depth, tb_rows, tb_cols, ctypes.byref(k), ctypes.byref(cu), buf, n)
    return k.value, cu.value, [tuple(buf[7 * i + m] for m in range(7)) for i in range(n)]
