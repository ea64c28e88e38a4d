"""ctypes binding of the C ABI in include/lbm_b200.h.

One library per precision (``liblbm_b200_dp.so`` / ``liblbm_b200_sp.so``), chosen the way the
reference chooses precision: by header at compile time (LB.c:59-65).  The libraries live
in-tree next to this file.  There is no fallback of any kind: a missing library or a missing
GPU raises.
"""
import ctypes
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
BLOB_BYTES = 256
DTYPE = {"dp": np.float64, "sp": np.float32}
CREAL = {"dp": ctypes.c_double, "sp": ctypes.c_float}

OK, ERR_ARG, ERR_CUDA, ERR_STATE, ERR_IO, ERR_FPE = range(6)


class LbmError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"lbm_b200 error {code}: {text}")
        self.code = code


def _structs(real):
    class Config(ctypes.Structure):
        _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int), ("tau", real), ("accel", real),
                    ("vecwidth", ctypes.c_int), ("device", ctypes.c_int),
                    ("row_lo", ctypes.c_int), ("row_hi", ctypes.c_int), ("inplace", ctypes.c_int), ("boundary_y", ctypes.c_int)]

    class Scene(ctypes.Structure):
        _fields_ = [("rho0", real), ("channel_walls", ctypes.c_int),
                    ("bar_i0", ctypes.c_int), ("bar_i1", ctypes.c_int),
                    ("bar_j0", ctypes.c_int), ("bar_j1", ctypes.c_int),
                    ("disc_ci", ctypes.c_longlong), ("disc_cj", ctypes.c_longlong), ("disc_r", ctypes.c_longlong),
                    ("perturb_seed", ctypes.c_uint), ("perturb_amp", real)]

    class SlabInfo(ctypes.Structure):
        _fields_ = [("nx_local", ctypes.c_int), ("ny", ctypes.c_int), ("row_lo", ctypes.c_int), ("row_hi", ctypes.c_int),
                    ("ldy", ctypes.c_size_t), ("plane", ctypes.c_size_t), ("halo_bytes", ctypes.c_size_t),
                    ("dev_f", ctypes.c_void_p * 2), ("dev_flags", ctypes.c_void_p)]

    class Tuning(ctypes.Structure):
        _fields_ = [("vec", ctypes.c_int), ("hints", ctypes.c_int), ("pdl", ctypes.c_int),
                    ("graph", ctypes.c_int), ("tb_depth", ctypes.c_int), ("tb_rows", ctypes.c_int), ("tb_cols", ctypes.c_int),
                    ("tb_chains", ctypes.c_int)]

    class Experimental(ctypes.Structure):      # include/lbm_b200_experimental.h (experiment build only)
        _fields_ = [("stage_smem", ctypes.c_int), ("rows_per_thread", ctypes.c_int), ("block", ctypes.c_int), ("smem_pad", ctypes.c_int)]

    return Config, Scene, SlabInfo, Tuning, Experimental


# every symbol include/lbm_b200.h declares (tests check the library exports each one)
SYMBOLS = [
    "lbm_sizeof_real", "lbm_last_error", "lbm_version", "lbm_create", "lbm_destroy",
    "lbm_scene_reference", "lbm_initialize_arrays", "lbm_upload", "lbm_download", "lbm_upload_slab", "lbm_download_slab", "lbm_download_window", "lbm_set_walls", "lbm_load_walls",
    "lbm_do_timestep", "lbm_run", "lbm_run_timed", "lbm_apply_source", "lbm_stream_collide", "lbm_sync",
    "lbm_compute_reynolds", "lbm_reynolds_column", "lbm_reynolds_column_static", "lbm_reynolds_from_column", "lbm_macroscopic",
    "lbm_print_lattice", "lbm_print_lattice_group", "lbm_checksum", "lbm_check_finite", "lbm_slab_info_get", "lbm_ipc_export", "lbm_ipc_connect", "lbm_connect_local",
    "lbm_group_run", "lbm_halo_pack", "lbm_halo_unpack", "lbm_step_rows", "lbm_stream_handle",
    "lbm_stream_join", "lbm_set_tuning", "lbm_get_tuning", "lbm_kernel_launches", "lbm_tb_schedule", "lbm_tb_schedule_chains",
]


class Lib:
    """A loaded precision variant of the library with typed prototypes."""

    def __init__(self, prec, experimental=False):
        if prec not in ("dp", "sp"):
            raise ValueError("prec must be 'dp' or 'sp'")
        self.prec, self.experimental = prec, bool(experimental)
        # experimental: the experiment build of the SAME library (exp/, make experimental: + the shared-memory staged kernels
        # behind lbm_set_experimental); LBM_B200_LIBDIR: one-off experiment builds for tools/.  There is still no fallback.
        libdir = os.path.join(HERE, "exp") if experimental else os.environ.get("LBM_B200_LIBDIR", HERE)
        self.path = os.path.join(libdir, f"liblbm_b200_{prec}.so")
        if not os.path.exists(self.path):
            raise ImportError(f"{self.path} is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(there is no CPU fallback)")
        self.dll = ctypes.CDLL(self.path)
        self.real = CREAL[prec]
        self.dtype = DTYPE[prec]
        self.Config, self.Scene, self.SlabInfo, self.Tuning, self.Experimental = _structs(self.real)
        d, vp, ci = self.dll, ctypes.c_void_p, ctypes.c_int
        d.lbm_last_error.restype = ctypes.c_char_p
        d.lbm_version.restype = ctypes.c_char_p
        d.lbm_scene_reference.restype = self.Scene
        d.lbm_reynolds_from_column.restype = self.real
        d.lbm_reynolds_from_column.argtypes = [vp, vp, ci, self.real]
        d.lbm_kernel_launches.restype = ctypes.c_longlong
        d.lbm_kernel_launches.argtypes = [vp]
        d.lbm_create.argtypes = [ctypes.POINTER(vp), ctypes.POINTER(self.Config)]
        d.lbm_tb_schedule.restype = ci
        d.lbm_tb_schedule.argtypes = [ci, ci, ci, ci, ci, ci, ctypes.POINTER(ci), vp, ci]
        d.lbm_tb_schedule_chains.restype = ci
        d.lbm_tb_schedule_chains.argtypes = [ci, ci, ci, ci, ci, ci, ci, ctypes.POINTER(ci), ctypes.POINTER(ci), vp, ci]
        for name, args in {
            "lbm_destroy": [vp], "lbm_initialize_arrays": [vp, ctypes.POINTER(self.Scene)],
            "lbm_upload": [vp, vp, vp, vp, ci], "lbm_download": [vp, vp, vp, vp, ci],
            "lbm_upload_slab": [vp, vp, vp, vp, ci], "lbm_download_slab": [vp, vp, vp, vp, ci],
            "lbm_download_window": [vp, ci, ci, ci, ci, ci, vp],
            "lbm_set_walls": [vp, vp, ci], "lbm_load_walls": [vp, ctypes.c_char_p],
            "lbm_do_timestep": [vp], "lbm_run": [vp, ci], "lbm_run_timed": [vp, ci, ctypes.POINTER(ctypes.c_float)],
            "lbm_apply_source": [vp, ci], "lbm_stream_collide": [vp, ci], "lbm_sync": [vp],
            "lbm_compute_reynolds": [vp, ctypes.POINTER(self.real)], "lbm_reynolds_column": [vp, vp, vp], "lbm_reynolds_column_static": [vp, vp, vp],
            "lbm_macroscopic": [vp, vp, vp, vp], "lbm_print_lattice": [vp, ctypes.c_char_p],
            "lbm_checksum": [vp, ci, ctypes.POINTER(ctypes.c_ulonglong)], "lbm_check_finite": [vp, ctypes.POINTER(ctypes.c_ulonglong)],
            "lbm_slab_info_get": [vp, ctypes.POINTER(self.SlabInfo)], "lbm_ipc_export": [vp, vp],
            "lbm_ipc_connect": [vp, vp, vp], "lbm_connect_local": [vp, vp, vp],
            "lbm_group_run": [ctypes.POINTER(vp), ci, ci], "lbm_print_lattice_group": [ctypes.POINTER(vp), ci, ctypes.c_char_p], "lbm_halo_pack": [vp, ci, vp, vp],
            "lbm_halo_unpack": [vp, ci, vp, vp], "lbm_step_rows": [vp, ci, ci, ci, ci, ci],
            "lbm_stream_handle": [vp, ci, ctypes.POINTER(vp)], "lbm_stream_join": [vp],
            "lbm_set_tuning": [vp, ctypes.POINTER(self.Tuning)], "lbm_get_tuning": [vp, ctypes.POINTER(self.Tuning)],
        }.items():
            fn = getattr(d, name)
            fn.argtypes = args
            fn.restype = ci
        if self.experimental:
            d.lbm_set_experimental.argtypes = [vp, ctypes.POINTER(self.Experimental)]
            d.lbm_set_experimental.restype = ci
        if d.lbm_sizeof_real() != np.dtype(self.dtype).itemsize:
            raise ImportError(f"{self.path}: precision mismatch")

    def check(self, rc):
        if rc != OK:
            raise LbmError(rc, self.dll.lbm_last_error().decode())


_LIBS = {}


def load(prec, experimental=False):
    key = (prec, bool(experimental))
    if key not in _LIBS:
        _LIBS[key] = Lib(prec, experimental)
    return _LIBS[key]
