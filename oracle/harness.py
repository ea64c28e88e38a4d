"""TEST INFRASTRUCTURE ONLY -- ctypes access to the CPU oracle and to the compiled reference.

``Oracle``  wraps oracle/liblbm_oracle_{dp,sp}.so (our restatement, lbm_oracle.c).
``Ref``     wraps oracle/_ref/liblbref_{dp,sp}_{avx,sse,serial}.so, i.e. the reference's
            own functions (LB.c:85-123) compiled by oracle/build_ref.sh.

Both expose the same small surface so tests can swap one for the other:
``initialize()``, ``run(nsteps)``, ``reynolds()``, arrays ``fA``, ``fB``, ``walls`` in the
reference layout ``f[s][i][j]`` with row stride ``ldy`` = NYPADDED (LB.c:71-72).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this module.  The product (latticeboltzmann_b200) never does.
"""
import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

VECWIDTH = {("dp", "avx"): 4, ("dp", "sse"): 2, ("dp", "serial"): 1,
            ("sp", "avx"): 8, ("sp", "sse"): 4, ("sp", "serial"): 1}
DTYPE = {"dp": np.float64, "sp": np.float32}
CREAL = {"dp": ctypes.c_double, "sp": ctypes.c_float}


def padded_ny(ny, align):
    """NYPADDED, LB.c:71."""
    return align * ((ny - 1) // align) + align


def aligned_zeros(n, dtype, align=64):
    """Zero-filled 1-D array whose base address is ``align``-byte aligned (LB.c:138-140 use 32)."""
    item = np.dtype(dtype).itemsize
    raw = np.zeros(n * item + align, dtype=np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * item].view(dtype)


def build_oracle():
    """Compile oracle/liblbm_oracle_*.so (and oracle/_ref when the reference checkout exists)."""
    subprocess.run(["make", "-s", "-C", HERE, "liblbm_oracle_dp.so", "liblbm_oracle_sp.so"], check=True)
    subprocess.run(["sh", os.path.join(HERE, "build_ref.sh")], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


class _Base:
    def __init__(self, prec, nx, ny, vw, tau=0.7, accel=0.005, rho0=0.1):
        assert prec in ("dp", "sp")
        self.prec, self.nx, self.ny, self.vw = prec, int(nx), int(ny), int(vw)
        self.dtype = DTYPE[prec]
        self.ldy = padded_ny(self.ny, self.vw)
        self.tau, self.accel, self.rho0 = tau, accel, rho0
        n = self.nx * self.ldy
        self.fA = aligned_zeros(9 * n, self.dtype)
        self.fB = aligned_zeros(9 * n, self.dtype)
        self.walls = aligned_zeros(n, np.int8)

    # views -----------------------------------------------------------------
    def f_view(self, which="A"):
        """(9, nx, ny) view without the padding columns."""
        f = self.fA if which == "A" else self.fB
        return f.reshape(9, self.nx, self.ldy)[:, :, :self.ny]

    def walls_view(self):
        return self.walls.reshape(self.nx, self.ldy)[:, :self.ny]

    def set_state(self, fA, walls, fB=None):
        """Load a (9,nx,ny) state and an (nx,ny) 0/1 mask."""
        self.f_view("A")[...] = fA
        self.f_view("B")[...] = 0 if fB is None else fB
        self.walls_view()[...] = walls

    def _p(self, a):
        return a.ctypes.data_as(ctypes.c_void_p)


class Oracle(_Base):
    """Our CPU restatement (oracle/lbm_oracle.c)."""

    class Params(ctypes.Structure):
        pass

    def __init__(self, prec, nx, ny, vw=None, source_col=0, outflow=False, **kw):
        vw = VECWIDTH[(prec, "avx")] if vw is None else vw
        super().__init__(prec, nx, ny, vw, **kw)
        path = os.path.join(HERE, f"liblbm_oracle_{prec}.so")
        if not os.path.exists(path):
            build_oracle()
        self.lib = ctypes.CDLL(path)
        real = CREAL[prec]

        class P(ctypes.Structure):
            _fields_ = [("nx", ctypes.c_int), ("ny", ctypes.c_int), ("ldy", ctypes.c_int),
                        ("vw", ctypes.c_int), ("tau", real), ("accel", real), ("rho0", real), ("source_col", ctypes.c_int),
                        ("outflow", ctypes.c_int)]
        self.p = P(self.nx, self.ny, self.ldy, self.vw, self.tau, self.accel, self.rho0, int(source_col), int(bool(outflow)))
        self.lib.lbo_compute_reynolds.restype = real
        assert self.lib.lbo_sizeof_real() == np.dtype(self.dtype).itemsize

    def initialize(self):
        self.lib.lbo_initialize_arrays(ctypes.byref(self.p), self._p(self.fA), self._p(self.fB), self._p(self.walls))

    def run(self, nsteps):
        assert nsteps % 2 == 0
        self.lib.lbo_run(ctypes.byref(self.p), self._p(self.fA), self._p(self.fB), self._p(self.walls), int(nsteps))

    def apply_source(self, which="A"):
        f = self.fA if which == "A" else self.fB
        self.lib.lbo_apply_source(ctypes.byref(self.p), self._p(f), self._p(self.walls))

    def stream_collide(self, src="A"):
        """One pull step src -> other buffer WITHOUT the source term (LB.c:195-196)."""
        a, b = (self.fA, self.fB) if src == "A" else (self.fB, self.fA)
        nyvec = self.vw * (self.ny // self.vw) if self.vw > 1 else 0
        if nyvec:
            self.lib.lbo_stream_collide_vec(ctypes.byref(self.p), 0, nyvec, self._p(a), self._p(b), self._p(self.walls))
        self.lib.lbo_stream_collide(ctypes.byref(self.p), nyvec, self.ny, self._p(a), self._p(b), self._p(self.walls))

    def reynolds(self):
        return float(self.lib.lbo_compute_reynolds(ctypes.byref(self.p), self._p(self.fA), self._p(self.walls)))

    def macroscopic(self):
        n = self.nx * self.ny
        rho, ux, uy = (np.zeros(n, self.dtype) for _ in range(3))
        self.lib.lbo_macroscopic(ctypes.byref(self.p), self._p(self.fA), self._p(rho), self._p(ux), self._p(uy))
        return tuple(a.reshape(self.nx, self.ny) for a in (rho, ux, uy))

    def print_lattice(self, path):
        return self.lib.lbo_print_lattice(ctypes.byref(self.p), self._p(self.fA), path.encode())


def ref_available(prec="dp", isa="avx"):
    return os.path.exists(os.path.join(REF_DIR, f"liblbref_{prec}_{isa}.so"))


class Ref(_Base):
    """The reference's own compiled functions (strict -O2 -ffp-contract=off build)."""

    def __init__(self, prec, nx, ny, isa="avx", **kw):
        super().__init__(prec, nx, ny, VECWIDTH[(prec, isa)], **kw)
        self.isa = isa
        self.lib = ctypes.CDLL(os.path.join(REF_DIR, f"liblbref_{prec}_{isa}.so"))
        real = CREAL[prec]
        ctypes.c_int.in_dll(self.lib, "lbref_nx").value = self.nx
        ctypes.c_int.in_dll(self.lib, "lbref_ny").value = self.ny
        real.in_dll(self.lib, "lbref_tau").value = self.tau
        real.in_dll(self.lib, "lbref_accel").value = self.accel
        real.in_dll(self.lib, "lbref_rho0").value = self.rho0
        self.lib.ComputeReynolds.restype = real
        if nx * self.ldy * 9 >= 2 ** 31:
            raise ValueError("reference indexes with int (LB.c:72): lattice too large")

    def initialize(self):
        if self.nx < 220 or self.ny < 105:
            raise ValueError("InitializeArrays writes the barrier at absolute rows 20..219, cols 100..104 (LB.c:567-573)")
        self.lib.InitializeArrays(self._p(self.fA), self._p(self.fB), self._p(self.walls))

    def run(self, nsteps):
        assert nsteps % 2 == 0
        for _ in range(nsteps // 2):
            self.lib.DoTimeStep(self._p(self.fA), self._p(self.fB), self._p(self.walls))

    def apply_source(self, which="A"):
        f = self.fA if which == "A" else self.fB
        self.lib.ApplySource(self._p(f), self._p(self.walls))

    def reynolds(self):
        return float(self.lib.ComputeReynolds(self._p(self.fA), self._p(self.walls)))

    def print_lattice(self, time_step):
        """The reference's own PrintLattice (LB.c:610-639): writes data/<time_step>.csv under the CURRENT directory."""
        self.lib.PrintLattice(int(time_step), self._p(self.fA))
