"""Closed-form input scenes on the host (numpy), identical to what ``lbm_initialize_arrays``
evaluates on the device.

``reference_scene`` is InitializeArrays (LB.c:551-606): channel walls on rows 0 and nx-1, the
barrier rows [20,220) x columns 100..104 (absolute indices, clipped to the lattice), fA at rest
with density rho0, fB zero.  ``cylinder_walls`` is BASELINE.json config 3's obstacle (the
reference has no cylinder; SURVEY.md section 8(d) fixes its integer definition).
"""
import numpy as np

_DT = {"dp": np.float64, "sp": np.float32}


def weights(dtype):
    """OMEGA0 / OMEGA14 / OMEGA58 formed in ``dtype`` like the REAL_T literals (LB.c:38-40)."""
    t = np.dtype(dtype).type
    return np.array([t(4) / t(9)] + [t(1) / t(9)] * 4 + [t(1) / t(36)] * 4, dtype=dtype)


def equilibrium_rest(nx, ny, prec="dp", rho0=0.1):
    """fA of LB.c:583-591: INITIALDENSITY*OMEGA_s at every site, walls included."""
    dt = _DT[prec]
    w = weights(dt)
    f = np.empty((9, nx, ny), dtype=dt)
    f[...] = (np.dtype(dt).type(rho0) * w)[:, None, None]
    return f


def channel_walls(nx, ny, barrier=True):
    walls = np.zeros((nx, ny), dtype=np.int8)
    if barrier:
        walls[20:min(220, nx), 100:min(105, ny)] = 1     # LB.c:567-573
    walls[0, :] = 1                                        # LB.c:575-578
    walls[nx - 1, :] = 1
    return walls


def cylinder_walls(nx, ny, ci, cj, r):
    """Channel rows plus the solid disc (i-ci)^2 + (j-cj)^2 <= r^2, integer arithmetic."""
    walls = channel_walls(nx, ny, barrier=False)
    i = np.arange(nx, dtype=np.int64)[:, None]
    j = np.arange(ny, dtype=np.int64)[None, :]
    walls[(i - ci) ** 2 + (j - cj) ** 2 <= r * r] = 1
    return walls


def reference_scene(nx, ny, prec="dp", rho0=0.1):
    return equilibrium_rest(nx, ny, prec, rho0), channel_walls(nx, ny)


def perturb_lcg(f, seed=12345, amplitude=0.05):
    """Multiply f by 1 + amplitude*u, u in [-1,1) from a 32-bit LCG -- the robustness input of
    SURVEY.md section 8(d).  Pure integer generator so every implementation sees identical bits."""
    n = f.size
    a, c = np.uint64(1664525), np.uint64(1013904223)
    x = np.empty(n, dtype=np.uint64)
    s = np.uint64(seed)
    # vectorised LCG: x_k = a^k s + c (a^k - 1)/(a - 1)  (mod 2^32), built by doubling
    x[0] = (a * s + c) & np.uint64(0xFFFFFFFF)
    filled, ak, ck = 1, a, c
    while filled < n:
        m = min(filled, n - filled)
        x[filled:filled + m] = (ak * x[:m] + ck) & np.uint64(0xFFFFFFFF)
        ck = (ak * ck + ck) & np.uint64(0xFFFFFFFF)
        ak = (ak * ak) & np.uint64(0xFFFFFFFF)
        filled += m
    u = (x.astype(np.float64) / 2.0 ** 31 - 1.0).reshape(f.shape)
    return (f.astype(np.float64) * (1.0 + amplitude * u)).astype(f.dtype)
