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


# ------------------------------------------------------------------------------------------------
# numpy statements of two device-side closed forms (lbm_kernels.cuh): the perturbed rest state of
# lbm_scene.perturb_seed and the decomposition-invariant checksum of lbm_checksum.  Host-side mirrors
# for tests and tools; nothing on the product path calls them.
# ------------------------------------------------------------------------------------------------
def hash32(x):
    """lowbias32 on uint32 arrays (hash32 of lbm_kernels.cuh)."""
    x = np.asarray(x, dtype=np.uint32).copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint32(16)
        x *= np.uint32(0x7feb352d)
        x ^= x >> np.uint32(15)
        x *= np.uint32(0x846ca68b)
        x ^= x >> np.uint32(16)
    return x


def perturbed_rest(rows, cols, prec="dp", rho0=0.1, seed=12345, amp=0.1):
    """fA of scene_fill_kernel for the sites (rows x cols) given as GLOBAL row / column index arrays:
    rho = rho0*(1 + amp*(h/2^32 - 1/2)), f_s = rho*omega_s, each operation rounded once in the precision."""
    dt = _DT[prec]
    t = np.dtype(dt).type
    rows = np.asarray(rows, dtype=np.int64)
    cols = np.asarray(cols, dtype=np.int64)
    with np.errstate(over="ignore"):
        hj = hash32((cols.astype(np.uint32) + np.uint32(seed)).astype(np.uint32))
        hi = (rows.astype(np.uint32) * np.uint32(0x9E3779B1)).astype(np.uint32)
    h = hash32(hi[:, None] ^ hj[None, :])
    u = h.astype(dt) * t(1.0 / 4294967296.0)
    rho = t(rho0) * (t(1) + t(amp) * (u - t(0.5)))
    w = weights(dt)
    return (rho[None, :, :] * w[:, None, None]).astype(dt)


def scene_walls(nx, rows, cols, channel=True, bar=(0, 0, 0, 0), disc=None):
    """walls of scene_mask_kernel for a window given by GLOBAL row / column index arrays."""
    rows = np.asarray(rows, dtype=np.int64)[:, None]
    cols = np.asarray(cols, dtype=np.int64)[None, :]
    w = np.zeros((rows.shape[0], cols.shape[1]), dtype=np.int8)
    if channel:
        w[((rows == 0) | (rows == nx - 1)) & (cols == cols)] = 1
    i0, i1, j0, j1 = bar
    w[(rows >= i0) & (rows < i1) & (cols >= j0) & (cols < j1)] = 1
    if disc is not None:
        ci, cj, r = disc
        w[(rows - ci) ** 2 + (cols - cj) ** 2 <= r * r] = 1
    return w


def mix64(z):
    z = np.asarray(z, dtype=np.uint64).copy()
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xbf58476d1ce4e5b9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94d049bb133111eb)
    return z ^ (z >> np.uint64(31))


def checksum(f, nx_global=None, row_lo=0):
    """lbm_checksum of a (9, rows, ny) array whose first row is global row ``row_lo`` of an ``nx_global``-row lattice."""
    f = np.ascontiguousarray(f)
    _, rows, ny = f.shape
    nx_global = rows if nx_global is None else nx_global
    bits = f.view(np.uint64 if f.dtype == np.float64 else np.uint32).astype(np.uint64)
    s = np.arange(9, dtype=np.uint64)[:, None, None]
    i = (np.arange(rows, dtype=np.uint64) + np.uint64(row_lo))[None, :, None]
    j = np.arange(ny, dtype=np.uint64)[None, None, :]
    with np.errstate(over="ignore"):
        idx = (s * np.uint64(nx_global) + i) * np.uint64(ny) + j
        return int(np.sum(mix64(bits ^ mix64(idx)), dtype=np.uint64))
