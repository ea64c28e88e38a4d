"""GPU (-m gpu): parity AT BASELINE.json's full sizes -- the weak-scaling unit 8192x32768 (configs[4]) with the library's
own defaults (in place, 2-D temporal blocking, k = 16, 16/32-KB strips), i.e. the code path no small golden case reaches.

Two independent checks, both bit-exact (no tolerance):
  * WINDOWS AGAINST THE ORACLE.  D2Q9 information travels one site per step (LB.c:235-243), so after T steps the state of
    a window [a,b) x [c,d) depends only on the initial state of the window widened by T sites.  The widened window is cut
    out of the closed-form initial state (lbm_scene.perturb_seed: a rough density field a host can reproduce anywhere),
    stepped by the CPU oracle as a small periodic lattice, and its interior compared with the same sites of the GPU's
    full lattice.  Windows sit on the tile seams of the wavefront schedule, on slab seams, on the periodic wraps, on the
    source column and on the obstacle.
  * CROSS-ENGINE CHECKSUM.  The in-place + temporal-blocking engine and the A/B double-buffer engine (different kernels,
    different schedules) must leave the same 64-bit checksum of fA (lbm_checksum).
"""
import numpy as np
import pytest

from latticeboltzmann_b200 import geometry

pytestmark = pytest.mark.gpu
VW = {"dp": 4, "sp": 8}
SEED, AMP = 12345, 0.1


def _scene(nx, ny):
    return dict(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10, perturb_seed=SEED, perturb_amp=AMP)


def _fetch(slabs, rows, cols, ny):
    """(9, len(rows), len(cols)) of buffer A for GLOBAL index arrays that are contiguous modulo nx / ny."""
    out = np.zeros((9, len(rows), len(cols)), dtype=slabs[0].dtype)
    col_runs, k = [], 0
    while k < len(cols):                                     # runs of consecutive columns (at most two: the y wrap)
        e = k
        while e + 1 < len(cols) and cols[e + 1] == cols[e] + 1:
            e += 1
        col_runs.append((k, e + 1))
        k = e + 1
    for lat in slabs:
        sel = np.nonzero((rows >= lat.row_lo) & (rows < lat.row_hi))[0]
        k = 0
        while k < len(sel):
            e = k
            while e + 1 < len(sel) and sel[e + 1] == sel[e] + 1 and rows[sel[e + 1]] == rows[sel[e]] + 1:
                e += 1
            r0, r1 = rows[sel[k]] - lat.row_lo, rows[sel[e]] - lat.row_lo + 1
            for c0, c1 in col_runs:
                out[:, sel[k]:sel[e] + 1, c0:c1] = lat.download_window(int(r0), int(r1), int(cols[c0]), int(cols[c1 - 1]) + 1)
            k = e + 1
    return out


def _check_window(slabs, prec, nx, ny, T, a, b, c, d, what):
    """interior rows [a,b) x columns [c,d) (global, may wrap) after T steps == the oracle on the widened window"""
    from oracle.harness import Oracle
    M = T + 2
    c0 = (c - M) // 8 * 8                                    # keep the aligned VECWIDTH groups of LB.c:437-440 where they are
    d1 = c0 + -(-(d + M - c0) // 8) * 8
    rows = np.arange(a - M, b + M) % nx
    cols = np.arange(c0, d1) % ny
    assert len(rows) < nx and len(cols) < ny
    f0 = geometry.perturbed_rest(rows, cols, prec, seed=SEED, amp=AMP)
    walls = geometry.scene_walls(nx, rows, cols, disc=(nx // 2, ny // 5, nx // 10))
    src = np.nonzero(cols == 0)[0]
    o = Oracle(prec, len(rows), len(cols), VW[prec], source_col=int(src[0]) if len(src) else -1)
    o.set_state(f0, walls)
    o.run(T)
    want = o.f_view("A")[:, M:M + (b - a), (c - c0):(c - c0) + (d - c)]
    got = _fetch(slabs, rows[M:M + (b - a)], cols[(c - c0):(c - c0) + (d - c)], ny)
    if got.tobytes() != want.tobytes():
        bad = np.argwhere(got != want)
        raise AssertionError(f"{what}: window rows [{a},{b}) x cols [{c},{d}) after {T} steps: {len(bad)} of {got.size} values differ, "
                             f"first (s,i,j) = {bad[0].tolist()}: {got[tuple(bad[0])]!r} vs {want[tuple(bad[0])]!r}")
    assert np.ptp(want.astype(np.float64)) > 0


def _windows(nx, ny, row_seams, col_seams):
    """interiors of 48 rows x 128 columns: on the seams given, on both periodic wraps, on column 0 (ApplySource), on the disc"""
    H, W = 48, 128
    ci, cj, r = nx // 2, ny // 5, nx // 10
    w = [(-H // 2, H // 2, 4096, 4096 + W, "x wrap (rows nx-1 | 0)"), (nx // 3, nx // 3 + H, -W // 2, W // 2, "y wrap + source column"),
         (-H // 2, H // 2, -W // 2, W // 2, "corner: both wraps"), (ci - r - H // 2, ci - r + H // 2, cj - W // 2, cj + W // 2, "disc rim"),
         (ci - H // 2, ci + H // 2, cj + r - W // 2, cj + r + W // 2, "disc rim (downstream)")]
    for rs in row_seams:
        for cs in col_seams:
            w.append((rs - H // 2, rs + H // 2, cs - W // 2, cs + W // 2, f"tile seam row {rs} col {cs}"))
    return w


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_scene_perturbation_and_checksum_match_numpy(prec):
    from latticeboltzmann_b200 import Lattice, make_ring
    nx, ny = 61, 264
    with Lattice(nx, ny, prec) as lat:
        lat.InitializeArrays(bar_i0=3, bar_i1=9, bar_j0=10, bar_j1=40, disc_ci=30, disc_cj=100, disc_r=9, perturb_seed=SEED, perturb_amp=AMP)
        got = lat.download("A")
        want = geometry.perturbed_rest(np.arange(nx), np.arange(ny), prec, seed=SEED, amp=AMP)
        assert got.tobytes() == want.tobytes()
        assert len(np.unique(got[0])) > nx * ny // 2          # really rough
        walls = geometry.scene_walls(nx, np.arange(nx), np.arange(ny), bar=(3, 9, 10, 40), disc=(30, 100, 9))
        assert np.array_equal(lat.download_walls(), walls)
        assert lat.checksum() == geometry.checksum(got)
        lat.run(6)
        after = lat.download("A")
        whole = lat.checksum()
        assert whole == geometry.checksum(after) and whole != geometry.checksum(got)
        assert lat.checksum("B") == geometry.checksum(lat.download("B"))
        assert np.array_equal(lat.download_window(7, 19, 33, 64), after[:, 7:19, 33:64])
    slabs = make_ring(nx, ny, 3, prec)
    try:
        for s in slabs:
            s.InitializeArrays(bar_i0=3, bar_i1=9, bar_j0=10, bar_j1=40, disc_ci=30, disc_cj=100, disc_r=9, perturb_seed=SEED, perturb_amp=AMP)
        from latticeboltzmann_b200 import group_run
        group_run(slabs, 6)
        assert sum(s.checksum() for s in slabs) % 2 ** 64 == whole     # the slabs of a ring add up to the single-GPU value
    finally:
        for s in slabs:
            s.close()


def test_check_finite_stands_in_for_the_fp_traps():
    from latticeboltzmann_b200 import Lattice, _capi
    fA, walls = geometry.reference_scene(40, 64, "dp")
    with Lattice(40, 64, "dp") as lat:
        lat.upload(fA, walls)
        lat.run(4)
        lat.check_finite()
        fA[3, 7, 9] = np.nan
        fA[5, 8, 1] = np.inf
        lat.upload(fA, walls)
        with pytest.raises(_capi.LbmError) as e:
            lat.check_finite()
        assert e.value.code == _capi.ERR_FPE and "2 non-finite" in str(e.value)


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_weak_scaling_unit_windows_match_oracle(prec):
    """8192x32768 (BASELINE configs[4], one GPU's share), library defaults, two macro-steps of k = 16."""
    from latticeboltzmann_b200 import Lattice, tb_schedule
    nx, ny, T = 8192, 32768, 32
    with Lattice(nx, ny, prec, inplace=True) as lat:
        tune = lat.get_tuning()
        assert tune["tb_depth"] == 16, tune
        k, launches = tb_schedule(nx, ny, tune["tb_depth"], tune["tb_rows"], tune["tb_cols"], prec)
        row_seams = sorted({l[1] for l in launches if l[0] == 0 and 0 < l[1] < nx})
        col_seams = sorted({l[3] for l in launches if l[0] == 0 and 0 < l[3] < ny})
        assert len(row_seams) >= 2 and len(col_seams) >= 2, "the default schedule of this lattice is 2-D"
        lat.InitializeArrays(**_scene(nx, ny))
        lat.run(T)
        for a, b, c, d, what in _windows(nx, ny, [row_seams[0], row_seams[len(row_seams) // 2], row_seams[-1]],
                                         [col_seams[0], col_seams[len(col_seams) // 2], col_seams[-1]]):
            _check_window([lat], prec, nx, ny, T, a, b, c, d, f"{prec} in place + temporal blocking, {what}")
        inplace_sum = lat.checksum()
    with Lattice(nx, ny, prec, inplace=False) as ab:            # the other engine: A/B double buffer, one pull kernel per step
        ab.InitializeArrays(**_scene(nx, ny))
        ab.run(T)
        assert ab.checksum() == inplace_sum, "in-place + temporal blocking and A/B disagree at 8192x32768"


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_slab_ring_at_full_width_windows_match_oracle(prec):
    """Four slabs of 1024 x 32768 (rows as wide as the weak-scaling unit's, library defaults for slabs: temporal blocking
    across the seams, 32-KB strips), stepped by lbm_group_run: windows on every slab seam and on the strip seams."""
    from latticeboltzmann_b200 import group_run, make_ring, tb_schedule
    nx, ny, T, P = 4096, 32768, 32, 4
    slabs = make_ring(nx, ny, P, prec, inplace=True)
    try:
        tune = slabs[0].get_tuning()
        assert tune["tb_depth"] == 16, tune
        _, launches = tb_schedule(nx // P, ny, tune["tb_depth"], tune["tb_rows"], tune["tb_cols"], prec, multi=True)
        col_seams = sorted({l[3] for l in launches if l[0] == 0 and 0 < l[3] < ny})
        assert len(col_seams) >= 2
        for s in slabs:
            s.InitializeArrays(**_scene(nx, ny))
        group_run(slabs, T)
        seams = [s.row_lo for s in slabs]
        for a, b, c, d, what in _windows(nx, ny, seams[1:], [col_seams[0], col_seams[-1]]):
            _check_window(slabs, prec, nx, ny, T, a, b, c, d, f"{prec} ring of {P} slabs, {what}")
        total = sum(s.checksum() for s in slabs) % 2 ** 64
    finally:
        for s in slabs:
            s.close()
    from latticeboltzmann_b200 import Lattice
    with Lattice(nx, ny, prec, inplace=False) as ab:
        ab.InitializeArrays(**_scene(nx, ny))
        ab.run(T)
        assert ab.checksum() == total, "the slab ring and the single-GPU A/B engine disagree"
