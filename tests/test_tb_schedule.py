"""CPU: the temporal-blocking (wavefront) schedule of the in-place engine is a valid order.

The in-place kernels have read-after-write dependencies only (DESIGN.md "Temporal blocking"): row i may
take level s once rows i-1, i, i+1 (periodic) have taken level s-1.  lbm_tb_schedule returns the launch
list the library really enqueues (host logic, no GPU); this test replays it on a row-level model and
checks every dependency, full coverage and the band sizes."""
import numpy as np
import pytest

from latticeboltzmann_b200.solver import tb_schedule

CASES = [(400, 4000, 16, 0), (400, 4000, 8, 148), (800, 4000, 12, 196), (96, 520, 46, 3), (96, 520, 4, 1), (11, 64, 4, 2),
         (10, 64, 4, 0), (7, 40, 2, 0), (6, 40, 2, 5), (5, 40, 2, 0), (3, 40, 2, 0), (8192, 32768, 4, 15), (401, 1000, 32, 37),
         (64, 64, 30, 2), (66, 64, 32, 2), (1000, 256, 1024, 0)]


@pytest.mark.parametrize("prec", ["dp", "sp"])
@pytest.mark.parametrize("nx,ny,depth,rows", CASES)
def test_wavefront_schedule_respects_dependencies(nx, ny, depth, rows, prec):
    k, launches = tb_schedule(nx, ny, depth, rows, prec)
    if 2 * (depth & ~1) + 2 > nx:                       # ring too small for that depth: clamped, possibly to none
        assert k < depth and (k == 0 or 2 * k + 2 <= nx)
    else:
        assert k == depth & ~1
    if k == 0:
        assert launches == []
        return
    level = np.zeros(nx, dtype=np.int64)                # levels completed per row
    for s, r0, r1 in launches:
        assert 0 <= s < k and 0 <= r0 < r1 <= r0 + nx and r0 < nx
        rows_ = np.arange(r0, r1) % nx
        assert len(set(rows_.tolist())) == len(rows_)
        for d in (-1, 0, 1):                            # RAW: the neighbours' previous level is done (and nothing is skipped)
            assert (level[(rows_ + d) % nx] >= s).all(), f"level {s} rows [{r0},{r1}): neighbour {d:+d} not at level {s}"
        assert (level[rows_] == s).all(), f"level {s} rows [{r0},{r1}) taken twice or out of order"
        level[rows_] += 1
    assert (level == k).all()                           # every row took every level exactly once
    # working set of a band: (H + k) rows stay within ~70 MB unless the caller forced the band height
    if rows == 0:
        row_bytes = 9 * ny * (8 if prec == "dp" else 4)
        tall = max(r1 - r0 for _, r0, r1 in launches)
        assert (tall + k) * row_bytes <= 70e6 or tall <= 2 * k + 2


def test_bad_arguments():
    from latticeboltzmann_b200 import LbmError
    with pytest.raises(LbmError):
        tb_schedule(0, 64, 4)
