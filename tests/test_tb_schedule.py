"""CPU: the temporal-blocking (wavefront) schedule of the in-place engine is a valid order.

The in-place kernels have read-after-write dependencies only (DESIGN.md "Temporal blocking"): site (i, j)
may take level s once its 3x3 neighbourhood (periodic in both axes) has taken level s-1.  lbm_tb_schedule
returns the launch list the library really enqueues (host logic, no GPU); this test replays it on a
site-level model and checks every dependency, full coverage and the tile sizes."""
import numpy as np
import pytest

from latticeboltzmann_b200.solver import tb_schedule

#        nx    ny   depth rows cols
CASES = [(400, 4000, 16, 0, 0), (400, 4000, 8, 148, 0), (800, 4000, 12, 196, 0), (96, 520, 46, 3, 0), (96, 520, 4, 1, 0),
         (11, 64, 4, 2, 0), (10, 64, 4, 0, 0), (7, 40, 2, 0, 0), (6, 40, 2, 5, 0), (5, 40, 2, 0, 0), (3, 40, 2, 0, 0),
         (401, 1000, 32, 37, 0), (64, 64, 30, 2, 0), (66, 64, 32, 2, 0), (1000, 256, 1024, 0, 0),
         # column strips as well (2-D tiles)
         (96, 520, 4, 20, 128), (96, 520, 6, 0, 104), (40, 1024, 8, 12, 256), (200, 2048, 16, 50, 512), (33, 136, 4, 9, 64),
         (24, 144, 8, 5, 16), (64, 4096, 4, 16, 0), (50, 264, 16, 10, 8)]


def _check(nx, ny, k, launches):
    level = np.zeros((nx, ny), dtype=np.int16)          # levels completed per site
    for s, r0, r1, j0, j1 in launches:
        if j0 == j1 == 0:
            j1 = ny
        assert 0 <= s < k and 0 <= r0 < r1 <= r0 + nx and r0 < nx and 0 <= j0 < j1 <= j0 + ny and j0 < ny
        assert j0 % 8 == 0 and (j1 % 8 == 0 or j1 == ny)
        rows, cols = np.arange(r0, r1) % nx, np.arange(j0, j1) % ny
        mine = level[np.ix_(rows, cols)]
        assert (mine == s).all(), f"level {s} rows [{r0},{r1}) cols [{j0},{j1}) taken twice or out of order"
        for di in (-1, 0, 1):                           # RAW: the 3x3 neighbourhood's previous level is done
            for dj in (-1, 0, 1):
                nb = level[np.ix_((rows + di) % nx, (cols + dj) % ny)]
                assert (nb >= s).all(), f"level {s} rows [{r0},{r1}) cols [{j0},{j1}): neighbour ({di:+d},{dj:+d}) behind"
        level[np.ix_(rows, cols)] = s + 1
    assert (level == k).all()                           # every site took every level exactly once


@pytest.mark.parametrize("prec", ["dp", "sp"])
@pytest.mark.parametrize("nx,ny,depth,rows,cols", CASES)
def test_wavefront_schedule_respects_dependencies(nx, ny, depth, rows, cols, prec):
    k, launches = tb_schedule(nx, ny, depth, rows, cols, prec)
    if 2 * (depth & ~1) + 2 > nx:                       # ring too small for that depth: clamped, possibly to none
        assert k < depth and (k == 0 or 2 * k + 2 <= nx)
    else:
        assert k == depth & ~1
    if k == 0:
        assert launches == []
        return
    _check(nx, ny, k, launches)
    if 0 < cols < ny and ny % 8 == 0 and 16 * k + 8 <= ny:
        assert any(j1 > j0 for _, _, _, j0, j1 in launches), "column strips were asked for and possible"
    # working set of a tile: (H + k) x (W + 8k) sites stay within ~70 MB unless the caller forced the sizes
    if rows == 0 and cols == 0:
        site = 9 * (8 if prec == "dp" else 4)
        worst = max((r1 - r0 + k) * ((j1 - j0 + 8 * k) if j1 > j0 else ny) for _, r0, r1, j0, j1 in launches)
        assert worst * site <= 70e6 or max(r1 - r0 for _, r0, r1, _, _ in launches) <= 2 * k + 2


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_wide_lattices_get_column_strips_by_default(prec):
    """8192x32768 (the weak-scaling unit): full-width bands would be ~19 rows, so the automatic plan cuts strips;
    the area of all launches is k * nx * ny and no tile (with its skew margins) exceeds the L2 budget."""
    nx, ny, k = 8192, 32768, 8
    used, launches = tb_schedule(nx, ny, k, 0, 0, prec)
    assert used == k and all(j1 > j0 for _, _, _, j0, j1 in launches)
    area = sum((r1 - r0) * (j1 - j0) for _, r0, r1, j0, j1 in launches)
    assert area == k * nx * ny
    site = 9 * (8 if prec == "dp" else 4)
    assert max((r1 - r0 + k) * (j1 - j0 + 8 * k) for _, r0, r1, j0, j1 in launches) * site <= 70e6


def test_bad_arguments():
    from latticeboltzmann_b200 import LbmError
    with pytest.raises(LbmError):
        tb_schedule(0, 64, 4)


# ---------------------------------------------------------------------------------------------------
# several slabs: every slab walks its own launch list; the only coupling is the event protocol of
# lbm_tb_schedule's `seam` bits (wait: both ring neighbours have announced as many events as I have;
# announce: after the launch).  Any interleaving the protocol allows must respect the global dependencies.
# ---------------------------------------------------------------------------------------------------
RINGS = [  # (slab heights, ny, depth, rows, cols)
    ([40, 40], 264, 4, 9, 64), ([41, 40, 40], 136, 6, 0, 0), ([30, 5, 30], 264, 8, 7, 64),      # middle slab too low: whole-slab launches
    ([3, 3, 3, 3], 64, 2, 0, 0), ([1, 1], 64, 2, 0, 0), ([50, 49], 520, 16, 12, 104), ([20, 21, 20, 21, 20], 40, 4, 5, 0),
]


@pytest.mark.parametrize("heights,ny,depth,rows,cols", RINGS)
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_slab_ring_event_protocol(heights, ny, depth, rows, cols, seed):
    """Two in-order queues per slab, as enqueue_macro builds them: stream 0 takes the launches without seam bits,
    stream 1 the seam chain; a chain launch starts only after every stream-0 launch listed before it (the event
    recorded when the chain of a column piece opens) and after both neighbours have announced as many events as
    this slab has.  Any interleaving allowed by that must respect the global 3x3 dependencies."""
    P, nx = len(heights), sum(heights)
    lo = np.concatenate([[0], np.cumsum(heights)[:-1]])
    main, chain = [], []
    for h in heights:
        k, launches = tb_schedule(h, ny, depth, rows, cols, "dp", multi=True)
        assert k == depth                               # never clamped per slab: all slabs walk the same events
        assert all(l[5] in (0, 3) for l in launches)
        main.append([(i, l) for i, l in enumerate(launches) if l[5] == 0])
        chain.append([(i, l) for i, l in enumerate(launches) if l[5] == 3])
    assert len({len(c) for c in chain}) == 1, "slabs disagree on the number of events"
    level = np.zeros((nx, ny), dtype=np.int16)
    mpos, cpos = [0] * P, [0] * P
    rng = np.random.default_rng(seed)

    def run(p, launch):
        s, r0, r1, j0, j1, seam = launch
        if j0 == j1 == 0:
            j1 = ny
        h = heights[p]
        rws = lo[p] + (np.arange(r0, r1) % h)           # rows past the slab wrap INSIDE the slab (its first rows)
        cls = np.arange(j0, j1) % ny
        assert len(set(rws.tolist())) == len(rws)
        assert (level[np.ix_(rws, cls)] == s).all(), f"slab {p} level {s} rows [{r0},{r1}) cols [{j0},{j1}) out of order"
        for di in (-1, 0, 1):
            for dj in (-1, 0, 1):
                assert (level[np.ix_((rws + di) % nx, (cls + dj) % ny)] >= s).all(), \
                    f"slab {p} level {s} rows [{r0},{r1}) cols [{j0},{j1}): neighbour ({di:+d},{dj:+d}) behind (seam {seam})"
        if not seam and P > 1:                          # stream-0 launches must not touch another slab's rows at all
            assert ((rws - 1 >= lo[p]) & (rws + 1 < lo[p] + h)).all()
        level[np.ix_(rws, cls)] = s + 1

    while any(mpos[p] < len(main[p]) or cpos[p] < len(chain[p]) for p in range(P)):
        runnable = []
        for p in range(P):
            if mpos[p] < len(main[p]):
                runnable.append((p, 0))
            if cpos[p] < len(chain[p]):
                idx = chain[p][cpos[p]][0]
                main_done = mpos[p] == len(main[p]) or main[p][mpos[p]][0] > idx
                if main_done and cpos[(p + 1) % P] >= cpos[p] and cpos[(p - 1) % P] >= cpos[p]:
                    runnable.append((p, 1))
        assert runnable, f"deadlock: main {mpos}, chain {cpos}"
        p, q = runnable[int(rng.integers(len(runnable)))]
        if q == 0:
            run(p, main[p][mpos[p]][1])
            mpos[p] += 1
        else:
            run(p, chain[p][cpos[p]][1])
            cpos[p] += 1
    assert (level == depth).all()


# ---------------------------------------------------------------------------------------------------
# BASELINE.json's full sizes: the 2-D site model above would need gigabytes, but the schedule is a product of two
# 1-D schedules (column pieces outermost, row pieces next, levels innermost), so it is enough that (i) the launch
# list really is that product and (ii) each axis' own piece order respects the 1-D dependencies (reach 1).
# ---------------------------------------------------------------------------------------------------
def _axis_check(n, k, pieces):
    """pieces: list over axis pieces (in schedule order) of {level: (lo, hi)}; 1-D level model with reach 1"""
    level = np.zeros(n, dtype=np.int16)
    for ranges in pieces:
        for s in range(k):
            if s not in ranges:
                continue
            lo, hi = ranges[s]
            idx = np.arange(lo, hi) % n
            assert (level[idx] == s).all()
            assert (level[(idx - 1) % n] >= s).all() and (level[(idx + 1) % n] >= s).all()
            level[idx] = s + 1
    assert (level == k).all()


@pytest.mark.parametrize("nx,ny,depth,prec", [(8192, 32768, 16, "dp"), (8192, 32768, 16, "sp"), (16384, 65536, 16, "dp"),
                                              (16384, 65536, 24, "dp"), (400, 4000, 16, "dp"), (800, 4000, 16, "sp")])
def test_full_size_schedules_are_valid_products(nx, ny, depth, prec):
    k, launches = tb_schedule(nx, ny, depth, 0, 0, prec)
    assert k == depth
    whole_rows = all(j0 == j1 == 0 for _, _, _, j0, j1 in launches)
    # (column piece, row piece) groups: inside one the level rises strictly; a new column piece starts where a group's
    # column range disagrees with the current piece's at a level both have (inverted pyramids are empty at level 0)
    groups = []
    for s, r0, r1, j0, j1 in launches:
        if not groups or s <= groups[-1]["last"]:
            groups.append(dict(cols={}, rows={}, last=-1))
        g = groups[-1]
        g["cols"][s] = (0, ny) if whole_rows else (j0, j1)
        g["rows"][s] = (r0, r1)
        g["last"] = s
    col_pieces, row_pieces_of = [], []
    for g in groups:
        if not col_pieces or any(col_pieces[-1].get(s, cr) != cr for s, cr in g["cols"].items()):
            col_pieces.append({})
            row_pieces_of.append([])
        col_pieces[-1].update(g["cols"])
        row_pieces_of[-1].append(g["rows"])
    _axis_check(ny, k, col_pieces)
    assert set(col_pieces[0]) == set(range(k))                     # the column pyramid (or the whole-row piece) exists at every level
    for cp, rp in zip(col_pieces, row_pieces_of):                  # every column piece sweeps the same row pieces, at the levels it has
        want = [{s: rng for s, rng in ref.items() if s in cp} for ref in row_pieces_of[0]]
        assert rp == [w for w in want if w]
    _axis_check(nx, k, row_pieces_of[0])
    area = sum((r1 - r0) * ((j1 - j0) if j1 > j0 else ny) for _, r0, r1, j0, j1 in launches)
    assert area == k * nx * ny


# ---------------------------------------------------------------------------------------------------
# concurrent tile chains (lbm_tuning.tb_chains): lanes of one slab on their own streams, seam chains between them.
# What orders them (enqueue_macro_lanes): each stream is in order; the chain of boundary b takes a column piece only after
# lanes b and b+1 have finished their tiles of that piece; a lane starts the next macro-step only after its two chains.
# ---------------------------------------------------------------------------------------------------
CHAIN_CASES = [  # nx, ny, chains, depth, rows, cols
    (400, 4000, 2, 16, 0, 0), (96, 264, 2, 4, 9, 64), (120, 520, 3, 6, 0, 104), (200, 2048, 4, 8, 10, 512), (80, 136, 2, 16, 0, 0),
    (64, 64, 2, 4, 0, 0), (401, 1000, 3, 32, 20, 0),
]


def _pieces(launches, C):
    """split the enqueue-ordered list into (column piece, stream) segments: a new column piece begins whenever lane 0 comes
    back after a chain launch"""
    piece, out, seen_chain = 0, [], False
    for l in launches:
        st = l[5]
        if st >= C:
            seen_chain = True
        elif st == 0 and seen_chain:
            piece, seen_chain = piece + 1, False
        out.append((piece, l))
    return out


@pytest.mark.parametrize("nx,ny,chains,depth,rows,cols", CHAIN_CASES)
@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_chain_schedule_is_a_valid_sequential_order(nx, ny, chains, depth, rows, cols, prec):
    from latticeboltzmann_b200.solver import tb_schedule_chains
    k, C, launches = tb_schedule_chains(nx, ny, chains, depth, rows, cols, prec)
    assert k == depth and 2 <= C <= chains
    _check(nx, ny, k, [l[:5] for l in launches])
    assert {l[5] for l in launches} == set(range(2 * C))
    assert all((l[6] == 1) == (l[5] == 2 * C - 1) for l in launches)          # only the last boundary is the outer one
    lanes_rows = [sorted({r for l in launches if l[5] == s for r in range(l[1], l[2])}) for s in range(C)]
    for a in range(C):                                  # lanes own disjoint row ranges
        for b in range(a + 1, C):
            assert not set(lanes_rows[a]) & set(lanes_rows[b])


def test_too_low_for_lanes_falls_back():
    from latticeboltzmann_b200.solver import tb_schedule_chains
    k, C, launches = tb_schedule_chains(60, 264, 2, 16)          # two lanes of 30 rows cannot hold a 16-level pyramid
    assert C == 1 and launches == []
    k, C, launches = tb_schedule_chains(200, 264, 4, 16)         # 2k + 4 = 36 rows per lane: five would fit, four were asked for
    assert C == 4


@pytest.mark.parametrize("nx,ny,chains,depth,rows,cols", CHAIN_CASES)
@pytest.mark.parametrize("seed", [0, 1, 2])
def test_chain_schedule_under_concurrency(nx, ny, chains, depth, rows, cols, seed):
    """Any interleaving of the 2C in-order streams that respects the two event rules must respect the 3x3 dependencies."""
    from latticeboltzmann_b200.solver import tb_schedule_chains
    k, C, launches = tb_schedule_chains(nx, ny, chains, depth, rows, cols, "dp")
    tagged = _pieces(launches, C)
    queues = [[(pc, l) for pc, l in tagged if l[5] == s] for s in range(2 * C)]
    npieces = tagged[-1][0] + 1
    # lane_done[l][pc]: number of launches of lane l belonging to pieces <= pc
    lane_upto = [[sum(1 for p, _ in queues[l] if p <= pc) for pc in range(npieces)] for l in range(C)]
    pos = [0] * (2 * C)
    level = np.zeros((nx, ny), dtype=np.int16)
    rng = np.random.default_rng(seed)
    total = len(launches)
    done = 0
    while done < total:
        ready = []
        for s in range(2 * C):
            if pos[s] >= len(queues[s]):
                continue
            pc, l = queues[s][pos[s]]
            if s >= C:                                  # chain of boundary b: lanes b and b+1 have finished piece pc
                b = s - C
                if pos[b] < lane_upto[b][pc] or pos[(b + 1) % C] < lane_upto[(b + 1) % C][pc]:
                    continue
            ready.append(s)
        assert ready, "deadlock"
        s = int(rng.choice(ready))
        pc, (lv, r0, r1, j0, j1, _, _) = queues[s][pos[s]]
        if j0 == j1 == 0:
            j1 = ny
        rws, cls = np.arange(r0, r1) % nx, np.arange(j0, j1) % ny
        assert (level[np.ix_(rws, cls)] == lv).all(), f"stream {s} level {lv} rows [{r0},{r1}) out of order"
        for di in (-1, 0, 1):
            for dj in (-1, 0, 1):
                assert (level[np.ix_((rws + di) % nx, (cls + dj) % ny)] >= lv).all(), \
                    f"stream {s} level {lv} rows [{r0},{r1}) cols [{j0},{j1}): neighbour ({di:+d},{dj:+d}) behind"
        level[np.ix_(rws, cls)] = lv + 1
        pos[s] += 1
        done += 1
    assert (level == k).all()
