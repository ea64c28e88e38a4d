"""CPU: the data movement of the in-place (AA) kernels and of the fused pair kernel, on a symbolic lattice.

Every population is a label (site, speed, step) = "post-collide f_speed of site after step"; a phase reads the
slots the CUDA kernels read (lbm_kernels.cuh: aa_load / aa_store, stream_collide_aa_pair_kernel,
stream_collide_aa_edges_kernel), checks that each holds exactly the label the reference's pull (LB.c:235-243)
would fetch, and writes the labels of its results where the kernels store them.  Arithmetic is not modelled
(relax_sites is shared by all kernels and checked bit for bit on the GPU); what is checked here is that no
slot is read before it is written, overwritten while still needed, or written by two producers -- for any
processing order of sites, tiles and CTAs."""
import numpy as np
import pytest

# pull offsets (LB.c:235-243): speed -> (row shift, column shift) of the site the population is pulled from
PULL = [(0, 0), (0, -1), (-1, 0), (0, 1), (1, 0), (-1, -1), (-1, 1), (1, 1), (1, -1)]
OPP = [0, 3, 4, 1, 2, 7, 8, 5, 6]


class Model:
    def __init__(self, nx, ny):
        self.nx, self.ny = nx, ny
        r, c = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
        self.site = r * ny + c
        self.mem = np.stack([self.label(self.site, s, 0) for s in range(9)])    # reference layout after step 0

    def label(self, site, s, t):
        return (t * 9 + s) * (self.nx * self.ny) + site

    def at(self, r, c):
        return (r % self.nx) * self.ny + (c % self.ny)

    def phase1(self, r, c, t):
        """site (r, c) takes step t from the reference layout (aa_load<1> / aa_store<1>)"""
        for s in range(9):
            rr, cc = (r + PULL[s][0]) % self.nx, (c + PULL[s][1]) % self.ny
            assert self.mem[s, rr, cc] == self.label(self.at(rr, cc), s, t - 1), (r, c, s)
        for k in range(9):
            rr, cc = (r + PULL[k][0]) % self.nx, (c + PULL[k][1]) % self.ny
            self.mem[k, rr, cc] = self.label(self.at(r, c), OPP[k], t)

    def phase2_read(self, r, c, t):
        for s in range(9):
            src = self.at(r + PULL[s][0], c + PULL[s][1])
            assert self.mem[OPP[s], r, c] == self.label(src, s, t - 1), (r, c, s)

    def phase2(self, r, c, t):
        """site (r, c) takes step t back to the reference layout (aa_load<2> / aa_store<2>)"""
        self.phase2_read(r, c, t)
        for s in range(9):
            self.mem[s, r, c] = self.label(self.at(r, c), s, t)

    def is_reference_layout(self, t):
        return all((self.mem[s] == self.label(self.site, s, t)).all() for s in range(9))

    def is_after_phase1(self, t):
        """slot k of site z holds o_opp(k) of step t of the site z - pull(k) that pulled through it"""
        for k in range(9):
            r = (np.arange(self.nx)[:, None] - PULL[k][0]) % self.nx
            c = (np.arange(self.ny)[None, :] - PULL[k][1]) % self.ny
            if not (self.mem[k] == self.label(r * self.ny + c, OPP[k], t)).all():
                return False
        return True


@pytest.mark.parametrize("nx,ny", [(1, 2), (2, 3), (3, 8), (7, 5), (12, 16)])
def test_in_place_phases_any_site_order(nx, ny):
    m = Model(nx, ny)
    rng = np.random.default_rng(nx * 100 + ny)
    sites = [(r, c) for r in range(nx) for c in range(ny)]
    for t in range(1, 7):
        for i in rng.permutation(len(sites)):
            (m.phase1 if t & 1 else m.phase2)(*sites[i], t)
        assert m.is_after_phase1(t) if t & 1 else m.is_reference_layout(t)


def _pair_tile(m, t, tr0, tr1, c0, c1):
    """one CTA of stream_collide_aa_pair_kernel: phase 2 (step t) of its tile row by row, results in a ring,
    phase 1 (step t + 1) of row r-1 once row r is in the ring; stores as the kernel's"""
    ring = {}
    for r in range(tr0, tr1):
        for c in range(c0, c1):
            m.phase2_read(r, c, t)
            ring[(r, c)] = [m.label(m.at(r, c), s, t) for s in range(9)]
            edge = c in (c0, c1 - 1) or r in (tr0, tr1 - 1)
            for k in range(9):
                if edge:
                    m.mem[k, r, c] = ring[(r, c)][k]
                    continue
                yr, yc = r - PULL[k][0], c - PULL[k][1]           # the slot's consumer in phase 1
                if k and (yr in (tr0, tr1 - 1) or yc in (c0, c1 - 1)):
                    m.mem[k, r, c] = ring[(r, c)][k]
        y = r - 1
        if y > tr0:
            for c in range(c0 + 1, c1 - 1):
                for s in range(9):
                    src = (y + PULL[s][0], c + PULL[s][1])
                    assert tr0 <= src[0] <= r and c0 <= src[1] < c1
                    assert ring[src][s] == m.label(m.at(*src), s, t)
                for k in range(9):
                    m.mem[k, y + PULL[k][0], c + PULL[k][1]] = m.label(m.at(y, c), OPP[k], t + 1)
        for key in [key for key in ring if key[0] < r - 2]:       # the kernel's ring holds three rows back at most
            del ring[key]


@pytest.mark.parametrize("nx,ny,tile_rows,tile_cols", [(12, 16, 5, 8), (9, 24, 3, 8), (16, 16, 16, 16), (7, 12, 4, 4),
                                                       (10, 8, 200, 128), (5, 8, 3, 2), (1, 4, 3, 4), (2, 4, 5, 4)])
def test_fused_pairs_leave_what_two_phases_leave(nx, ny, tile_rows, tile_cols):
    m = Model(nx, ny)
    rng = np.random.default_rng(nx * 1000 + ny * 10 + tile_rows)
    for r in range(nx):
        for c in range(ny):
            m.phase1(r, c, 1)
    tiles = [(a, min(a + tile_rows, nx), b, min(b + tile_cols, ny)) for a in range(0, nx, tile_rows) for b in range(0, ny, tile_cols)]
    for t in (2, 4, 6):
        for i in rng.permutation(len(tiles)):                     # CTAs in any order: tiles touch their own sites only
            _pair_tile(m, t, *tiles[i])
        rims = [(r, c) for a, b, c0, c1 in tiles for r in range(a, b) for c in range(c0, c1) if r in (a, b - 1) or c in (c0, c1 - 1)]
        assert len(set(rims)) == len(rims)
        for i in rng.permutation(len(rims)):                      # stream_collide_aa_edges_kernel
            m.phase1(*rims[i], t + 1)
        assert m.is_after_phase1(t + 1)
    for r in range(nx):
        for c in range(ny):
            m.phase2(r, c, 8)
    assert m.is_reference_layout(8)


# the wavefront (temporal blocking) schedule with the in-place phases: the library's own launch list, sites of a
# launch in any order -- every read must find the label of the step before, although neighbouring tiles are
# several steps apart (the claim behind DESIGN.md "temporal blocking": read-after-write dependencies only)
@pytest.mark.parametrize("nx,ny,depth,rows,cols", [(24, 144, 8, 5, 16), (33, 136, 4, 9, 64), (11, 64, 4, 2, 0),
                                                   (40, 64, 6, 7, 0), (40, 64, 6, 7, 1 << 20), (14, 40, 6, 2, 1 << 20)])
def test_wavefront_schedule_with_in_place_phases(nx, ny, depth, rows, cols):
    from latticeboltzmann_b200.solver import tb_schedule
    k, launches = tb_schedule(nx, ny, depth, rows, cols, "dp")
    assert k >= 2
    m = Model(nx, ny)
    rng = np.random.default_rng(nx + ny + depth)
    for macro in range(2):
        for s, r0, r1, j0, j1 in launches:
            if j0 == j1 == 0:
                j1 = ny
            sites = [(r % nx, c % ny) for r in range(r0, r1) for c in range(j0, j1)]
            t = macro * k + s + 1
            for i in rng.permutation(len(sites)):
                (m.phase1 if t & 1 else m.phase2)(*sites[i], t)
        assert m.is_reference_layout((macro + 1) * k)
