"""GPU (-m gpu): the boundary options the reference names and never implements -- open y ends
(LB.c:42-43, lbm_config.boundary_y = LBM_Y_OUTFLOW), "reflect" (slip) walls (LB.c:21, wall values 2 / 3) and geometry
from a file (lbm_load_walls, instead of the hard-wired barrier of LB.c:567-578).

PARITY UNPINNED: the reference has no code for these, so the CUDA path is compared (bit for bit) with the CPU oracle's
restatement of the definitions in include/lbm_b200.h, not with anything the reference computes.  With the options off
the same kernels are pinned by the golden vectors (tests/test_gpu_parity.py)."""
import numpy as np
import pytest

from tests import cases

pytestmark = pytest.mark.gpu
VW = {("dp", "avx"): 4, ("dp", "sse"): 2, ("dp", "serial"): 1, ("sp", "avx"): 8, ("sp", "sse"): 4, ("sp", "serial"): 1}


def _inputs(prec, nx, ny, slip):
    c = dict(prec=prec, isa="avx", nx=nx, ny=ny, scene="random", perturb=23)
    fA, walls = cases.make_inputs(c)
    if slip:
        walls[0, :] = 2                                   # the channel walls reflect (normal x) ...
        walls[nx - 1, :] = 2
        walls[nx // 3:nx // 3 + 5, ny // 2] = 3           # ... and so does a thin plate across the flow (normal y)
        walls[nx // 2, ny // 4:ny // 4 + 9] = 2           # a plate along the flow, and random bounce-back sites everywhere
    return fA, walls


def _same(got, want, what):
    if got.tobytes() != want.tobytes():
        bad = np.argwhere(got != want)
        raise AssertionError(f"{what}: {len(bad)} values differ, first (s,i,j) = {bad[0].tolist()}: {got[tuple(bad[0])]!r} vs {want[tuple(bad[0])]!r}")


@pytest.mark.parametrize("slip", [False, True], ids=["bounce", "slip"])
@pytest.mark.parametrize("outflow", [False, True], ids=["periodic", "outflow"])
@pytest.mark.parametrize("prec,isa,nx,ny", [("dp", "avx", 40, 264), ("sp", "avx", 40, 264), ("dp", "sse", 33, 130), ("sp", "sse", 33, 132),
                                            ("dp", "serial", 17, 45), ("sp", "serial", 17, 45), ("dp", "avx", 24, 131), ("sp", "avx", 24, 133)])
def test_boundary_options_match_oracle(prec, isa, nx, ny, outflow, slip):
    from latticeboltzmann_b200 import Lattice, LbmError, group_run, make_ring
    from oracle.harness import Oracle
    if not outflow and not slip:
        pytest.skip("the reference's own configuration: tests/test_gpu_parity.py")
    vw, T = VW[(prec, isa)], 24
    fA, walls = _inputs(prec, nx, ny, slip)
    o = Oracle(prec, nx, ny, vw, outflow=outflow)
    o.set_state(fA, walls)
    o.run(T)
    want_a, want_b = o.f_view("A").copy(), o.f_view("B").copy()
    if outflow:                                           # really different from the periodic lattice
        per = Oracle(prec, nx, ny, vw)
        per.set_state(fA, walls)
        per.run(T)
        assert per.f_view("A").tobytes() != want_a.tobytes()
    # A/B engine: every vector width the build offers, single slab and a ring of three
    vecs = [v for v in (1, 2, 4) if v <= max(vw, 1) and v * (8 if prec == "dp" else 4) <= 32]
    for v in vecs:
        with Lattice(nx, ny, prec, vecwidth=vw, outflow=outflow) as lat:
            lat.set_tuning(vec=v)
            lat.upload(fA, walls)
            lat.run(T)
            _same(lat.download("A"), want_a, f"A/B fA vec {v}")
            _same(lat.download("B"), want_b, f"A/B fB vec {v}")
            assert np.array_equal(lat.download_walls(), walls)
    slabs = make_ring(nx, ny, 3, prec, vecwidth=vw, outflow=outflow)
    try:
        for s in slabs:
            s.upload(fA, walls)
        group_run(slabs, T)
        got = np.zeros_like(want_a)
        for s in slabs:
            s.download("A", out=got)
        _same(got, want_a, "A/B ring of 3")
    finally:
        for s in slabs:
            s.close()
    # in place: plain, temporally blocked (bands and 2-D tiles), ring -- open ends need the vector kernels on every column
    if outflow and (vw == 1 or ny % vw):
        with pytest.raises(LbmError):
            Lattice(nx, ny, prec, vecwidth=vw, outflow=True, inplace=True)
        return
    for tune in (dict(), dict(tb_depth=4, tb_rows=5), dict(tb_depth=6, tb_rows=7, tb_cols=64, pdl=1), dict(vec=1, tb_depth=2, hints=1),
                 dict(vec=min(2, max(vw, 1)), tb_depth=4, tb_rows=3, tb_cols=16),
                 # concurrent tile chains (lanes on their own streams): the open ends and the slip walls across lane boundaries
                 dict(tb_depth=4, tb_rows=3, tb_chains=2), dict(tb_depth=2, tb_rows=2, tb_cols=32, tb_chains=3, pdl=1)):
        with Lattice(nx, ny, prec, vecwidth=vw, outflow=outflow, inplace=True) as lat:
            lat.set_tuning(**tune)
            lat.upload(fA, walls)
            lat.run(4)
            lat.run(T - 4)
            _same(lat.download("A"), want_a, f"in-place fA with {tune}")
    for tune in (dict(), dict(tb_depth=4, tb_rows=3, tb_cols=32)):
        slabs = make_ring(nx, ny, 3, prec, vecwidth=vw, outflow=outflow, inplace=True)
        try:
            for s in slabs:
                s.set_tuning(**tune)
                s.upload(fA, walls)
            group_run(slabs, T)
            got = np.zeros_like(want_a)
            for s in slabs:
                s.download("A", out=got)
            _same(got, want_a, f"in-place ring of 3 with {tune}")
        finally:
            for s in slabs:
                s.close()


def test_scene_slip_channel_and_physics():
    """lbm_scene.channel_walls = 2 (reflect walls) on the device == the oracle; a slip channel runs faster than a no-slip one."""
    from latticeboltzmann_b200 import Lattice, geometry
    from oracle.harness import Oracle
    nx, ny, T = 32, 256, 400
    res = {}
    for kind in (1, 2):
        with Lattice(nx, ny, "dp", inplace=True) as lat:
            lat.InitializeArrays(channel_walls=kind, bar_i0=0, bar_i1=0)
            walls = lat.download_walls()
            assert set(np.unique(walls[[0, nx - 1]])) == {kind} and not walls[1:nx - 1].any()
            lat.run(T)
            got = lat.download("A")
            rho, ux, uy = lat.macroscopic()
        o = Oracle("dp", nx, ny)
        o.set_state(geometry.equilibrium_rest(nx, ny, "dp"), walls)
        o.run(T)
        assert got.tobytes() == o.f_view("A").tobytes()
        res[kind] = float(uy[1:nx - 1].mean())
        if kind == 2:                                     # plug flow: no shear at a reflecting wall
            assert abs(uy[1, 5:].mean() - uy[nx // 2, 5:].mean()) < 0.05 * abs(uy[nx // 2, 5:].mean())
    assert res[2] > 1.2 * res[1] > 0


def test_load_walls_from_file(tmp_path):
    from latticeboltzmann_b200 import Lattice, LbmError, make_ring
    from oracle.harness import Oracle
    nx, ny = 19, 72
    fA, walls = _inputs("dp", nx, ny, slip=True)
    dense = tmp_path / "dense.txt"
    dense.write_text("# geometry: 0 fluid, 1 bounce-back, 2 / 3 reflect\n" + "\n".join("".join(str(v) for v in row) for row in walls) + "\n")
    csv = tmp_path / "csv.txt"
    csv.write_text("\n".join(", ".join(str(v) for v in row) for row in walls))          # no trailing newline
    o = Oracle("dp", nx, ny)
    o.set_state(fA, walls)
    o.run(10)
    for path in (dense, csv):
        with Lattice(nx, ny, "dp") as lat:
            lat.upload(fA, np.zeros_like(walls))
            lat.load_walls(path)
            assert np.array_equal(lat.download_walls(), walls)
            lat.run(10)
            assert lat.download("A").tobytes() == o.f_view("A").tobytes()
    slabs = make_ring(nx, ny, 3, "dp")                     # every slab reads the whole file and keeps its rows
    try:
        got = np.zeros_like(walls)
        for s in slabs:
            s.load_walls(dense)
            part = np.zeros_like(walls)
            s._call("lbm_download", None, None, part.ctypes.data_as(__import__("ctypes").c_void_p), ny)
            got[s.row_lo:s.row_hi] = part[s.row_lo:s.row_hi]
        assert np.array_equal(got, walls)
    finally:
        for s in slabs:
            s.close()
    with Lattice(nx, ny, "dp") as lat:
        bad = tmp_path / "bad.txt"
        for text in ("\n".join("".join(str(v) for v in row) for row in walls[:-1]),                       # a row short
                     "\n".join("".join(str(v) for v in row) for row in walls).replace("2", "7", 1),       # not a mask value
                     "\n".join("".join(str(v) for v in row)[:-1] for row in walls)):                      # a column short
            bad.write_text(text)
            with pytest.raises(LbmError):
                lat.load_walls(bad)
        with pytest.raises(LbmError):
            lat.load_walls(tmp_path / "missing.txt")
