"""GPU (-m gpu): the CUDA path, called through the C ABI, against the CPU oracle on the same
inputs, against the golden vectors of the compiled reference, and through size-independent
properties at BASELINE.json's full sizes.

Tolerances: NONE.  f_i, the Reynolds number, rho/u and the wall mask are compared bit for bit
(stricter than north_star's 1e-12 DP / 1e-5 SP), because the kernels evaluate every operation
in the reference's order with round-to-nearest intrinsics."""
import json
import os
import subprocess

import numpy as np
import pytest

from tests import cases

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
VW = {("dp", "avx"): 4, ("dp", "sse"): 2, ("dp", "serial"): 1, ("sp", "avx"): 8, ("sp", "sse"): 4, ("sp", "serial"): 1}


def _lattice(c, **kw):
    from latticeboltzmann_b200 import Lattice
    return Lattice(c["nx"], c["ny"], c["prec"], vecwidth=VW[(c["prec"], c["isa"])], **kw)


def _oracle(c, fA, walls, steps):
    from oracle.harness import Oracle
    o = Oracle(c["prec"], c["nx"], c["ny"], VW[(c["prec"], c["isa"])])
    o.set_state(fA, walls)
    o.run(steps)
    return o


def _assert_same(got, want, what):
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape and got.dtype == want.dtype, f"{what}: {got.shape}/{got.dtype} vs {want.shape}/{want.dtype}"
    if got.tobytes() != want.tobytes():
        bad = np.argwhere(~((got == want) & (np.signbit(got) == np.signbit(want))))
        at = tuple(bad[0]) if len(bad) else None
        raise AssertionError(f"{what}: {len(bad)} elements differ, first at {at}: "
                             f"{got[at]!r} vs {want[at]!r}" if at else f"{what}: bytes differ")


# ---------------------------------------------------------------------------------------------
# against the golden vectors of the compiled reference (every case, incl. 10,000 steps)
# ---------------------------------------------------------------------------------------------
ENGINES = {"default": {}, "pdl": dict(pdl=1), "graph": dict(graph=3), "graph+pdl": dict(graph=4, pdl=1), "hints": dict(hints=1, vec=2),
           # one buffer, in-place (AA) kernels (lbm_config.inplace): only buffer A exists
           "inplace": dict(inplace=1), "inplace+graph+pdl": dict(inplace=1, graph=4, pdl=1),
           # + temporal blocking in L2: bands of tb_rows rows advanced tb_depth steps each (wavefront schedule)
           "inplace+tb": dict(inplace=1, tb_depth=4, tb_rows=3),
           "inplace+tb+graph+pdl": dict(inplace=1, tb_depth=6, tb_rows=5, graph=1, pdl=1),
           # 2-D tiles: column strips as well as row bands (falls back to bands where ny is not a multiple of 8);
           # hints: the last level of every macro-step stores with the evict-first policy
           "inplace+tb2d": dict(inplace=1, tb_depth=4, tb_rows=16, tb_cols=256, graph=1, pdl=1),
           "inplace+tb2d+hints": dict(inplace=1, tb_depth=6, tb_rows=9, tb_cols=64, pdl=1, hints=3, vec=2),
           # concurrent tile chains: the rows cut into lanes on their own streams, seam chains between them (graph: a multi-stream capture)
           "inplace+tb+chains": dict(inplace=1, tb_depth=4, tb_rows=3, tb_chains=2, pdl=1),
           "inplace+tb2d+chains+graph": dict(inplace=1, tb_depth=4, tb_rows=9, tb_cols=64, tb_chains=3, graph=1, pdl=1)}


def _split(engine_kw):
    """(constructor kwargs, tuning kwargs) of an ENGINES entry"""
    kw = dict(engine_kw)
    ctor = {"inplace": True} if kw.pop("inplace", 0) else {}
    return ctor, kw


@pytest.mark.parametrize("engine", list(ENGINES))
@pytest.mark.parametrize("name", list(cases.CASES))
def test_cuda_matches_reference_golden(name, engine):
    c, g = cases.CASES[name], GOLDEN[name]
    if engine in ("pdl", "graph", "hints", "inplace+tb2d", "inplace+tb2d+hints", "inplace+tb2d+chains+graph") and c["steps"] > 1000:
        pytest.skip("10,000-step cases: not under the plain pdl/graph engines, nor under the deliberately tiny 2-D tiles "
                    "(tens of millions of launches); the in-place defaults cover temporal blocking at full length")
    fA, walls = cases.make_inputs(name)
    ctor, tune = _split(ENGINES[engine])
    with _lattice(c, **ctor) as lat:
        lat.set_tuning(**tune)
        lat.upload(fA, walls)
        lat.run(c["steps"])
        a = lat.download("A")
        b = None if ctor else lat.download("B")
        assert cases.digest(a) == g["fA_sha256"], f"fA differs from the reference after {c['steps']} steps"
        if b is not None:
            assert cases.digest(b) == g["fB_sha256"], "fB differs (odd step with its source applied, LB.c:201)"
        assert float(lat.ComputeReynolds()).hex() == g["reynolds_hex"]
        _assert_same(lat.download_walls(), walls, "wall mask round trip")
        if c.get("full"):
            z = np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"))
            _assert_same(a, z["fA"], "fA")
            if b is not None:
                _assert_same(b, z["fB"], "fB")


# ---------------------------------------------------------------------------------------------
# against the oracle, over kernel variants and awkward shapes
# ---------------------------------------------------------------------------------------------
SHAPES = [
    dict(nx=3, ny=8), dict(nx=1, ny=16), dict(nx=2, ny=9), dict(nx=5, ny=33), dict(nx=40, ny=64),
    dict(nx=33, ny=257), dict(nx=64, ny=1000), dict(nx=7, ny=4099),
]


@pytest.mark.parametrize("prec", ["dp", "sp"])
@pytest.mark.parametrize("isa", ["avx", "sse", "serial"])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: f"{s['nx']}x{s['ny']}")
def test_cuda_matches_oracle_shapes(prec, isa, shape):
    c = dict(prec=prec, isa=isa, scene="random", perturb=31, steps=8, **shape)
    fA, walls = cases.make_inputs(c)
    if c["nx"] < 3:
        walls[:] = 0                           # no channel rows when the lattice is 1-2 rows high
        walls[0, ::5] = 1
    o = _oracle(c, fA, walls, c["steps"])
    for engine, ekw in ENGINES.items():
        ctor, tune = _split(ekw)
        with _lattice(c, **ctor) as lat:
            lat.set_tuning(**tune)
            lat.upload(fA, walls)
            lat.run(c["steps"])
            _assert_same(lat.download("A"), o.f_view("A"), f"fA engine={engine}")
            if not ctor:
                _assert_same(lat.download("B"), o.f_view("B"), f"fB engine={engine}")


def _variants(prec):
    vecs = (1, 2, 4)
    out = []
    for v in vecs:
        for hints in (0, 1):
            out.append(dict(vec=v, hints=hints))
    out += [dict(vec=1, hints=0, pdl=1, graph=2), dict(vec=vecs[-1], hints=1, pdl=1, graph=0), dict(vec=2, hints=0, pdl=0, graph=1)]
    return out


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_every_kernel_variant_is_bit_identical(prec):
    c = dict(prec=prec, isa="avx", nx=96, ny=520, scene="random", perturb=17, steps=10)
    fA, walls = cases.make_inputs(c)
    o = _oracle(c, fA, walls, c["steps"])
    with _lattice(c) as lat:
        for tune in _variants(prec):
            lat.set_tuning(**dict(dict(pdl=0, graph=0), **tune))
            lat.upload(fA, walls)
            lat.run(c["steps"])
            _assert_same(lat.download("A"), o.f_view("A"), f"fA with {tune}")
            _assert_same(lat.download("B"), o.f_view("B"), f"fB with {tune}")
    # the in-place kernels: every vector width, a few CTA shapes, PDL/graph
    vecs = (1, 2, 4)
    with _lattice(c, inplace=True) as lat:
        for v in vecs:
            for extra in (dict(), dict(pdl=1, graph=2), dict(pdl=1), dict(tb_depth=4, tb_rows=11, tb_cols=40), dict(hints=1, tb_depth=2, tb_rows=30)):
                tune = dict(dict(vec=v, hints=0, pdl=0, graph=0, tb_depth=0, tb_rows=0, tb_cols=0), **extra)
                lat.set_tuning(**tune)
                lat.upload(fA, walls)
                lat.run(4)
                lat.run(c["steps"] - 4)
                _assert_same(lat.download("A"), o.f_view("A"), f"in-place fA with {tune}")
        # wavefront schedules: depth k, band height H (0 = automatic), with and without graph replay, with and without
        # evict-first stores on the last level
        for k, h, g, w in ((2, 0, 0, 0), (4, 7, 0, 0), (8, 16, 1, 0), (46, 3, 0, 0), (10, 200, 1, 0), (4, 1, 1, 0), (6, 90, 0, 0),
                           (4, 20, 0, 128), (6, 0, 1, 104), (2, 7, 0, 64), (8, 30, 1, 8), (16, 40, 0, 200), (2, 95, 1, 256)):
            tune = dict(vec=vecs[-1] if w == 0 else vecs[(k // 2) % len(vecs)], hints=(k // 2) & 1,
                        pdl=g, graph=g, tb_depth=k, tb_rows=h, tb_cols=w)
            lat.set_tuning(**tune)
            lat.upload(fA, walls)
            lat.run(c["steps"])
            _assert_same(lat.download("A"), o.f_view("A"), f"wavefront fA with {tune}")
            lat.upload(fA, walls)
            lat.run(2)
            lat.run(c["steps"] - 2)
            _assert_same(lat.download("A"), o.f_view("A"), f"wavefront fA (2 + {c['steps'] - 2} steps) with {tune}")


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_staged_kernel_is_bit_identical(prec):
    """The experiment build (latticeboltzmann_b200/exp/, include/lbm_b200_experimental.h): the in-place step with the next
    row staged in shared memory by cp.async (stage_smem 1) or, for whole-row launches of phase 2, by TMA bulk copies
    (stage_smem 2).  north_star's "TMA or shared-memory staging": bit-identical, measured slower (profiles/r02_run1)."""
    c = dict(prec=prec, isa="avx", nx=96, ny=520, scene="random", perturb=17, steps=10)
    fA, walls = cases.make_inputs(c)
    o = _oracle(c, fA, walls, c["steps"])
    with _lattice(c, inplace=True, experimental=True) as lat:
        for rpt, mode, extra in ((2, 1, {}), (5, 1, dict(tb_depth=4, tb_rows=9)), (3, 1, dict(pdl=1, graph=2)), (96, 1, {}), (7, 1, dict(block=64)),
                                 (4, 1, dict(block=256)), (2, 1, dict(tb_depth=4, tb_rows=20, tb_cols=128, pdl=1)), (1000, 1, dict(block=32)),
                                 # 2: phase 2 through TMA bulk copies (whole rows), phase 1 through cp.async
                                 (2, 2, {}), (3, 2, dict(tb_depth=4, tb_rows=9, pdl=1)), (96, 2, dict(block=64)),
                                 (7, 2, dict(block=32, pdl=1, graph=2)), (4, 2, dict(tb_depth=4, tb_rows=20, tb_cols=128))):
            tune = dict(dict(vec=1, hints=0, pdl=0, graph=0, tb_depth=0, tb_rows=0, tb_cols=0), **extra)
            blk = tune.pop("block", 0)                        # the CTA size is a knob of the experiment build
            lat.set_tuning(**tune)
            lat.set_experimental(stage_smem=mode, rows_per_thread=rpt, block=blk)
            lat.upload(fA, walls)
            l0 = lat.launches
            lat.run(4)
            lat.run(c["steps"] - 4)
            _assert_same(lat.download("A"), o.f_view("A"), f"staged in-place fA with {tune}, stage_smem {mode}, rows_per_thread {rpt}")
            assert lat.launches > l0
        lat.set_experimental(stage_smem=0, rows_per_thread=0)        # and off again: the product kernels of the same build
        lat.upload(fA, walls)
        lat.run(c["steps"])
        _assert_same(lat.download("A"), o.f_view("A"), "experiment build, staging off")


@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_reference_function_surface(prec):
    """DoTimeStep == ApplySource/StreamCollide x2 (LB.c:186-210); InitializeArrays on the device ==
    the host scene; ComputeReynolds, macroscopic fields and PrintLattice as the oracle computes them."""
    from latticeboltzmann_b200 import Lattice
    from oracle.harness import Oracle
    nx, ny = 230, 136
    o = Oracle(prec, nx, ny)
    o.initialize()
    with Lattice(nx, ny, prec) as a, Lattice(nx, ny, prec) as b:
        a.InitializeArrays()
        _assert_same(a.download("A"), o.f_view("A"), "InitializeArrays fA")
        _assert_same(a.download_walls(), o.walls_view(), "InitializeArrays walls")
        b.upload(o.f_view("A"), o.walls_view())
        for _ in range(3):
            a.DoTimeStep()
            b.ApplySource("A"); b.StreamCollide("A"); b.ApplySource("B"); b.StreamCollide("B")
        o.run(6)
        _assert_same(a.download("A"), o.f_view("A"), "DoTimeStep x3")
        _assert_same(b.download("A"), o.f_view("A"), "ApplySource/StreamCollide x6")
        _assert_same(a.download("B"), o.f_view("B"), "fB")
        assert a.ComputeReynolds() == o.reynolds()
        for got, want, nm in zip(a.macroscopic(), o.macroscopic(), ("rho", "ux", "uy")):
            _assert_same(got, want, nm)


def test_print_lattice_bytes(tmp_path):
    from latticeboltzmann_b200 import Lattice
    from oracle.harness import Oracle
    for prec in ("dp", "sp"):
        nx, ny = 230, 120
        o = Oracle(prec, nx, ny)
        o.initialize()
        o.run(40)
        with Lattice(nx, ny, prec) as lat:
            lat.InitializeArrays()
            lat.run(40)
            lat.PrintLattice(tmp_path / f"gpu_{prec}.csv")
        o.print_lattice(str(tmp_path / f"cpu_{prec}.csv"))
        assert (tmp_path / f"gpu_{prec}.csv").read_bytes() == (tmp_path / f"cpu_{prec}.csv").read_bytes()


def test_print_lattice_over_slabs_and_driver_dump(tmp_path):
    """SAVELATTICE of a multi-slab run (LB.c:155-159): lbm_print_lattice_group writes ONE file for the whole lattice, identical
    to the single-slab file; the C driver's DUMP_EVERY writes data/<step>.csv at the reference's observation points."""
    from latticeboltzmann_b200 import Lattice, group_run, make_ring, print_lattice_group
    nx, ny = 230, 120
    with Lattice(nx, ny, "dp") as lat:
        lat.InitializeArrays()
        lat.run(40)
        lat.PrintLattice(tmp_path / "one.csv")
    for inplace in (False, True):
        slabs = make_ring(nx, ny, 3, "dp", inplace=inplace)
        try:
            for s in slabs:
                s.InitializeArrays()
            group_run(slabs, 40)
            print_lattice_group(slabs, tmp_path / "ring.csv")
        finally:
            for s in slabs:
                s.close()
        assert (tmp_path / "ring.csv").read_bytes() == (tmp_path / "one.csv").read_bytes()
    (tmp_path / "data").mkdir()
    exe = os.path.join(ROOT, "driver", "lbm_b200_dp")
    out = subprocess.run([exe, str(nx), str(ny), "80", "1", "0", "20"], capture_output=True, text=True, timeout=300, cwd=tmp_path)
    assert out.returncode == 0, out.stderr
    assert sorted(os.listdir(tmp_path / "data")) == ["0.csv", "20.csv", "40.csv", "60.csv"]
    assert (tmp_path / "data" / "40.csv").read_bytes() == (tmp_path / "one.csv").read_bytes()
    out = subprocess.run([exe, str(nx), str(ny), "80", "1", "0", "5"], capture_output=True, text=True, timeout=300, cwd=tmp_path)   # odd: every 10
    assert out.returncode == 0 and "70.csv" in os.listdir(tmp_path / "data") and "5.csv" not in os.listdir(tmp_path / "data")


def test_error_behaviour():
    from latticeboltzmann_b200 import Lattice, LbmError
    with Lattice(8, 16) as lat:
        with pytest.raises(LbmError):
            lat.run(2)                                   # nothing loaded
        f, w = np.zeros((9, 8, 16)), np.zeros((8, 16), dtype=np.int8)
        w[3, 3] = 4
        with pytest.raises(LbmError):
            lat.upload(f + 0.1, w)                       # mask values: 0 fluid, 1 bounce-back, 2 / 3 reflect
        w[3, 3] = 1
        lat.upload(f + 0.1, w)
        with pytest.raises(LbmError):
            lat.run(3)                                   # odd step count (LB.c:148)
        lat.run(0)
    with pytest.raises(LbmError):
        Lattice(0, 16)
    # an in-place context has one buffer: everything that names buffer B or a single step is refused
    with Lattice(8, 16, inplace=True) as lat:
        f, w = np.zeros((9, 8, 16)) + 0.1, np.zeros((8, 16), dtype=np.int8)
        with pytest.raises(LbmError):
            lat.upload(f, w, fB=f)
        lat.upload(f, w)
        lat.run(2)
        for bad in (lambda: lat.download("B"), lambda: lat.StreamCollide("A"), lambda: lat.ApplySource("B"), lambda: lat.run(1)):
            with pytest.raises(LbmError):
                bad()
        lat.ApplySource("A")
        assert lat.slab_info().dev_f[1] is None


# ---------------------------------------------------------------------------------------------
# decomposition invariance: N logical slabs on one GPU == one slab, bit for bit
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("inplace", [False, True, dict(tb_depth=4, tb_rows=7, tb_cols=128), dict(tb_depth=8, tb_rows=0, tb_cols=0, pdl=1)],
                         ids=["ab", "inplace", "inplace+tb2d", "inplace+tb"])
@pytest.mark.parametrize("prec,nslabs", [("dp", 2), ("dp", 3), ("dp", 8), ("sp", 4)])
def test_slab_ring_equals_single_slab(prec, nslabs, inplace):
    from latticeboltzmann_b200 import Lattice, group_run, make_ring
    c = dict(prec=prec, isa="avx", nx=200, ny=1000, scene="cylinder", perturb=None, steps=60)
    fA, walls = cases.make_inputs(c)
    with Lattice(c["nx"], c["ny"], prec) as one:
        one.upload(fA, walls)
        one.run(c["steps"])
        want_a, want_b, want_re = one.download("A"), one.download("B"), one.ComputeReynolds()
    ring = make_ring(c["nx"], c["ny"], nslabs, prec, inplace=bool(inplace))
    try:
        for s in ring:
            if isinstance(inplace, dict):
                s.set_tuning(**inplace)                   # temporal blocking across the seams of lbm_group_run
            s.upload(fA, walls)
        group_run(ring, 20)
        group_run(ring, 40)
        got_a, got_b = np.zeros_like(want_a), np.zeros_like(want_b)
        for s in ring:
            s.download("A", out=got_a)
            if not inplace:
                s.download("B", out=got_b)
        _assert_same(got_a, want_a, "fA over the ring")
        if not inplace:
            _assert_same(got_b, want_b, "fB over the ring")
        cols = [s.reynolds_column() for s in ring]
        re = ring[0].reynolds_from_column(np.concatenate([u for u, _ in cols]), np.concatenate([f for _, f in cols]))
        assert re == want_re
    finally:
        for s in ring:
            s.close()


def test_uneven_slabs_and_single_row_slabs():
    from latticeboltzmann_b200 import Lattice, group_run, make_ring
    c = dict(prec="dp", isa="avx", nx=11, ny=64, scene="random", perturb=9, steps=10)
    fA, walls = cases.make_inputs(c)
    o = _oracle(c, fA, walls, c["steps"])
    for bounds, inplace in [(b, ip) for b in ([(0, 1), (1, 2), (2, 11)], [(0, 5), (5, 6), (6, 11)], [(k, k + 1) for k in range(11)])
                            for ip in (False, True, dict(tb_depth=2, tb_rows=2, tb_cols=16), dict(tb_depth=4, tb_rows=0, tb_cols=0))]:
        ring = make_ring(c["nx"], c["ny"], len(bounds), "dp", bounds=bounds, inplace=bool(inplace))
        try:
            for s in ring:
                if isinstance(inplace, dict):
                    s.set_tuning(**inplace)               # slabs too low for the row pyramid: whole-slab launches per event
                s.upload(fA, walls)
            group_run(ring, c["steps"])
            got = np.zeros_like(fA)
            for s in ring:
                s.download("A", out=got)
            _assert_same(got, o.f_view("A"), f"ring {bounds} inplace={inplace}")
        finally:
            for s in ring:
                s.close()


# ---------------------------------------------------------------------------------------------
# full BASELINE sizes: properties that need no CPU run
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("inplace", [False, True], ids=["ab", "inplace"])
@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_full_size_properties_400x4000(prec, inplace):
    """BASELINE config 2 at its full 10,000 steps: mass is conserved to rounding (source and
    bounce-back are conservative), wall-site f0 keeps its initial value (LB.c:246-255), the flow is
    finite and positive, and two chunked runs equal one long run bit for bit."""
    from latticeboltzmann_b200 import Lattice
    nx, ny, steps = 400, 4000, 10000
    with Lattice(nx, ny, prec, inplace=inplace) as a, Lattice(nx, ny, prec) as b:
        a.InitializeArrays()
        b.InitializeArrays()
        f0 = a.download("A")
        walls = a.download_walls().astype(bool)
        a.run(steps)
        for _ in range(10):
            b.run(steps // 10)
        fa, fb = a.download("A"), b.download("A")
        assert fa.tobytes() == fb.tobytes()
        assert np.isfinite(fa).all() and (fa[:, ~walls] > 0).all()
        m0, m1 = f0.astype(np.float64).sum(), fa.astype(np.float64).sum()
        assert abs(m1 - m0) / m0 < (1e-11 if prec == "dp" else 2e-3)
        np.testing.assert_array_equal(fa[0][walls], f0[0][walls])
        assert np.isfinite(a.ComputeReynolds())


def test_c_driver_prints_reference_reynolds():
    """driver/main.c (the reference's main over the C ABI) on the shipped scenario: the final line's
    Re must be the compiled reference's value to all printed digits (LB.c:173, %.10le)."""
    for tag, case, extra in (("dp", "builtin_dp_10000", []), ("sp", "builtin_sp_10000", []),
                             ("dp", "builtin_dp_10000", ["1", "0", "0", "0"]), ("sp", "builtin_sp_10000", ["1", "0", "0", "0"])):   # INPLACE=0: fA/fB pair
        exe = os.path.join(ROOT, "driver", f"lbm_b200_{tag}")
        out = subprocess.run([exe, "400", "2000", "10000"] + extra, capture_output=True, text=True, timeout=300)
        assert out.returncode == 0, out.stderr
        last = out.stdout.strip().splitlines()[-1]
        want = "%.10le" % float.fromhex(GOLDEN[case]["reynolds_hex"])
        assert last.startswith("Runtime:") and last.endswith("Re " + want), last
