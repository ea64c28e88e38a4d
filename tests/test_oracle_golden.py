"""CPU: the oracle restatement (oracle/lbm_oracle.c) against the golden vectors produced by the
compiled reference (tests/golden/make_golden.py), and -- where oracle/_ref is present -- against
the reference's own functions directly.  Bit-exact in SP and DP."""
import json
import os

import numpy as np
import pytest

from oracle.harness import VECWIDTH, Oracle, Ref, ref_available
from tests import cases

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden.json")))
ORACLE_BUDGET = 1.0e8    # site updates the CPU suite spends per case


def _small(name):
    c = cases.CASES[name]
    return c["nx"] * c["ny"] * c["steps"] <= ORACLE_BUDGET


def _run_oracle(name):
    c = cases.CASES[name]
    fA, walls = cases.make_inputs(name)
    o = Oracle(c["prec"], c["nx"], c["ny"], VECWIDTH[(c["prec"], c["isa"])])
    o.set_state(fA, walls)
    o.run(c["steps"])
    return o


def test_golden_file_covers_all_cases():
    assert set(GOLDEN) == set(cases.CASES)
    for name, g in GOLDEN.items():
        fA, walls = cases.make_inputs(name)
        assert cases.digest(fA) == g["input_sha256"], "input recipe drifted from the golden file"
        assert cases.digest(walls) == g["walls_sha256"]


@pytest.mark.parametrize("name", [n for n in cases.CASES if _small(n)])
def test_oracle_matches_golden(name):
    o, g = _run_oracle(name), GOLDEN[name]
    assert cases.digest(o.f_view("A")) == g["fA_sha256"]
    assert cases.digest(o.f_view("B")) == g["fB_sha256"]
    assert float(o.reynolds()).hex() == g["reynolds_hex"]


@pytest.mark.parametrize("name", [n for n, c in cases.CASES.items() if c.get("full")])
def test_oracle_matches_golden_arrays(name):
    o = _run_oracle(name)
    z = np.load(os.path.join(os.path.dirname(__file__), "golden", name + ".npz"))
    np.testing.assert_array_equal(o.f_view("A"), z["fA"])
    np.testing.assert_array_equal(o.f_view("B"), z["fB"])


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (reference checkout absent)")
@pytest.mark.parametrize("prec", ["dp", "sp"])
@pytest.mark.parametrize("isa", ["avx", "sse", "serial"])
def test_oracle_matches_compiled_reference(prec, isa):
    """InitializeArrays, ApplySource, DoTimeStep and ComputeReynolds of the reference itself."""
    nx, ny = 231, 133                       # ragged for every VECWIDTH
    r, o = Ref(prec, nx, ny, isa), Oracle(prec, nx, ny, VECWIDTH[(prec, isa)])
    r.initialize()
    o.initialize()
    assert r.fA.tobytes() == o.fA.tobytes() and r.walls.tobytes() == o.walls.tobytes()
    r.apply_source("A")
    o.apply_source("A")
    assert r.fA.tobytes() == o.fA.tobytes()
    r.run(30)
    o.run(30)
    assert r.f_view("A").tobytes() == o.f_view("A").tobytes()
    assert r.f_view("B").tobytes() == o.f_view("B").tobytes()
    assert r.reynolds() == o.reynolds()


@pytest.mark.skipif(not ref_available(), reason="oracle/_ref not built (reference checkout absent)")
@pytest.mark.parametrize("prec", ["dp", "sp"])
def test_oracle_print_lattice_matches_compiled_reference(prec, tmp_path, monkeypatch):
    """PrintLattice (LB.c:610-639) of the reference itself, byte for byte: pins the CSV leg of the oracle, which is what
    lbm_print_lattice is compared with on the GPU (tests/test_gpu_parity.py::test_print_lattice_bytes)."""
    nx, ny = 231, 133
    r, o = Ref(prec, nx, ny, "avx"), Oracle(prec, nx, ny, VECWIDTH[(prec, "avx")])
    r.initialize()
    o.initialize()
    r.run(40)
    o.run(40)
    (tmp_path / "data").mkdir()
    monkeypatch.chdir(tmp_path)                       # the reference writes data/<step>.csv relative to the cwd (LB.c:613)
    r.print_lattice(40)
    assert o.print_lattice(str(tmp_path / "oracle.csv")) == 0
    ref_bytes = (tmp_path / "data" / "40.csv").read_bytes()
    assert len(ref_bytes) > nx * ny * 12 and ref_bytes.count(b"\n") == nx
    assert (tmp_path / "oracle.csv").read_bytes() == ref_bytes
    rho, ux, uy = o.macroscopic()
    assert np.any(ux * ux + uy * uy > 0)            # the file is not all zeros


def test_known_answer_reynolds_10k():
    """The reference's only printed physics number (LB.c:173) after the shipped 10,000 steps,
    400x2000: strict -O2 AVX builds, measured with the compiled reference."""
    assert GOLDEN["builtin_dp_10000"]["reynolds"].startswith("8.5513048345592")
    assert GOLDEN["builtin_sp_10000"]["reynolds"].startswith("8.5506")


def test_oracle_conserves_mass_and_wall_f0():
    name = "pert_dp_sse_60"
    c = cases.CASES[name]
    fA, walls = cases.make_inputs(name)
    o = _run_oracle(name)
    out = o.f_view("A").astype(np.float64)
    # bounce-back + BGK + source conserve mass; wall sites never get f0 written (LB.c:246-255)
    assert abs(out.sum() - fA.astype(np.float64).sum()) < 1e-9 * fA.sum()
    w = walls.astype(bool)
    np.testing.assert_array_equal(out[0][w], fA[0][w])
