"""Deterministic input recipes shared by tests/golden/make_golden.py (which runs the compiled
reference) and by the parity tests (which run the oracle and the CUDA path on the same inputs)."""
import hashlib

import numpy as np

from latticeboltzmann_b200 import geometry

# name -> recipe.  `scene`: builtin (InitializeArrays), cylinder, random (3 % random solids + channel rows);
# `perturb`: LCG seed for a +-5 % density perturbation, or None.
CASES = {
    # small enough to commit the full output arrays (tests/golden/*.npz)
    "tiny_dp_avx":      dict(prec="dp", isa="avx", nx=24, ny=40, steps=20, scene="random", perturb=7, full=True),
    "tiny_sp_avx":      dict(prec="sp", isa="avx", nx=24, ny=40, steps=20, scene="random", perturb=7, full=True),
    "tiny_dp_ragged":   dict(prec="dp", isa="avx", nx=17, ny=43, steps=10, scene="random", perturb=3, full=True),
    "tiny_sp_ragged":   dict(prec="sp", isa="avx", nx=17, ny=43, steps=10, scene="random", perturb=3, full=True),
    # digests only
    "builtin_dp_100":   dict(prec="dp", isa="avx", nx=400, ny=2000, steps=100, scene="builtin", perturb=None),
    "builtin_sp_100":   dict(prec="sp", isa="avx", nx=400, ny=2000, steps=100, scene="builtin", perturb=None),
    "builtin_dp_1000":  dict(prec="dp", isa="avx", nx=400, ny=2000, steps=1000, scene="builtin", perturb=None),
    "builtin_sp_1000":  dict(prec="sp", isa="avx", nx=400, ny=2000, steps=1000, scene="builtin", perturb=None),
    "builtin_dp_10000": dict(prec="dp", isa="avx", nx=400, ny=2000, steps=10000, scene="builtin", perturb=None),
    "builtin_sp_10000": dict(prec="sp", isa="avx", nx=400, ny=2000, steps=10000, scene="builtin", perturb=None),
    "c2_dp_200":        dict(prec="dp", isa="avx", nx=400, ny=4000, steps=200, scene="builtin", perturb=None),
    "c2_sp_200":        dict(prec="sp", isa="avx", nx=400, ny=4000, steps=200, scene="builtin", perturb=None),
    # BASELINE.json configs 2 and 3 at their full 10,000 steps
    "c2_dp_10000":      dict(prec="dp", isa="avx", nx=400, ny=4000, steps=10000, scene="builtin", perturb=None),
    "c2_sp_10000":      dict(prec="sp", isa="avx", nx=400, ny=4000, steps=10000, scene="builtin", perturb=None),
    "cyl_dp_10000":     dict(prec="dp", isa="avx", nx=800, ny=4000, steps=10000, scene="cylinder", perturb=None),
    "cyl_dp_200":       dict(prec="dp", isa="avx", nx=800, ny=4000, steps=200, scene="cylinder", perturb=None),
    "pert_dp_sse_60":   dict(prec="dp", isa="sse", nx=230, ny=126, steps=60, scene="builtin", perturb=11),
    "pert_sp_sse_60":   dict(prec="sp", isa="sse", nx=230, ny=124, steps=60, scene="builtin", perturb=11),
    "pert_dp_serial":   dict(prec="dp", isa="serial", nx=64, ny=96, steps=40, scene="random", perturb=5),
    "pert_sp_serial":   dict(prec="sp", isa="serial", nx=64, ny=96, steps=40, scene="random", perturb=5),
    "ragged_dp_avx":    dict(prec="dp", isa="avx", nx=230, ny=131, steps=40, scene="builtin", perturb=13),
    "ragged_sp_avx":    dict(prec="sp", isa="avx", nx=230, ny=131, steps=40, scene="builtin", perturb=13),
}

CYLINDER = dict(ci_frac=(1, 2), cj_frac=(1, 5), r_frac=(1, 10))   # centre (nx/2, ny/5), radius nx/10


def cylinder_params(nx, ny):
    return nx // 2, ny // 5, nx // 10


def random_walls(nx, ny, seed):
    """channel rows + ~3 % solid sites from the integer LCG (no floating point in the recipe)."""
    walls = geometry.channel_walls(nx, ny, barrier=False)
    u = geometry.perturb_lcg(np.ones((nx, ny)), seed=seed + 1000, amplitude=1.0)   # in [0,2)
    walls[u < 0.06] = 1
    return walls


def make_inputs(case):
    c = CASES[case] if isinstance(case, str) else case
    nx, ny, prec = c["nx"], c["ny"], c["prec"]
    fA = geometry.equilibrium_rest(nx, ny, prec)
    if c["scene"] == "builtin":
        walls = geometry.channel_walls(nx, ny)
    elif c["scene"] == "cylinder":
        walls = geometry.cylinder_walls(nx, ny, *cylinder_params(nx, ny))
    elif c["scene"] == "random":
        walls = random_walls(nx, ny, c["perturb"] or 0)
    else:
        raise ValueError(c["scene"])
    if c["perturb"] is not None:
        fA = geometry.perturb_lcg(fA, seed=c["perturb"])
    return fA, walls


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
