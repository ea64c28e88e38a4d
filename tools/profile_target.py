"""Short, deterministic workload for ncu: a few steps of the step kernel on the BASELINE lattices.
usage: profile_target.py [dp|sp] [small|large] [nsteps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "dp"
size = sys.argv[2] if len(sys.argv) > 2 else "small"
nsteps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
nx, ny = (400, 4000) if size == "small" else (8192, 32768)
with Lattice(nx, ny, prec) as lat:
    lat.InitializeArrays()
    ms = lat.run_timed(nsteps)
    print(f"{prec} {nx}x{ny}: {nsteps} steps in {ms:.3f} ms -> {nx * ny * nsteps / ms / 1e3:.1f} MLUPS, tuning {lat.get_tuning()}")
