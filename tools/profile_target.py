"""Short, deterministic workload for ncu: a few steps of the step kernel on the BASELINE lattices.
usage: profile_target.py [dp|sp] [small|large|c3] [nsteps] [ab|inplace|inplace:<hints>|staged:<mode>:<rows_per_thread>]
(inplace = one buffer, AA kernels, library-default temporal blocking -- what bench.py's headline runs)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "dp"
size = sys.argv[2] if len(sys.argv) > 2 else "small"
nsteps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
engine = sys.argv[4] if len(sys.argv) > 4 else "ab"
nx, ny = {"small": (400, 4000), "large": (8192, 32768), "c3": (800, 4000)}[size]
with Lattice(nx, ny, prec, inplace=(engine != "ab"), experimental=engine.startswith("staged")) as lat:
    lat.InitializeArrays()
    if engine.startswith("inplace"):
        lat.set_tuning(graph=0)          # individual launches, so that ncu sees every kernel
        if ":" in engine:                # L2 eviction-priority hints of the temporal-blocking levels (lbm_tuning.hints)
            lat.set_tuning(graph=0, hints=int(engine.split(":")[1]))
    if engine.startswith("staged"):      # experiment build: next row staged in shared memory (cp.async / TMA), no temporal blocking
        _, mode, rpt = engine.split(":")
        lat.set_tuning(graph=0, pdl=1, tb_depth=0, vec=1)
        lat.set_experimental(stage_smem=int(mode), rows_per_thread=int(rpt))
    l0 = lat.launches
    ms = lat.run_timed(nsteps)
    print(f"{prec} {nx}x{ny} {engine}: {nsteps} steps in {ms:.3f} ms -> {nx * ny * nsteps / ms / 1e3:.1f} MLUPS, "
          f"{lat.launches - l0} launches, tuning {lat.get_tuning()}")
