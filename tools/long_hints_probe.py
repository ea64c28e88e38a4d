"""Round 2: do the L2 eviction hints (half the DRAM traffic) pay in a LONG, power-capped run?  1000 timesteps per measurement on
the weak-scaling unit, hints 0 / 3 alternating.   usage: long_hints_probe.py [dp|sp|both]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

precs = ("dp", "sp") if len(sys.argv) < 2 or sys.argv[1] == "both" else (sys.argv[1],)
nx, ny, T = 8192, 32768, 1000
for prec in precs:
    with Lattice(nx, ny, prec, inplace=True) as lat:
        lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        lat.run_timed(64)
        for rep in range(2):
            for hints in (0, 3):
                lat.set_tuning(hints=hints)
                ms = lat.run_timed(T)
                print(json.dumps(dict(prec=prec, hints=hints, rep=rep, timesteps=T, mlups=round(nx * ny * T / ms / 1e3, 1))), flush=True)
