"""How fast does the step run when A+B fit the L2?  Lattices of decreasing size, same kernels.
Tells what an in-place (AA) layout -- half the working set -- could buy on the BASELINE small lattices."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

B = {"dp": 144, "sp": 72}
for prec in ("dp", "sp"):
    for nx, ny in ((800, 4000), (400, 4000), (400, 2000), (200, 2000), (400, 1000), (200, 1000), (100, 1000)):
        ws = 2 * 9 * nx * ny * (8 if prec == "dp" else 4) / 1e6
        T = 4000
        best = None
        for tune in (dict(pdl=0, graph=0), dict(pdl=1, graph=10)):
            for vec in ((1, 2) if prec == "dp" else (2, 4)):
                with Lattice(nx, ny, prec) as lat:
                    lat.InitializeArrays()
                    lat.set_tuning(vec=vec, **tune)
                    lat.run_timed(400)
                    ms = min(lat.run_timed(T) for _ in range(2))
                rec = dict(prec=prec, nx=nx, ny=ny, working_set_mb=round(ws, 1), vec=vec, **tune, us_per_step=round(ms * 1e3 / T, 2),
                           mlups=round(nx * ny * T / ms / 1e3, 1), gbs=round(nx * ny * T / ms / 1e3 * B[prec] / 1e3, 1))
                if best is None or rec["mlups"] > best["mlups"]:
                    best = rec
        print(json.dumps(best), flush=True)
