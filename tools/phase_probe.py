"""Kernel-only timing of ONE in-place phase, repeated (not a physical run: the same phase again and again; values stay
finite because each application is a mass-conserving relaxation).  Separates phase 1 (neighbour loads/stores) from
phase 2 (local) and shows what rows_per_thread (software pipelining) does to each."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

for prec, nx, ny in (("dp", 400, 2000), ("sp", 400, 4000)):
    with Lattice(nx, ny, prec, inplace=True) as lat:
        lat.InitializeArrays()
        lat.run(10)
        for vec in ((1, 2, 4) if prec == "dp" else (2, 4, 8)):
            for rpt in (0, -1):                       # -1: aligned accesses + warp shuffles (neighbour = 0)
                if vec == 1 and rpt:
                    continue
                lat.set_tuning(vec=vec, rows_per_thread=max(rpt, 0), neighbour=0 if rpt < 0 else 1, pdl=1, graph=0, tb_depth=0)
                for phase in (0, 1):
                    n = 3000
                    for _ in range(200):
                        lat._call("lbm_step_rows", phase, 0, nx, 1, 0)
                    lat.sync()
                    t0 = time.perf_counter()
                    for _ in range(n):
                        lat._call("lbm_step_rows", phase, 0, nx, 1, 0)
                    lat.sync()
                    dt = time.perf_counter() - t0
                    print(json.dumps(dict(prec=prec, nx=nx, ny=ny, vec=vec, rows_per_thread=rpt, phase=phase + 1,
                                          us_per_launch=round(dt / n * 1e6, 2), glups=round(nx * ny * n / dt / 1e9, 1))), flush=True)
