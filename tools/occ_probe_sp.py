"""SP in-place phases at a given register cap (experiment libraries via LBM_B200_LIBDIR): kernel-only timing per phase
(tools/phase_probe.py's method) and the real run."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

tag = os.environ.get("LBM_B200_LIBDIR", "default").rstrip("/").split("/")[-1]
nx, ny = 400, 4000
with Lattice(nx, ny, "sp", inplace=True) as lat:
    lat.InitializeArrays()
    lat.run(10)
    for vec in (2, 4):
        lat.set_tuning(vec=vec, vec_phase1=0, pdl=1, graph=0, tb_depth=0)
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
            print(json.dumps(dict(lib=tag, vec=vec, phase=phase + 1, us_per_launch=round(dt / n * 1e6, 2))), flush=True)
    for v1 in (0, 2):
        lat.set_tuning(vec=4, vec_phase1=v1, pdl=1, graph=10, tb_depth=0)
        lat.InitializeArrays()
        lat.run_timed(400)
        ms = min(lat.run_timed(4000) for _ in range(3))
        print(json.dumps(dict(lib=tag, vec=4, vec_phase1=v1, mlups=round(nx * ny * 4000 / ms / 1e3, 1))), flush=True)
