"""One in-place phase repeated (kernel-only, as tools/phase_probe.py), for the shared-memory-staged kernels:
stage_smem = 1 (cp.async, both phases) and 2 (TMA bulk copies, phase 2) against the one-row-per-thread default."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402


def timeit(lat, phase, nx, n=1500):
    for _ in range(100):
        lat._call("lbm_step_rows", phase, 0, nx, 1, 0)
    lat.sync()
    t0 = time.perf_counter()
    for _ in range(n):
        lat._call("lbm_step_rows", phase, 0, nx, 1, 0)
    lat.sync()
    return (time.perf_counter() - t0) / n


out = []
for prec, nx, ny in (("dp", 400, 2000), ("sp", 400, 4000)):
    with Lattice(nx, ny, prec, inplace=True, experimental=True) as lat:
        lat.InitializeArrays()
        lat.run(10)
        combos = [dict(vec=1, rows_per_thread=0, stage_smem=0, block=128)]
        if prec == "sp":
            combos.append(dict(vec=4, rows_per_thread=0, stage_smem=0, block=128))
        for blk in (128, 256):
            for rpt in (2, 4, 8, 16):
                combos.append(dict(vec=1, rows_per_thread=rpt, stage_smem=1, block=blk))
        for blk in (64, 128):
            for rpt in (4, 8, 16, 32):
                combos.append(dict(vec=1, rows_per_thread=rpt, stage_smem=2, block=blk))
        for cmb in combos:
            lat.set_tuning(pdl=1, graph=0, tb_depth=0, vec=cmb['vec'])
            lat.set_experimental(stage_smem=cmb['stage_smem'], rows_per_thread=cmb['rows_per_thread'], block=cmb['block'])
            for phase in (0, 1):
                if cmb["stage_smem"] == 2 and phase == 0:
                    continue
                dt = timeit(lat, phase, nx)
                rec = dict(prec=prec, nx=nx, ny=ny, phase=phase + 1, us=round(dt * 1e6, 2), glups=round(nx * ny / dt / 1e9, 1), **cmb)
                out.append(rec)
                print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "stage_phase_probe.json"), "w"), indent=0)
