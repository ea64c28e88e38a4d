"""lbm_tuning.stage_smem (next row staged in shared memory by cp.async, stream_collide_aa_staged_kernel; 2 = TMA bulk copies in
phase 2) against the one-row-per-thread kernels: bits first, then MLUPS over rows_per_thread."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

out = []
for prec, nx, ny, T in (("dp", 400, 2000, 4000), ("dp", 400, 4000, 2000), ("sp", 400, 4000, 2000), ("dp", 8192, 32768, 48)):
    with Lattice(nx, ny, prec, inplace=True, experimental=True) as lat:
        def init():
            if nx <= 2000:
                lat.InitializeArrays()
            else:
                lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        init()
        d = lat.get_tuning()
        lat.run_timed(min(T, 400))
        ms = min(lat.run_timed(T) for _ in range(3))
        rec = dict(prec=prec, nx=nx, ny=ny, engine="default", tb_depth=d["tb_depth"], vec=d["vec"], mlups=round(nx * ny * T / ms / 1e3, 1))
        out.append(rec); print(json.dumps(rec), flush=True)
        init(); lat.run(32); want = lat.download("A")
        for mode, blk in ((1, 128), (1, 256), (2, 128), (2, 64)):     # 1: cp.async both phases; 2: TMA bulk copies in phase 2
            for rpt in (4, 8, 16):
                lat.set_tuning(vec=1)
                lat.set_experimental(stage_smem=mode, rows_per_thread=rpt, block=blk)
                init(); lat.run(32)
                same = bool(np.array_equal(lat.download("A").view(np.uint8), want.view(np.uint8)))
                lat.run_timed(min(T, 400))
                ms = min(lat.run_timed(T) for _ in range(3))
                rec = dict(prec=prec, nx=nx, ny=ny, engine="staged", stage_smem=mode, rows_per_thread=rpt, block=blk, bit_identical=same,
                           mlups=round(nx * ny * T / ms / 1e3, 1))
                out.append(rec); print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "stage_probe.json"), "w"), indent=0)
