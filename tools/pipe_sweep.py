"""Software-pipelined in-place kernels (lbm_tuning.rows_per_thread) on the L2-resident workloads."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

out = []
for prec, nx, ny, T in (("dp", 400, 4000, 4000), ("dp", 400, 2000, 4000), ("dp", 800, 4000, 2000), ("sp", 400, 4000, 4000), ("sp", 800, 4000, 2000)):
    with Lattice(nx, ny, prec, inplace=True) as lat:
        lat.InitializeArrays()
        base = lat.get_tuning()
        for vec in (1, 2) if prec == "dp" else (1, 2, 4):
            for rpt in (0, 2, 3, 4, 5, 6, 8, 12):
                if vec == 4 and rpt:
                    continue
                for block in (128, 256):
                    lat.set_tuning(vec=vec, rows_per_thread=rpt, block=block)
                    lat.run_timed(400)
                    ms = min(lat.run_timed(T) for _ in range(2))
                    rec = dict(prec=prec, nx=nx, ny=ny, vec=vec, rows_per_thread=rpt, block=block, tb_depth=base["tb_depth"],
                               us_per_step=round(ms * 1e3 / T, 2), mlups=round(nx * ny * T / ms / 1e3, 1))
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
    rs = [r for r in out if r["prec"] == prec and r["nx"] == nx and r["ny"] == ny]
    for r in sorted(rs, key=lambda r: -r["mlups"])[:4] + [r for r in rs if r["rows_per_thread"] == 0 and r["block"] == 128]:
        print("best", r)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "pipe_sweep.json"), "w"), indent=0)
