"""2-D temporal blocking (row bands x column strips) on the wide lattices: MLUPS over depth, strip width, tile budget.
usage: tb2d_sweep.py [dp|sp|both] [nx ny]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

precs = ("dp", "sp") if len(sys.argv) < 2 or sys.argv[1] == "both" else (sys.argv[1],)
nx, ny = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (8192, 32768)
out = []
for prec in precs:
    R = 8 if prec == "dp" else 4
    with Lattice(nx, ny, prec, inplace=True) as lat:
        lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        combos = [(0, 0, 0, 4, 0)]
        fine = len(sys.argv) > 4 and sys.argv[4] == "fine"
        ks = ((16, 24, 32) if prec == "dp" else (8, 16)) if fine else (8, 16)
        ws = ((1024, 2048, 4096) if prec == "dp" else (2048, 4096, 8192)) if fine else ((2048, 4096, 8192) if prec == "dp" else (4096, 8192, 16384))
        for k in ks:
            for w in ws:
                for mb in ((60, 75, 90, 110) if fine else (30, 45, 60)):
                    h = int(mb * 1e6 / (9 * R * (w + 8 * k))) - k
                    if h < 2 * k:
                        continue
                    for vec in (((1,) if prec == "dp" else (2, 4)) if fine else ((1, 2) if prec == "dp" else (2, 4))):
                        for g in ((0,) if fine else (0, 1)):
                            combos.append((k, h, w, vec, g))
        for k, h, w, vec, g in combos:
            T = 48
            lat.set_tuning(vec=vec, hints=0, pdl=1 if k else 0, graph=g,
                           tb_depth=k, tb_rows=h, tb_cols=w)
            lat.run_timed(max(k, 2))
            l0 = lat.launches
            ms = lat.run_timed(T)
            nl = lat.launches - l0
            rec = dict(prec=prec, nx=nx, ny=ny, k=k, H=h, W=w, vec=vec, graph=g, tile_mb=round((h + k) * (w + 8 * k) * 9 * R / 1e6, 1),
                       launches_per_step=round(nl / T, 1), us_per_launch=round(ms * 1e3 / max(nl, 1), 2), ms_per_step=round(ms / T, 3),
                       mlups=round(nx * ny * T / ms / 1e3, 1))
            out.append(rec)
            print(json.dumps(rec), flush=True)
    for r in sorted((r for r in out if r["prec"] == prec), key=lambda r: -r["mlups"])[:6]:
        print("best", r)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "tb2d_sweep.json"), "w"), indent=0)
