"""Round 2: 2-D temporal-blocking tiles on the weak-scaling unit -- strip width x tile budget x depth x evict-first stores on
the last level (lbm_tuning.hints).  usage: tile_sweep.py [dp|sp|both] [quick|hints]   (hints: the default schedule with lbm_tuning.hints 0..3 only)"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

precs = ("dp", "sp") if len(sys.argv) < 2 or sys.argv[1] == "both" else (sys.argv[1],)
quick = len(sys.argv) > 2 and sys.argv[2] == "quick"
hints_only = len(sys.argv) > 2 and sys.argv[2] == "hints"
nx, ny, T = 8192, 32768, 64
out = []
for prec in precs:
    R = 8 if prec == "dp" else 4
    with Lattice(nx, ny, prec, inplace=True) as lat:
        lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        d = lat.get_tuning()
        combos = [(d["tb_depth"], 0, 0, h) for h in ((0, 1, 2, 3, 0, 3) if hints_only else (0,))]
        for k in (() if hints_only else (16,) if quick else (16, 32)):
            for w in ((16384 // R, 32768 // R) if quick else (8192 // R, 16384 // R, 32768 // R, 65536 // R)):
                for mb in ((40, 60) if quick else (24, 32, 40, 50, 60, 75)):
                    h = int(mb * 1e6 / (9 * R * (w + 8 * k))) - k
                    if h < 2 * k:
                        continue
                    for hints in (0, 1):
                        combos.append((k, h, w, hints))
        for k, h, w, hints in combos:
            lat.set_tuning(hints=hints, tb_depth=k, tb_rows=h, tb_cols=w)
            lat.run_timed(k)
            l0 = lat.launches
            ms = min(lat.run_timed(T) for _ in range(2))
            nl = (lat.launches - l0) // 2
            rec = dict(prec=prec, k=k, H=h, W=w, hints=hints, tile_mb=round((h + k) * (w + 8 * k) * 9 * R / 1e6, 1) if w else None,
                       launches_per_step=round(nl / T, 1), mlups=round(nx * ny * T / ms / 1e3, 1))
            out.append(rec)
            print(json.dumps(rec), flush=True)
    for r in sorted((r for r in out if r["prec"] == prec), key=lambda r: -r["mlups"])[:8]:
        print("best", r)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "tile_sweep.json"), "w"), indent=0)
