"""Round 2: does a second, independent chain of launches fill the drain every level of a temporal-blocking tile pays?
One GPU, the lattice cut into n in-process slabs (lbm_group_run: each slab runs its own tile chain on its own streams, seams
ordered by events), each chain with 1/n of the L2 tile budget.  n = 1 is the shipped single-chain schedule.
usage: two_chain_probe.py [dp|sp] [nx ny]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import make_ring, group_run  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "dp"
nx, ny = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (8192, 32768)
R = 8 if prec == "dp" else 4
T, k = 64, 16
out = []
for n in (1, 2, 3, 4):
    slabs = make_ring(nx, ny, n, prec, inplace=True)
    try:
        for s in slabs:
            s.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        combos = [(0, 0)]
        for mb in (72, 60, 48):
            for w in (16384 // R, 32768 // R):
                h = int(mb / n * 1e6 / (9 * R * (w + 8 * k))) - k
                if h >= 2 * k:
                    combos.append((h, w))
        for h, w in combos:
            for s in slabs:
                s.set_tuning(tb_depth=k, tb_rows=h, tb_cols=w)
            best = 1e9
            for rep in range(3):
                for s in slabs:
                    s.sync()
                t0 = time.perf_counter()
                group_run(slabs, T)
                for s in slabs:
                    s.sync()
                if rep:
                    best = min(best, time.perf_counter() - t0)
            rec = dict(prec=prec, nx=nx, ny=ny, chains=n, H=h, W=w, tile_mb_per_chain=round((h + k) * (w + 8 * k) * 9 * R / 1e6, 1) if w else None,
                       mlups=round(nx * ny * T / best / 1e6, 1))
            out.append(rec)
            print(json.dumps(rec), flush=True)
    finally:
        for s in slabs:
            s.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"two_chain_{prec}.json"), "w"), indent=0)
