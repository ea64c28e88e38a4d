"""Temporal blocking in L2 (in-place engine, wavefront schedule): MLUPS over depth k and band height H.
usage: tb_sweep.py [small|large|all]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

BYTES = {"dp": 144, "sp": 72}
which = sys.argv[1] if len(sys.argv) > 1 else "all"
LATTICES = []
if which in ("small", "all"):
    LATTICES += [(400, 4000, 4000), (800, 4000, 2000), (400, 2000, 4000)]
if which in ("large", "all"):
    LATTICES += [(8192, 32768, 40)]


def main():
    out = []
    for prec in ("dp", "sp"):
        for nx, ny, T in LATTICES:
            row_mb = 9 * ny * (8 if prec == "dp" else 4) / 1e6
            with Lattice(nx, ny, prec, inplace=True) as lat:
                lat.InitializeArrays()
                vecs = ((1, 2) if nx <= 800 else (4,)) if prec == "dp" else (4,)
                combos = [(0, 0)]
                for k in (4, 8, 12, 16, 24, 32):
                    for mb in (30, 38, 45, 52, 60, 70):
                        h = int(mb / row_mb) - k
                        if h >= 2 and 2 * k + 2 <= nx and h < nx:
                            combos.append((k, h))
                for vec in vecs:
                    for k, h in combos:
                        for g in ((1,) if nx <= 800 else (0, 1)):
                            lat.set_tuning(vec=vec, hints=0, pdl=1 if nx <= 800 else g,
                                           graph=10 * g, tb_depth=k, tb_rows=h)
                            lat.run_timed(max(2 * k, T // 10) & ~1)
                            l0 = lat.launches
                            ms = min(lat.run_timed(T) for _ in range(2))
                            nl = (lat.launches - l0) // 2
                            mlups = nx * ny * T / (ms * 1e-3) / 1e6
                            rec = dict(prec=prec, nx=nx, ny=ny, vec=vec, k=k, H=h, band_mb=round((h + k) * row_mb, 1), graph=g,
                                       launches_per_step=round(nl / T, 2), us_per_step=round(ms * 1e3 / T, 2), mlups=round(mlups, 1),
                                       gbs=round(mlups * BYTES[prec] / 1e3, 1))
                            out.append(rec)
                            print(json.dumps(rec), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"tb_sweep_{which}.json"), "w"), indent=0)
    for prec in ("dp", "sp"):
        for nx, ny, _ in LATTICES:
            rs = [r for r in out if r["prec"] == prec and r["nx"] == nx and r["ny"] == ny]
            base = [r for r in rs if r["k"] == 0]
            for r in sorted(base, key=lambda r: -r["mlups"])[:1] + sorted((r for r in rs if r["k"]), key=lambda r: -r["mlups"])[:4]:
                print("best", r)


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"tb_sweep took {time.time() - t0:.1f} s")
