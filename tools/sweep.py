"""Kernel-variant sweep on one GPU: MLUPS and fraction of the measured HBM peak for every
(vec, neighbour, hints, block, ctas_per_sm) on the small (400x4000) and large (8192x32768) lattices.
Writes gpurun_out/sweep.json; the best variants become the library defaults."""
import itertools
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

BYTES = {"dp": 144, "sp": 72}


def main():
    quick = "--quick" in sys.argv
    peak = 6523.7
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    out = []
    lattices = [(400, 4000, 2000), (8192, 32768, 20)]
    for prec in ("dp", "sp"):
        vecs = (1, 2, 4) if prec == "dp" else (1, 2, 4, 8)
        for nx, ny, T in lattices:
            with Lattice(nx, ny, prec) as lat:
                lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
                lat.run_timed(T)
                combos = list(itertools.product(vecs, (0, 1), (0, 1), (128, 256), (0,), (0,)))
                combos += list(itertools.product(vecs, (0, 1), (0,), (128, 256, 512), (0, 1, 2, 3, 4, 6, 8), (1,)))
                for vec, nb, hints, block, ctas, pers in combos:
                    if quick and (block != 256 or ctas not in (0,)):
                        continue
                    if pers and vec == 1 and nb == 0:
                        continue                     # vec 1 has no shuffle variant
                    lat.set_tuning(vec=vec, neighbour=nb, hints=hints, block=block, ctas_per_sm=ctas, persistent=pers)
                    lat.run_timed(max(2, T // 10))
                    ms = min(lat.run_timed(T) for _ in range(2))
                    mlups = nx * ny * T / (ms * 1e-3) / 1e6
                    rec = dict(prec=prec, nx=nx, ny=ny, T=T, vec=vec, neighbour=nb, hints=hints, block=block, ctas_per_sm=ctas, persistent=pers,
                               ms=round(ms, 3), mlups=round(mlups, 1), gbs=round(mlups * BYTES[prec] / 1e3, 1),
                               frac=round(mlups * BYTES[prec] / 1e3 / peak, 4))
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sweep.json"), "w"), indent=0)
    for prec in ("dp", "sp"):
        for nx, ny, _ in lattices:
            for pers in (0, 1):
              best = sorted((r for r in out if r["prec"] == prec and r["nx"] == nx and r["persistent"] == pers), key=lambda r: -r["mlups"])[:4]
              print(f"== best {prec} {nx}x{ny} persistent={pers}")
              for r in best:
                print("  ", {k: r[k] for k in ("persistent", "vec", "neighbour", "hints", "block", "ctas_per_sm", "mlups", "gbs", "frac")})


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"sweep took {time.time() - t0:.1f} s")
