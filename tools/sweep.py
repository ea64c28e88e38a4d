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
                MODES = {"stream": dict(persistent=0, pdl=0, graph=0), "pdl": dict(persistent=0, pdl=1, graph=0),
                         "graph": dict(persistent=0, pdl=0, graph=10), "graph+pdl": dict(persistent=0, pdl=1, graph=10),
                         "persistent": dict(persistent=1, pdl=0, graph=0)}
                combos = list(itertools.product(vecs, (0, 1), (0, 1), (128, 256), (0,), ("stream",)))
                combos += list(itertools.product(vecs, (0, 1), (0,), (64, 128, 256), (0,), ("pdl", "graph", "graph+pdl")))
                combos += list(itertools.product(vecs, (0, 1), (0,), (128, 256, 512), (0, 2, 4), ("persistent",)))
                for vec, nb, hints, block, ctas, mode in combos:
                    if quick and block != 128:
                        continue
                    if vec == 1 and nb == 0:
                        continue                     # vec 1 has no shuffle variant (identical code to nb 1)
                    lat.set_tuning(vec=vec, neighbour=nb, hints=hints, block=block, ctas_per_sm=ctas, **MODES[mode])
                    lat.run_timed(max(2, T // 10))
                    ms = min(lat.run_timed(T) for _ in range(2))
                    mlups = nx * ny * T / (ms * 1e-3) / 1e6
                    rec = dict(prec=prec, nx=nx, ny=ny, T=T, vec=vec, neighbour=nb, hints=hints, block=block, ctas_per_sm=ctas, mode=mode,
                               ms=round(ms, 3), mlups=round(mlups, 1), gbs=round(mlups * BYTES[prec] / 1e3, 1),
                               frac=round(mlups * BYTES[prec] / 1e3 / peak, 4))
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sweep.json"), "w"), indent=0)
    for prec in ("dp", "sp"):
        for nx, ny, _ in lattices:
            for mode in ("stream", "pdl", "graph", "graph+pdl", "persistent"):
              best = sorted((r for r in out if r["prec"] == prec and r["nx"] == nx and r["mode"] == mode), key=lambda r: -r["mlups"])[:3]
              print(f"== best {prec} {nx}x{ny} {mode}")
              for r in best:
                print("  ", {k: r[k] for k in ("mode", "vec", "neighbour", "hints", "block", "ctas_per_sm", "mlups", "gbs", "frac")})


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"sweep took {time.time() - t0:.1f} s")
