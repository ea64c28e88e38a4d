"""In-place (one buffer, AA kernels) against the A/B double buffer on one GPU: MLUPS per lattice, precision,
vector width, CTA size and launch engine.  Prints one JSON record per setting and the best per lattice;
writes gpurun_out/aa_sweep.json."""
import itertools
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

BYTES = {"dp": 144, "sp": 72}
LATTICES = [(400, 2000, 4000), (400, 4000, 4000), (800, 4000, 2000), (8192, 32768, 20)]
ENG = {"stream": dict(pdl=0, graph=0), "graph+pdl": dict(pdl=1, graph=10)}


def main():
    only = sys.argv[1] if len(sys.argv) > 1 else None
    out = []
    for prec in ("dp", "sp"):
        vecs = (1, 2, 4) if prec == "dp" else (1, 2, 4, 8)
        for nx, ny, T in LATTICES:
            if only == "small" and nx > 800:
                continue
            for inplace in (True, False):
                with Lattice(nx, ny, prec, inplace=inplace) as lat:
                    lat.InitializeArrays()
                    lat.run_timed(max(2, T // 10))
                    blocks = (128, 256) if inplace else (128,)
                    vv = vecs if inplace else ((1, 2) if prec == "dp" and nx <= 800 else (4,))
                    for vec, block, eng in itertools.product(vv, blocks, ENG):
                        if nx > 800 and eng != "stream":
                            continue
                        lat.set_tuning(vec=vec, neighbour=1, hints=0, block=block, ctas_per_sm=0, persistent=0, **ENG[eng])
                        lat.run_timed(max(2, T // 10))
                        ms = min(lat.run_timed(T) for _ in range(2))
                        mlups = nx * ny * T / (ms * 1e-3) / 1e6
                        rec = dict(prec=prec, nx=nx, ny=ny, inplace=int(inplace), vec=vec, block=block, engine=eng,
                                   us_per_step=round(ms * 1e3 / T, 2), mlups=round(mlups, 1), gbs=round(mlups * BYTES[prec] / 1e3, 1))
                        out.append(rec)
                        print(json.dumps(rec), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "aa_sweep.json"), "w"), indent=0)
    for prec in ("dp", "sp"):
        for nx, ny, _ in LATTICES:
            for inplace in (1, 0):
                best = sorted((r for r in out if r["prec"] == prec and r["nx"] == nx and r["ny"] == ny and r["inplace"] == inplace), key=lambda r: -r["mlups"])[:2]
                for r in best:
                    print("best", r)


if __name__ == "__main__":
    t0 = time.time()
    main()
    print(f"aa_sweep took {time.time() - t0:.1f} s")
