"""Fused pair kernel (lbm_tuning.fuse_rows: phase 2 + next phase 1 in one pass, intermediates through shared memory)
against the library defaults: same bits first, then MLUPS over tile height and CTA width.
usage: pair_probe.py [small|large|both]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "small"
out = []


def probe(prec, nx, ny, T, check_steps, tiles):
    with Lattice(nx, ny, prec, inplace=True) as lat:
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
        init(); lat.run(check_steps); want = lat.download("A")
        for blk in (128, 64):
            for tr in tiles:
                lat.set_tuning(fuse_rows=tr, block=blk, pdl=1, graph=0, tb_depth=0)
                init(); lat.run(check_steps)
                got = lat.download("A")
                same = bool(np.array_equal(got.view(np.uint8), want.view(np.uint8)))
                nbad = int((got != want).sum()) if not same else 0
                first = [int(v) for v in np.argwhere(got != want)[0]] if nbad else None
                l0 = lat.launches
                lat.run_timed(min(T, 400))
                ms = min(lat.run_timed(T) for _ in range(3))
                rec = dict(prec=prec, nx=nx, ny=ny, engine="pairs", fuse_rows=tr, block=blk, bit_identical=same, differ=nbad, first=first,
                           mlups=round(nx * ny * T / ms / 1e3, 1))
                out.append(rec); print(json.dumps(rec), flush=True)


if what in ("small", "both"):
    probe("dp", 400, 4000, 2000, 64, (4, 8, 12, 16, 24, 32))
    probe("sp", 400, 4000, 2000, 64, (4, 8, 16, 32))
    probe("dp", 400, 2000, 2000, 64, (8, 16))
    probe("dp", 800, 4000, 1000, 32, (8, 16, 32))
if what == "quick":
    probe("dp", 400, 4000, 2000, 64, (8, 12, 16))
    probe("sp", 400, 4000, 2000, 64, (8,))
    probe("dp", 8192, 32768, 48, 8, (16, 32, 64))
if what in ("large", "both"):
    probe("dp", 8192, 32768, 48, 8, (16, 32, 64, 128))
    probe("sp", 8192, 32768, 48, 8, (32, 64))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"pair_probe_{what}.json"), "w"), indent=0)
