"""Dataflow kernel (lbm_tuning.dataflow: one launch per macro-step of the wavefront schedule) against the launch-per-
(band, level) engine: MLUPS over items size, depth, band height, load path.  First checks, per lattice, that the two
engines leave the same bits.   usage: df_sweep.py [small|large|both]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

what = sys.argv[1] if len(sys.argv) > 1 else "small"
out = []


def timed(lat, T, reps=3):
    lat.run_timed(T if T < 400 else 400)
    return min(lat.run_timed(T) for _ in range(reps))


def sweep(prec, nx, ny, T, base, combos, check_steps):
    R = 8 if prec == "dp" else 4
    with Lattice(nx, ny, prec, inplace=True) as lat:
        def init():
            if nx <= 1000:
                lat.InitializeArrays()
            else:
                lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        init()
        dflt = lat.get_tuning()
        ms = timed(lat, T)
        rec = dict(prec=prec, nx=nx, ny=ny, engine="default", tuning={k: dflt[k] for k in ("vec", "tb_depth", "tb_rows", "tb_cols", "graph")},
                   mlups=round(nx * ny * T / ms / 1e3, 1))
        out.append(rec); print(json.dumps(rec), flush=True)
        want = None
        if check_steps:
            init(); lat.run(check_steps); want = lat.download("A")
        for kw in combos:
            tune = dict(base, **kw)
            lat.set_tuning(**tune)
            ok = None
            if check_steps and kw.get("check"):
                init(); lat.run(check_steps)
                ok = bool(np.array_equal(lat.download("A").view(np.uint8), want.view(np.uint8)))
            tune.pop("check", None)
            l0 = lat.launches
            ms = timed(lat, T)
            nl = lat.launches - l0
            rec = dict(prec=prec, nx=nx, ny=ny, engine="dataflow" if tune.get("dataflow") else "launches",
                       **{k: tune[k] for k in ("vec", "tb_depth", "tb_rows", "dataflow", "hints", "block", "ctas_per_sm")},
                       launches=nl, bit_identical=ok, mlups=round(nx * ny * T / ms / 1e3, 1))
            out.append(rec); print(json.dumps(rec), flush=True)
    rs = [r for r in out if r["prec"] == prec and r["nx"] == nx and r["ny"] == ny]
    for r in sorted(rs, key=lambda r: -r["mlups"])[:5]:
        print("best", r, flush=True)


class Tune(dict):
    pass


def combos_small(prec):
    v = 1 if prec == "dp" else 4
    k0 = 16 if prec == "dp" else 8
    cs = [dict(vec=v, tb_depth=k0, dataflow=0, graph=1, check=1),        # launch-based reference point with these bands
          dict(vec=v, tb_depth=k0, dataflow=1, check=1), dict(vec=v, tb_depth=k0, dataflow=1, hints=1, check=1)]
    for u in (2, 4, 8):
        cs.append(dict(vec=v, tb_depth=k0, dataflow=u))
    for k in (k0, 2 * k0, 4 * k0):
        for h in (0, 48, 96):
            for u in (1, 2):
                cs.append(dict(vec=v, tb_depth=k, tb_rows=h, dataflow=u))
    cs += [dict(vec=v, tb_depth=k0, dataflow=1, block=64), dict(vec=v, tb_depth=k0, dataflow=2, block=64),
           dict(vec=v, tb_depth=k0, dataflow=1, ctas_per_sm=6), dict(vec=v, tb_depth=k0, dataflow=1, ctas_per_sm=4),
           dict(vec=2, tb_depth=k0, dataflow=1, check=1), dict(vec=2, tb_depth=k0, dataflow=2)]
    if prec == "sp":
        cs += [dict(vec=4, tb_depth=k0, dataflow=1, block=256), dict(vec=4, tb_depth=16, tb_rows=0, dataflow=1, hints=1)]
    return cs


BASE = dict(neighbour=1, hints=0, block=128, ctas_per_sm=0, persistent=0, pdl=1, graph=0, tb_rows=0, tb_cols=1 << 20,
            vec_phase1=0, rows_per_thread=0)
if what in ("small", "both"):
    for prec in ("dp", "sp"):
        sweep(prec, 400, 4000, 2000, BASE, combos_small(prec), 64)
    sweep("dp", 800, 4000, 1000, BASE, [dict(vec=1, tb_depth=16, dataflow=1, check=1), dict(vec=1, tb_depth=16, dataflow=2),
                                         dict(vec=1, tb_depth=32, dataflow=1), dict(vec=1, tb_depth=32, tb_rows=48, dataflow=1)], 32)
if what == "v2":                                                   # second pass: item size x CTA size around the first pass' best
    for prec in ("dp", "sp"):
        v, k0 = (1, 16) if prec == "dp" else (4, 8)
        cs = [dict(vec=v, tb_depth=k0, dataflow=0, graph=1, check=1)]
        for blk in (128, 64, 32):
            for u in (1, 2, 3, 4, 6, 8, 16):
                if blk * u <= 1024:
                    cs.append(dict(vec=v, tb_depth=k0, dataflow=u, block=blk, check=int(u in (1, 3))))
        cs += [dict(vec=2, tb_depth=k0, dataflow=u) for u in (1, 2, 3)]
        cs += [dict(vec=v, tb_depth=k0, dataflow=u, hints=1) for u in (2, 4)]
        sweep(prec, 400, 4000, 2000, BASE, cs, 64)
if what in ("large", "both"):
    for prec in ("dp", "sp"):
        v = 1 if prec == "dp" else 4
        cs = [dict(vec=v, tb_depth=8, dataflow=1, check=1)]
        for k in (8, 12, 16):
            for h in (0, 8, 16):
                cs.append(dict(vec=v, tb_depth=k, tb_rows=h, dataflow=1))
        cs += [dict(vec=v, tb_depth=8, dataflow=2), dict(vec=v, tb_depth=8, dataflow=4), dict(vec=v, tb_depth=8, dataflow=1, hints=1)]
        sweep(prec, 8192, 32768, 48, BASE, cs, 16)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"df_sweep_{what}.json"), "w"), indent=0)
