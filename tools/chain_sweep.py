"""Round 2: concurrent tile chains (lbm_tuning.tb_chains) inside ONE context: chains x tile budget per chain x hints, on the
weak-scaling unit and on the small temporally blocked lattices (CUDA-graph path).  Every point restarts from the same initial
state and must end with the checksum of the single-chain run.   `resident`: lattices that fit in L2 whole (no temporal blocking by default): one launch per step against 16-level macro-steps of 2-3
chains.   usage: chain_sweep.py [dp|sp|both] [large|small|resident|all]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

precs = ("dp", "sp") if len(sys.argv) < 2 or sys.argv[1] == "both" else (sys.argv[1],)
which = sys.argv[2] if len(sys.argv) > 2 else "all"
out = []
for prec in precs:
    R = 8 if prec == "dp" else 4
    lattices = []
    if which in ("large", "all"):
        lattices.append((8192, 32768, 64))
    if which in ("small", "all"):
        lattices += [(400, 4000, 1600), (800, 4000, 800)]
    if which in ("resident", "all"):
        lattices += [(400, 2000, 3200)] if prec == "dp" else [(400, 4000, 1600), (400, 2000, 3200)]
    for nx, ny, T in lattices:
        k = 16
        with Lattice(nx, ny, prec, inplace=True) as lat:
            sc = dict(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=min(nx, ny) // 10, perturb_seed=7, perturb_amp=0.05)
            base = lat.get_tuning()
            resident = 9 * R * nx * ny <= 60e6
            combos = [(1, 0, 0)]
            if resident and which in ("resident", "all"):
                combos = [(0, 0, 0), (1, 0, ny), (2, 0, ny), (3, 0, ny), (2, 60, ny), (2, 40, ny)]      # chains 0: tb off (the default there)
            for C in (() if resident else (2, 3, 4)):
                combos.append((C, 0, 0))
                if nx >= 8192:
                    w = 16384 // R
                    for mb in ((16, 20, 24, 30, 36, 44) if C == 2 else (16, 24)):
                        h = int(mb * 1e6 / (9 * R * (w + 8 * k))) - k
                        if h >= 2 * k:
                            combos.append((C, h, w))
                else:
                    for mb in (16, 24, 32):
                        h = int(mb * 1e6 / (9 * R * ((ny + 15) // 16 * 16))) - k
                        if h >= 2:
                            combos.append((C, h, 0))
            ref = None
            for C, h, w in combos:
                for hints in ((0, 3) if (C <= 2 and h == 0) else (0,)):
                    lat.InitializeArrays(**sc)
                    lat.set_tuning(**{**base, "tb_depth": k if C else 0, "tb_chains": max(C, 1), "tb_rows": h, "tb_cols": w, "hints": hints})
                    lat.run_timed(k)
                    lat.InitializeArrays(**sc)
                    ms = lat.run_timed(T)
                    cs = lat.checksum()
                    ms = min(ms, lat.run_timed(T))
                    if ref is None:
                        ref = cs
                    rec = dict(prec=prec, nx=nx, ny=ny, chains=C, H=h, W=w, hints=hints, mlups=round(nx * ny * T / ms / 1e3, 1),
                               same_as_single_chain=(cs == ref))
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "chain_sweep.json"), "w"), indent=0)
bad = [r for r in out if not r["same_as_single_chain"]]
print("ALL CHECKSUMS AGREE" if not bad else f"MISMATCH in {len(bad)} points")
sys.exit(1 if bad else 0)
