"""Round 2: how many CTAs should be resident?  A launch of the in-place step loses about one CTA lifetime at its drain (the
next level's CTAs are resident but wait at griddepcontrol.wait), and a CTA's lifetime under load is queueing in L2: bytes in
flight / bandwidth.  The experiment build's smem_pad caps the CTAs per SM without touching the kernel.
usage: residency_probe.py [dp|sp|both]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

precs = ("dp", "sp") if len(sys.argv) < 2 or sys.argv[1] == "both" else (sys.argv[1],)
out = []
for prec in precs:
    for nx, ny in ((400, 2000), (400, 4000)) if prec == "dp" else ((400, 4000), (800, 4000)):
        with Lattice(nx, ny, prec, inplace=True, experimental=True) as lat:
            lat.InitializeArrays()
            base = lat.get_tuning()
            for block in (64, 128, 256):
                for pad in (0, 24 << 10, 28 << 10, 32 << 10, 37 << 10, 45 << 10, 48 << 10):
                    per_sm = min(2048 // block, 32, 228 * 1024 // (pad + 1024)) if pad else None
                    lat.set_tuning(**base)
                    lat.set_experimental(stage_smem=0, rows_per_thread=0, block=block, smem_pad=pad)
                    lat.run_timed(200)
                    ms = min(lat.run_timed(2000) for _ in range(2))
                    rec = dict(prec=prec, nx=nx, ny=ny, block=block, smem_pad=pad, ctas_per_sm_cap=per_sm, mlups=round(nx * ny * 2000 / ms / 1e3, 1))
                    out.append(rec)
                    print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "residency_probe.json"), "w"), indent=0)
