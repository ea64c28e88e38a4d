"""In-place kernels with capped grids (lbm_tuning.ctas_per_sm: CTAs walk several rows, the per-thread prologue -- parameter
loads, halo pointers, ~60 of ~260 instructions per DP site -- is paid once per thread instead of once per site)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

out = []
for prec, nx, ny, T in (("dp", 400, 2000, 4000), ("dp", 400, 4000, 2000), ("sp", 400, 4000, 2000), ("dp", 8192, 32768, 48)):
    with Lattice(nx, ny, prec, inplace=True) as lat:
        if nx <= 2000:
            lat.InitializeArrays()
        else:
            lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        d = lat.get_tuning()
        for cap in (0, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 16, 18, 24, 32):
            lat.set_tuning(ctas_per_sm=cap)
            lat.run_timed(min(T, 400))
            ms = min(lat.run_timed(T) for _ in range(3))
            rec = dict(prec=prec, nx=nx, ny=ny, tb_depth=d["tb_depth"], vec=d["vec"], graph=d["graph"], ctas_per_sm=cap, mlups=round(nx * ny * T / ms / 1e3, 1))
            out.append(rec); print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "cap_probe.json"), "w"), indent=0)
