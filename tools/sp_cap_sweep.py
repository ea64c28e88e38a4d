"""SP in-place engine after the 64-register cap: which vector width / depth the temporally blocked lattices want now."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

out = []
for nx, ny, T in ((800, 4000, 2000), (1600, 4000, 1000), (8192, 32768, 48)):
    with Lattice(nx, ny, "sp", inplace=True) as lat:
        if nx <= 2000:
            lat.InitializeArrays()
        else:
            lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        d = lat.get_tuning()
        combos = [dict()] + [dict(vec=v, tb_depth=k, block=b) for v in (2, 4) for k in (8, 16) for b in (128, 256)]
        if nx > 2000:
            combos += [dict(vec=4, tb_depth=8, tb_cols=w) for w in (4096, 16384)] + [dict(vec=8, tb_depth=8)]
        for kw in combos:
            lat.set_tuning(**dict(dict(vec=d["vec"], tb_depth=d["tb_depth"], block=128, tb_cols=0, tb_rows=0), **kw))
            t = lat.get_tuning()
            lat.run_timed(T // 4 & ~1)
            ms = min(lat.run_timed(T) for _ in range(3))
            rec = dict(nx=nx, ny=ny, default=not kw, vec=t["vec"], k=t["tb_depth"], block=t["block"], tb_cols=t["tb_cols"], graph=t["graph"],
                       mlups=round(nx * ny * T / ms / 1e3, 1))
            out.append(rec); print(json.dumps(rec), flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "sp_cap_sweep.json"), "w"), indent=0)
