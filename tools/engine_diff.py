"""Where do two engines of the library disagree on a big lattice?  (debug tool: checksums per engine / tuning, then a
row-block-wise comparison of the downloads)   usage: engine_diff.py [dp|sp] [nx] [ny] [steps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else "dp"
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
ny = int(sys.argv[3]) if len(sys.argv) > 3 else 32768
T = int(sys.argv[4]) if len(sys.argv) > 4 else 32
scene = dict(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=min(nx, ny) // 10, perturb_seed=12345, perturb_amp=0.1)
configs = [("inplace default", True, {}), ("inplace tb_depth=0", True, dict(tb_depth=0)), ("inplace tb rows only", True, dict(tb_cols=1 << 20)),
           ("ab default", False, {}), ("ab vec=1", False, dict(vec=1)), ("ab vec=2", False, dict(vec=2))]
lats, sums = {}, {}
for name, inplace, tune in configs:
    lat = Lattice(nx, ny, prec, inplace=inplace)
    if tune:
        lat.set_tuning(**tune)
    lat.InitializeArrays(**scene)
    s0 = lat.checksum()
    lat.run(T)
    sums[name] = lat.checksum()
    print(f"{name:24s} init {s0:20d} after {T} steps {sums[name]:20d}  tuning {lat.get_tuning()}", flush=True)
    if name in ("inplace default", "ab default"):
        lats[name] = lat
    else:
        lat.close()
a, b = lats["inplace default"], lats["ab default"]
if sums["inplace default"] != sums["ab default"]:
    H = 256
    total = 0
    for r0 in range(0, nx, H):
        r1 = min(nx, r0 + H)
        x, y = a.download_window(r0, r1, 0, ny), b.download_window(r0, r1, 0, ny)
        bad = np.argwhere(x != y)
        if len(bad):
            total += len(bad)
            print(f"rows [{r0},{r1}): {len(bad)} differ; planes {sorted(set(bad[:, 0].tolist()))}; rows {r0 + bad[:, 1].min()}..{r0 + bad[:, 1].max()}; "
                  f"cols {bad[:, 2].min()}..{bad[:, 2].max()}; first {bad[0].tolist()} {x[tuple(bad[0])]!r} vs {y[tuple(bad[0])]!r}", flush=True)
        if total > 5_000_000:
            break
    print("total differing elements:", total)
else:
    print("engines agree")
