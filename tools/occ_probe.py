"""DP in-place kernels at a given register cap (experiment libraries under latticeboltzmann_b200/exp/, selected
with LBM_B200_LIBDIR): L2-resident 400x2000 and temporally blocked 400x4000 / 800x4000."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

tag = os.environ.get("LBM_B200_LIBDIR", "default").rstrip("/").split("/")[-1]
for nx, ny, T in ((400, 2000, 4000), (400, 4000, 4000), (800, 4000, 2000)):
    with Lattice(nx, ny, "dp", inplace=True) as lat:
        lat.InitializeArrays()
        for vec in (1, 2):
            lat.set_tuning(vec=vec, block=128)
            lat.run_timed(400)
            ms = min(lat.run_timed(T) for _ in range(3))
            print(json.dumps(dict(lib=tag, nx=nx, ny=ny, vec=vec, us_per_step=round(ms * 1e3 / T, 2), mlups=round(nx * ny * T / ms / 1e3, 1))), flush=True)
