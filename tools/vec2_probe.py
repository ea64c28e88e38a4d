"""DP vec 2 under register caps (experiment libraries via LBM_B200_LIBDIR): temporally blocked 400x4000 and the
L2-resident 400x2000 lattice, phase mixes (vec / vec_phase1) and the shuffle variant."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice  # noqa: E402

tag = os.environ.get("LBM_B200_LIBDIR", "default").rstrip("/").split("/")[-1]
for nx, ny, T in ((400, 4000, 2000), (400, 2000, 4000), (8192, 32768, 48)):
    with Lattice(nx, ny, "dp", inplace=True) as lat:
        if nx <= 2000:
            lat.InitializeArrays()
        else:
            lat.InitializeArrays(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=nx // 10)
        for vec, v1, nb in ((1, 0, 1), (2, 0, 1), (2, 1, 1), (1, 2, 1), (2, 0, 0), (1, 2, 0)):
            lat.set_tuning(vec=vec, vec_phase1=v1, neighbour=nb)
            lat.run_timed(min(T, 400))
            ms = min(lat.run_timed(T) for _ in range(3))
            print(json.dumps(dict(lib=tag, nx=nx, ny=ny, vec=vec, vec_phase1=v1, neighbour=nb, mlups=round(nx * ny * T / ms / 1e3, 1))), flush=True)
