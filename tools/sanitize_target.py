"""Small in-place / temporally blocked / ring runs, every kernel variant that computes addresses from wrapped rows or
columns, each compared bit for bit with the A/B result (written for compute-sanitizer memcheck; that tool is closed on
this pool, so the cross-check of variants is what remains)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from latticeboltzmann_b200 import Lattice, group_run, make_ring  # noqa: E402
from tests import cases  # noqa: E402

for prec in ("dp", "sp"):
    for nx, ny in ((48, 264), (17, 43)):
        c = dict(prec=prec, isa="avx", nx=nx, ny=ny, scene="random", perturb=5, steps=12)
        fA, walls = cases.make_inputs(c)
        want = None
        for kw, tune in (({}, {}), (dict(inplace=True), dict(tb_depth=0)), (dict(inplace=True), dict(tb_depth=4, tb_rows=9, tb_cols=64, vec=2)),
                         (dict(inplace=True), dict(tb_depth=6, tb_rows=5, tb_cols=0, vec=1, rows_per_thread=3)),
                         (dict(inplace=True), dict(tb_depth=4, tb_rows=7, tb_cols=40, neighbour=0, vec_phase1=2))):
            with Lattice(nx, ny, prec, **kw) as lat:
                lat.set_tuning(**tune)
                lat.upload(fA, walls)
                lat.run(c["steps"])
                got = lat.download("A")
            if want is None:
                want = got
            assert got.tobytes() == want.tobytes(), (prec, nx, ny, kw, tune)
        for tune in (dict(tb_depth=0), dict(tb_depth=4, tb_rows=5, tb_cols=64), dict(tb_depth=2, tb_rows=0, tb_cols=0)):
            ring = make_ring(nx, ny, 3, prec, inplace=True)
            for s in ring:
                s.set_tuning(**tune)
                s.upload(fA, walls)
            group_run(ring, c["steps"])
            got = np.zeros_like(want)
            for s in ring:
                s.download("A", out=got)
                s.close()
            assert got.tobytes() == want.tobytes(), (prec, nx, ny, "ring", tune)
print("sanitize target: all variants agree")
