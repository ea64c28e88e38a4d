"""What the L2 of this B200 delivers to a streaming kernel whose working set is resident: the roof of the temporally
blocked in-place engine (its tiles live in L2 for k timesteps, HBM is not what binds it -- VERDICT r1, weak 7).

Two probes, CUDA events, best of N after warm-up:
  copy   b.copy_(a) over a working set of W MB (a and b together), repeated back to back; bytes = read + write.
         W from 8 MB to 512 MB: the plateau below ~60 MB is the L2 roof, the tail is HBM (cf. MEASURED_PEAKS.json hbm_gbs).
  lbm    the library's own local in-place phase (phase 2: 9 aligned loads + 9 aligned stores per site, no neighbours)
         on lattices of the same working-set sizes -- the same access shape as the product kernel.
Output: one JSON line per point and gpurun_out/l2_peak.json; profiles/ keeps the copy bench.py reads (`l2_gbs`).
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def copy_probe(mb, reps=200):
    n = int(mb * 1e6 / 2 / 8)
    a = torch.ones(n, dtype=torch.float64, device="cuda")
    b = torch.empty_like(a)
    for _ in range(20):
        b.copy_(a)
    best = 1e9
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            b.copy_(a)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / reps)
    return 2 * n * 8 / (best * 1e-3) / 1e9


def lbm_probe(prec, nx, ny, reps=400):
    from latticeboltzmann_b200 import Lattice
    with Lattice(nx, ny, prec, inplace=True) as lat:
        lat.InitializeArrays()
        lat.set_tuning(pdl=1, graph=0, tb_depth=0)
        out = {}
        for phase in (1, 2):
            for _ in range(50):
                lat._call("lbm_step_rows", phase - 1, 0, nx, 1, 0)
            best = 1e9
            for _ in range(3):
                lat.sync()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                # the library launches on its own stream: bracket with a device-wide sync and time on the host side of it
                import time
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                for _ in range(reps):
                    lat._call("lbm_step_rows", phase - 1, 0, nx, 1, 0)
                lat.sync()
                best = min(best, (time.perf_counter() - t0) / reps)
            b = 144 if prec == "dp" else 72
            out[f"phase{phase}_gbs"] = round(nx * ny * b / best / 1e9, 1)
            out[f"phase{phase}_glups"] = round(nx * ny / best / 1e9, 2)
        return out


def main():
    res = {"copy": [], "lbm": []}
    for mb in (8, 16, 24, 32, 40, 48, 56, 64, 80, 96, 128, 192, 256, 512):
        rec = {"working_set_mb": mb, "gbs": round(copy_probe(mb), 1)}
        res["copy"].append(rec)
        print(json.dumps(rec), flush=True)
    for prec, nx, ny in (("dp", 100, 2000), ("dp", 200, 2000), ("dp", 300, 2000), ("dp", 400, 2000), ("dp", 400, 4000),
                         ("sp", 200, 4000), ("sp", 400, 4000), ("sp", 800, 4000)):
        rec = {"prec": prec, "nx": nx, "ny": ny, "working_set_mb": round(9 * nx * ny * (8 if prec == "dp" else 4) / 1e6, 1)}
        rec.update(lbm_probe(prec, nx, ny))
        res["lbm"].append(rec)
        print(json.dumps(rec), flush=True)
    res["l2_gbs"] = max(r["gbs"] for r in res["copy"] if r["working_set_mb"] <= 48)
    res["how"] = "torch b.copy_(a) float64, working set (a+b) <= 48 MB, 200 back-to-back copies, best of 5, CUDA events; read+write bytes"
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(res, open(os.path.join(ROOT, "gpurun_out", "l2_peak.json"), "w"), indent=1)
    print(json.dumps({"l2_gbs": res["l2_gbs"]}))


if __name__ == "__main__":
    main()
