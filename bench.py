#!/usr/bin/env python
"""bench.py -- MLUPS of the D2Q9 BGK timestep path on B200, against the roof that binds it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

One bench "step" = one job of T lattice timesteps.  The headline workload is the SAME at every N, so the driver's
scaling efficiency compares like with like:

  headline  BASELINE configs[4], weak scaling: 8192x32768 sites PER GPU (N = 1: that lattice on one B200; N > 1: row slabs
            over the ranks, halo rows stored peer-to-peer over NVLink), DP, T = 1,000 timesteps per step (the reference's
            PRINTSTATSEVERY, LB.c:52: one observation interval), SP beside it under "sp".
  sub-records (named, each with roofline / e2e / parity where they apply):
            N = 1 : "c2_400x4000" (configs[1], DP and SP, 10,000 timesteps), "c3_800x4000_cylinder" (configs[2], DP),
                    "ab_double_buffer" (the A/B engine on the weak-scaling unit: the HBM-streaming regime),
                    "strong_16384x65536" (configs[3] on one GPU, in place: 77 GB)
            N >= 2: "strong_16384x65536" (configs[3]: that lattice slab-partitioned over the N ranks)

value     lattice-site updates of all ranks / device time (CUDA events, max over ranks), MLUPS, lattice resident in HBM
e2e       the same metric through the C ABI with pinned HOST buffers: lbm_upload[_slab] -> lbm_run -> lbm_download[_slab]
          inside the timed region, every step
roofline  the dominant kernel's algorithmic bytes (2*9*sizeof(real) per site update) per launch / its average
          duration over the timed region.  The default engine (in place + temporal blocking) keeps its tiles in L2 for k
          timesteps, so what binds it is L2 bandwidth: bound = "l2", peak = the measured L2 roof (profiles/l2_peak.json,
          tools/l2_peak_probe.py), frac <= 1; the DRAM side (measured bytes from the committed ncu capture against
          MEASURED_PEAKS.json hbm_gbs) is reported beside it under "hbm".  The A/B engine streams the lattice through HBM
          every step: bound = "hbm".
parity    config.parity: before anything is timed at N >= 2 the ring-parity pass (latticeboltzmann_b200/ringcheck.py: the
          slab ring over the N GPUs == the single-GPU lattice, bit for bit, 17 cases) and a cross-engine checksum of the timed
          lattice (in place + temporal blocking vs A/B: two independent engines); at N = 1 the golden digests of the
          compiled reference for configs[1], configs[2] (through the e2e path) and the same cross-engine checksum.
cpu_baseline / --impl reference = the reference's own AVX+OpenMP build (oracle/_ref, Makefile:2 flags), timed by its own
          "Runtime:" line on this box's host cores.
"""
import argparse
import hashlib
import json
import os
import re
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BYTES_PER_UPDATE = {"dp": 144, "sp": 72}      # 9 reads + 9 writes of real_t, SURVEY.md 8(d)
DTYPE_NAME = {"dp": "f64", "sp": "f32"}
C1 = dict(nx=400, ny=2000, timesteps=10000)                # BASELINE.json configs[0] (the reference's own CPU case)
C2 = dict(nx=400, ny=4000, timesteps=10000)                # configs[1]
C3 = dict(nx=800, ny=4000, timesteps=10000)                # configs[2]: channel + cylinder
WEAK = dict(nx_per_gpu=8192, ny=32768, timesteps=1000)     # configs[4]
STRONG = dict(nx=16384, ny=65536, timesteps=100)           # configs[3]
CHECK_STEPS = 32                                           # cross-engine checksum: two macro-steps of k = 16
T_START = time.time()


def elapsed():
    return time.time() - T_START


def read_json(*path):
    try:
        with open(os.path.join(ROOT, *path)) as fh:
            return json.load(fh)
    except Exception:
        return None


def measured_peak():
    d = read_json("MEASURED_PEAKS.json")
    if d and "hbm_gbs" in d:
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def l2_roof():
    d = read_json("profiles", "l2_peak.json")
    if d and d.get("l2_gbs"):
        return float(d["l2_gbs"]), ("measured: L2-resident in-place update, persistent sweeps with nothing ordering them "
                                    "(profiles/l2_peak.json, tools/l2_peak_probe.py; one launch per sweep reaches %s GB/s)" % d.get("l2_launch_per_sweep_gbs"))
    return None, "unmeasured (profiles/l2_peak.json absent)"


def ncu_traffic(key):
    """DRAM bytes per launch of the step kernel and per algorithmic byte, from the committed ncu capture of this build."""
    d = read_json("profiles", "traffic.json") or {}
    return d.get(key)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm = sorted(float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit())
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 7 for n, v in zip(names, r[3:7]) if v.lower().startswith("active")})
        pw = [float(r[2]) for r in self.rows if len(r) >= 7 and r[2].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": reasons}


# ------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the reference's own AVX+OpenMP program (oracle/_ref)
# ------------------------------------------------------------------------------------------------
def run_reference_binary(prec, nx, ny, steps, threads):
    exe = os.path.join(ROOT, "oracle", "_ref", f"lbref_fast_{prec}_{nx}x{ny}_{steps}")
    if not os.path.exists(exe):
        raise FileNotFoundError(f"{exe} missing: run `sh oracle/build_ref.sh` where /root/reference exists")
    env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="spread", OMP_PLACES="cores")
    out = subprocess.run([exe], capture_output=True, text=True, env=env, cwd="/tmp", timeout=1800).stdout
    m = re.search(r"Runtime: ([0-9.]+) Re (\S+)", out)
    if not m:
        raise RuntimeError("reference binary printed no Runtime line:\n" + out[-500:])
    secs = float(m.group(1))                                  # the program's own timer, LB.c:146-166
    return nx * ny * steps / secs / 1e6, secs


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def mem_available():
    try:
        return int(re.search(r"MemAvailable:\s+(\d+)", open("/proc/meminfo").read()).group(1)) * 1024
    except Exception:
        return 0


REF_FLAGS = "LB.c compiled -Ofast -funroll-loops -march=core-avx-i -fopenmp, prec_double_avx.h"


def cpu_baseline():
    """BASELINE configs[0]: the reference's own CPU-runnable case, 400x2000 DP."""
    thr = host_threads()
    mlups, secs = run_reference_binary("dp", C1["nx"], C1["ny"], 1000, thr)
    sample = "configs[0] 400x2000 DP, 1,000 of the 10,000 timesteps"
    if secs < 3.0:      # many-core host: afford the full job so the sample is ~10-30 s of CPU work
        mlups, secs = run_reference_binary("dp", C1["nx"], C1["ny"], 10000, thr)
        sample = "configs[0] 400x2000 DP, all 10,000 timesteps"
    return {"value": round(mlups, 2), "unit": "MLUPS", "cores": thr, "kind": "reference", "sample": f"{sample} ({secs:.2f} s; {REF_FLAGS})"}


def reference_arm(args):
    """The reference's CPU implementation on the headline workload: the weak-scaling unit 8192x32768 DP (its `I()` and
    `allocSize` widened to size_t by oracle/build_ref.sh -- int overflows at 2.4e9 elements, LB.c:72,136 -- nothing else
    touched).  Each step is a bounded sample: 10 timesteps of that lattice.  The reference is a single-node CPU program:
    whatever N, it runs the per-GPU unit."""
    if int(os.environ.get("RANK", "0")) != 0:
        return 0
    thr = host_threads()
    nx, ny, ts = WEAK["nx_per_gpu"], WEAK["ny"], 10
    need = 2.1 * 9 * nx * ny * 8
    if args.reference_lattice == "small" or mem_available() < need or not os.path.exists(os.path.join(ROOT, "oracle", "_ref", f"lbref_fast_dp_{nx}x{ny}_{ts}")):
        nx, ny, ts = C2["nx"], C2["ny"], 1000
        sample = ("each step = 1,000 timesteps of the 400x4000 DP lattice (configs[1]) -- " +
                  ("--reference-lattice small" if args.reference_lattice == "small" else
                   f"host memory {mem_available() / 1e9:.0f} GB cannot hold the {need / 1e9:.0f} GB of the 8192x32768 A+B buffers"))
    else:
        sample = "each step = 10 timesteps of the 8192x32768 DP lattice (a bounded sample of the 1,000-timestep job)"
    for _ in range(min(args.warmup, 1)):      # every run is a fresh process: one warm-up run pages the binary in
        run_reference_binary("dp", nx, ny, ts, thr)
    t_tot, upd = 0.0, 0
    for _ in range(args.steps):
        mlups, secs = run_reference_binary("dp", nx, ny, ts, thr)
        t_tot += secs
        upd += nx * ny * ts
        if elapsed() > args.budget_s:
            break
    value = upd / t_tot / 1e6
    line = {"impl": "reference", "metric": "MLUPS", "value": round(value, 2), "unit": "MLUPS", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(1e3 * t_tot / args.steps, 3),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": headline_workload(args.gpus), "sample": sample, "flags": REF_FLAGS},
            "cpu_baseline": {"value": round(value, 2), "unit": "MLUPS", "cores": thr, "kind": "reference", "sample": sample},
            "e2e": {"value": round(value, 2), "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)
    return 0


def headline_workload(n):
    return (f"BASELINE configs[4]: weak scaling, {WEAK['nx_per_gpu']}x{WEAK['ny']} sites per GPU, DP, {n} B200, "
            f"{WEAK['timesteps']} timesteps per step")


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def timed_jobs(run_timed, launches_of, timesteps, warmup, steps, reinit=None, sync=None):
    """W untimed + K timed jobs of `timesteps` lattice steps; returns (ms list, launches in the timed regions)."""
    for _ in range(warmup):
        if reinit:
            reinit()
        run_timed(timesteps)
    ms, launches = [], 0
    for _ in range(steps):
        if reinit:
            reinit()
        if sync:
            sync()
        l0 = launches_of()
        ms.append(run_timed(timesteps))
        launches += launches_of() - l0
    return ms, launches


def step_kernel_launches(launches, jobs):
    """each lbm_run adds ApplySource + halo refresh (2 launches, 3 on a flag-ordered slab); the rest are step launches"""
    return max(jobs, launches - 2 * jobs)


def roofline(prec, engine, sites, timesteps_total, kernel_launches, total_ms, traffic_key=None):
    """Dominant kernel = the fused stream-collide step.  Algorithmic bytes per launch = 18 reals per site update x the site
    updates one launch does on average (temporal blocking: several tile launches per timestep)."""
    hbm, hbm_src = measured_peak()
    l2, l2_src = l2_roof()
    per_launch_s = total_ms * 1e-3 / kernel_launches
    bytes_per_launch = BYTES_PER_UPDATE[prec] * sites * timesteps_total / kernel_launches
    achieved = bytes_per_launch / per_launch_s / 1e9
    tr = ncu_traffic(traffic_key) if traffic_key else None
    rec = {"kernel": "lbm::stream_collide_aa_kernel" if engine == "inplace" else "lbm::stream_collide_vec_kernel",
           "achieved": round(achieved, 1), "unit": "GB/s", "algorithmic_gbs": round(achieved, 1),
           "algorithmic_bytes_per_launch": int(round(bytes_per_launch)), "timesteps_per_launch": round(timesteps_total / kernel_launches, 4),
           "avg_launch_us": round(per_launch_s * 1e6, 3), "us_per_timestep": round(total_ms * 1e3 / timesteps_total, 3)}
    blocked = engine == "inplace" and kernel_launches > timesteps_total
    resident = engine == "inplace" and BYTES_PER_UPDATE[prec] // 2 * sites <= 60e6
    if (blocked or resident) and l2:
        # every load and store of the step kernel is served by L2 (tiles resident for k timesteps): L2 bandwidth is the roof
        rec.update({"bound": "l2", "peak": l2, "frac": round(achieved / l2, 4), "peak_source": l2_src})
        ratio = tr.get("dram_bytes_per_algorithmic_byte") if isinstance(tr, dict) else None
        rec["hbm"] = {"peak": hbm, "peak_source": hbm_src, "dram_bytes_per_algorithmic_byte": ratio,
                      "achieved_dram_gbs": round(achieved * ratio, 1) if ratio else None,
                      "frac": round(achieved * ratio / hbm, 4) if ratio else None,
                      "source": tr.get("source") if isinstance(tr, dict) else "no ncu capture committed for this key"}
    else:
        rec.update({"bound": "hbm", "peak": hbm, "frac": round(achieved / hbm, 4), "peak_source": hbm_src,
                    "frac_of_nominal_8TBs": round(achieved / 8000.0, 4)})
        if blocked or resident:
            rec["note"] = "L2-resident engine but no measured L2 roof committed: fraction is against HBM and may exceed 1"
    rec["traffic"] = tr.get("dram_bytes_per_launch") if isinstance(tr, dict) else tr
    return rec


def golden_digest(name):
    d = read_json("tests", "golden", "golden.json")
    return d[name]["fA_sha256"] if d and name in d else None


def weak_scene(nx, ny, perturb=False):
    sc = dict(bar_i0=0, bar_i1=0, disc_ci=nx // 2, disc_cj=ny // 5, disc_r=min(nx, ny) // 10)
    if perturb:
        sc.update(perturb_seed=12345, perturb_amp=0.1)
    return sc


def pinned(shape, dtype):
    import torch
    return torch.empty(shape, dtype=dtype, pin_memory=True).numpy()


def e2e_single(lat, prec, nx, ny, T, reps, f_in, w_in, f_out):
    """pinned host buffers -> lbm_upload -> lbm_run -> lbm_download, wall clock around the three calls, every step"""
    e2e_t, parts = [], [0.0, 0.0, 0.0]
    for k in range(1 + reps):
        lat.sync()
        t0 = time.perf_counter()
        lat.upload(f_in, w_in)
        t1 = time.perf_counter()
        lat.run(T)
        t2 = time.perf_counter()
        lat.download("A", out=f_out)
        t3 = time.perf_counter()
        if k:
            e2e_t.append(t3 - t0)
            parts = [parts[0] + t1 - t0, parts[1] + t2 - t1, parts[2] + t3 - t2]
    n = len(e2e_t)
    h2d, d2h = int(f_in.nbytes + w_in.nbytes), int(f_out.nbytes)
    return {"value": round(nx * ny * T * n / sum(e2e_t) / 1e6, 1), "unit": "MLUPS", "steps": n,
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": round(1e3 * sum(e2e_t) / n, 3),
            "host_ms": {"upload": round(1e3 * parts[0] / n, 2), "enqueue": round(1e3 * parts[1] / n, 2), "wait_and_download": round(1e3 * parts[2] / n, 2)},
            "note": f"pinned host buffers; upload {h2d / (parts[0] / n) / 1e9:.1f} GB/s H2D"}


def small_config(name, nx, ny, T, prec, scene_kw, golden, args, out, engines=("inplace",)):
    """BASELINE configs[1] / configs[2]: device-resident value, e2e through host buffers, golden stamp of the e2e output"""
    import numpy as np
    import torch
    from latticeboltzmann_b200 import Lattice, geometry
    K, W = min(args.steps, 5), min(max(args.warmup, 3), 3)
    for engine in engines:
        with Lattice(nx, ny, prec, inplace=(engine == "inplace")) as lat:
            init = lambda: lat.InitializeArrays(**scene_kw)  # noqa: E731
            ms, launches = timed_jobs(lat.run_timed, lambda: lat.launches, T, W, K, init, lat.sync)
            total_ms = sum(ms)
            res = {"value": round(nx * ny * T * K / (total_ms * 1e-3) / 1e6, 1), "unit": "MLUPS", "ms_per_step": round(total_ms / K, 3),
                   "steps": K, "warmup": W, "timesteps_per_step": T, "engine": engine, "launches": launches,
                   "roofline": roofline(prec, engine, nx * ny, T * K, step_kernel_launches(launches, K), total_ms, traffic_key=f"{prec}_{engine}_{nx}x{ny}"),
                   "tuning": lat.get_tuning()}
            if engine == "inplace":
                tdt = torch.float64 if prec == "dp" else torch.float32
                f_in, w_in, f_out = pinned((9, nx, ny), tdt), pinned((nx, ny), torch.int8), pinned((9, nx, ny), tdt)
                f_in[...] = geometry.equilibrium_rest(nx, ny, prec)
                if scene_kw:
                    w_in[...] = geometry.cylinder_walls(nx, ny, scene_kw["disc_ci"], scene_kw["disc_cj"], scene_kw["disc_r"])
                else:
                    w_in[...] = geometry.channel_walls(nx, ny)
                res["e2e"] = e2e_single(lat, prec, nx, ny, T, 3, f_in, w_in, f_out)
                want = golden_digest(golden)
                got = hashlib.sha256(np.ascontiguousarray(f_out).tobytes()).hexdigest()
                res["parity"] = (f"bit-exact vs compiled reference ({golden})" if want == got else
                                 f"unchecked (no golden {golden})" if want is None else f"MISMATCH vs compiled reference ({golden})")
                res["reynolds"] = lat.ComputeReynolds()
            out[f"{prec}_{engine}" if len(engines) > 1 or engine != "inplace" else prec] = res
    return out


def bench_single(args):
    import torch
    from latticeboltzmann_b200 import Lattice

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (there is no CPU fallback; use --impl reference for the CPU arm)")
    nx, ny, T = WEAK["nx_per_gpu"], WEAK["ny"], WEAK["timesteps"]
    sampler, clocks = ClockSampler(0), None
    head, parity, skipped = {}, [], []

    # ---- headline: the weak-scaling unit, DP then SP, library defaults (in place + 2-D temporal blocking) ----
    for prec in ("dp", "sp"):
        K, W = (args.steps, args.warmup) if prec == "dp" else (min(args.steps, 5), min(args.warmup, 3))
        with Lattice(nx, ny, prec, inplace=True) as lat:
            lat.InitializeArrays(**weak_scene(nx, ny))
            if prec == "dp":
                sampler.start()
            ms, launches = timed_jobs(lat.run_timed, lambda: lat.launches, T, W, K, None, lat.sync)
            if prec == "dp":
                clocks = sampler.stop()
            total_ms = sum(ms)
            res = {"value": round(nx * ny * T * K / (total_ms * 1e-3) / 1e6, 1), "unit": "MLUPS", "ms_per_step": round(total_ms / K, 3),
                   "steps": K, "warmup": W, "launches": launches, "ms_min": round(min(ms), 3), "ms_max": round(max(ms), 3),
                   "roofline": roofline(prec, "inplace", nx * ny, T * K, step_kernel_launches(launches, K), total_ms, traffic_key=f"{prec}_inplace_{nx}x{ny}"),
                   "reynolds": lat.ComputeReynolds(), "tuning": lat.get_tuning()}
            # e2e: the whole 19.3 GB (DP) lattice up from pinned host memory, T timesteps, and down again, every step
            tdt = torch.float64 if prec == "dp" else torch.float32
            need = 9 * nx * ny * (8 if prec == "dp" else 4)
            if mem_available() > 1.5 * need:
                f_host, w_host = pinned((9, nx, ny), tdt), lat.download_walls()
                lat.download("A", out=f_host)
                res["e2e"] = e2e_single(lat, prec, nx, ny, T, min(args.steps, 3 if prec == "dp" else 2), f_host, w_host, f_host)
                del f_host
            else:
                res["e2e"] = {"value": None, "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0, "note": "skipped: host memory too small"}
            # parity at the timed lattice: two independent engines must leave the same checksum of fA
            lat.InitializeArrays(**weak_scene(nx, ny, perturb=True))
            lat.run(CHECK_STEPS)
            s_inplace = lat.checksum()
        with Lattice(nx, ny, prec, inplace=False) as ab:
            ab.InitializeArrays(**weak_scene(nx, ny, perturb=True))
            ab.run(CHECK_STEPS)
            same = ab.checksum() == s_inplace
            parity.append(f"{prec} in-place+TB {'==' if same else '!= (MISMATCH)'} A/B checksum @{nx}x{ny} ({CHECK_STEPS} steps, rough initial state)")
            # the A/B engine of the same library on the same lattice: the HBM-streaming regime SURVEY.md 8(d) describes
            ab.InitializeArrays(**weak_scene(nx, ny))
            Ta = 100
            ms, launches = timed_jobs(ab.run_timed, lambda: ab.launches, Ta, 1, 3, None, ab.sync)
            res["ab"] = {"value": round(nx * ny * Ta * 3 / (sum(ms) * 1e-3) / 1e6, 1), "unit": "MLUPS", "ms_per_step": round(sum(ms) / 3, 3),
                         "timesteps_per_step": Ta, "steps": 3, "warmup": 1, "engine": "ab",
                         "roofline": roofline(prec, "ab", nx * ny, Ta * 3, step_kernel_launches(launches, 3), sum(ms), traffic_key=f"{prec}_ab_{nx}x{ny}"),
                         "tuning": ab.get_tuning()}
        head[prec] = res

    # ---- BASELINE configs[1] and configs[2] ----
    c2, c3 = {}, {}
    small_config("c2", C2["nx"], C2["ny"], C2["timesteps"], "dp", {}, "c2_dp_10000", args, c2, engines=("inplace", "ab"))
    small_config("c2", C2["nx"], C2["ny"], C2["timesteps"], "sp", {}, "c2_sp_10000", args, c2, engines=("inplace", "ab"))
    cyl = dict(bar_i0=0, bar_i1=0, disc_ci=C3["nx"] // 2, disc_cj=C3["ny"] // 5, disc_r=C3["nx"] // 10)
    if elapsed() < args.budget_s:
        small_config("c3", C3["nx"], C3["ny"], C3["timesteps"], "dp", cyl, "cyl_dp_10000", args, c3)
    else:
        skipped.append("c3_800x4000_cylinder")
    for rec in (c2.get("dp_inplace"), c2.get("sp_inplace"), c3.get("dp")):
        if rec:
            parity.append(rec["parity"])

    # ---- BASELINE configs[3] on one GPU: only the in-place engine (one 77.3 GB buffer) fits ----
    strong = None
    if elapsed() < args.budget_s:
        sx, sy, st = STRONG["nx"], STRONG["ny"], STRONG["timesteps"]
        with Lattice(sx, sy, "dp", inplace=True) as lat:
            lat.InitializeArrays(**weak_scene(sx, sy))
            ms, launches = timed_jobs(lat.run_timed, lambda: lat.launches, st, 1, 3, None, lat.sync)
            strong = {"value": round(sx * sy * st * 3 / (sum(ms) * 1e-3) / 1e6, 1), "unit": "MLUPS", "ms_per_step": round(sum(ms) / 3, 3),
                      "lattice": f"{sx}x{sy}", "timesteps_per_step": st, "steps": 3, "warmup": 1, "scaling": "strong",
                      "engine": "in-place single buffer (77.3 GB; A+B would need 154.6 GB)",
                      "roofline": roofline("dp", "inplace", sx * sy, st * 3, step_kernel_launches(launches, 3), sum(ms), traffic_key="dp_inplace_8192x32768"),
                      "reynolds": lat.ComputeReynolds()}
    else:
        skipped.append("strong_16384x65536")

    try:
        cpu = cpu_baseline()
    except Exception as exc:
        cpu = {"value": None, "unit": "MLUPS", "cores": host_threads(), "kind": "reference", "sample": f"failed: {exc}"}
    dp = head["dp"]
    ok = all("MISMATCH" not in p and "unchecked" not in p for p in parity)
    line = {"metric": "MLUPS", "value": dp["value"], "unit": "MLUPS", "n_gpus": 1, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dp["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": headline_workload(1), "lattice": f"{nx}x{ny}", "timesteps_per_step": T,
                       "scene": "channel walls + disc obstacle (InitializeArrays closed form), tau 0.7, accel 0.005",
                       "engine": "in-place single buffer (lbm_config.inplace=1, AA kernels) + 2-D temporal blocking in L2, two concurrent tile chains (library defaults); "
                                 "the A/B double-buffer engine of the same library is under 'ab_double_buffer'",
                       "l2": "19.3 GB lattice, far larger than L2: nothing is flushed because keeping two 24 MB (SP: 36 MB) tiles L2-resident for 16 timesteps IS the "
                             "engine (HBM sees each site once per several steps; the roof that binds is L2 bandwidth, see roofline)",
                       "parity": ("bit-exact: " if ok else "NOT CLEAN: ") + "; ".join(parity),
                       "skipped": skipped},
            "roofline": dp["roofline"], "cpu_baseline": cpu, "e2e": dp["e2e"], "gpu_launches": dp["launches"],
            "clocks": clocks, "reynolds": dp["reynolds"], "tuning": dp["tuning"],
            "sp": {k: head["sp"][k] for k in ("value", "ms_per_step", "steps", "warmup", "roofline", "e2e", "reynolds", "tuning")},
            "ab_double_buffer": {p: head[p]["ab"] for p in ("dp", "sp")},
            "c2_400x4000": c2, "c3_800x4000_cylinder": c3, "strong_16384x65536": strong,
            "elapsed_s": round(elapsed(), 1)}
    print(json.dumps(line), flush=True)
    return 0


def bench_multi(args):
    """torchrun, one rank per GPU: weak scaling, 8192x32768 sites per GPU, DP headline (+SP), strong scaling beside it."""
    import torch
    import torch.distributed as dist
    from latticeboltzmann_b200.ringcheck import ring_parity
    from latticeboltzmann_b200.slab import DistributedLattice

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    # NCCL prints its version banner on fd 1 whatever NCCL_DEBUG says: point fd 1 at stderr until the JSON line
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    inplace = args.engine == "inplace"
    parity, skipped = [], []

    def max_over_ranks(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def min_over_ranks(x):
        return -max_over_ranks(-x)

    def ring_sum(dl):
        sums = [None] * world
        dist.all_gather_object(sums, dl.lat.checksum())
        return sum(sums) % 2 ** 64

    # ---- parity before anything is timed: the ring over these GPUs == one GPU, bit for bit ----
    if not args.no_parity:
        t0 = time.time()
        okc, total = ring_parity(local, log=(lambda s: print(s, file=sys.stderr, flush=True)) if rank == 0 else None)
        parity.append(f"ring over {world} GPUs {okc}/{total} cases bit-exact vs single GPU{'' if okc == total else ' (MISMATCH)'} ({time.time() - t0:.0f} s)")

    def timed(dl, T, W, K, sampler=None):
        for _ in range(W):
            dist.barrier()
            dl.run_timed(T)
        if sampler:
            sampler.start()
        total_ms, launches, wall = 0.0, 0, 0.0
        for _ in range(K):
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            l0, t0 = dl.lat.launches, time.perf_counter()
            ms = dl.run_timed(T)
            torch.cuda.synchronize()
            wall += max_over_ranks(time.perf_counter() - t0)
            total_ms += max_over_ranks(ms)                       # device time, max over ranks
            launches += dl.lat.launches - l0
        return total_ms, launches, wall, (sampler.stop() if sampler else None)

    # ---- headline: weak scaling, 8192x32768 per GPU ----
    nx, ny, T = WEAK["nx_per_gpu"] * world, WEAK["ny"], WEAK["timesteps"]
    rows = nx // world
    out, clocks = {}, None
    for prec in ("dp", "sp"):
        K, W = (args.steps, args.warmup) if prec == "dp" else (min(args.steps, 5), min(args.warmup, 3))
        dl = DistributedLattice(nx, ny, prec, transport=args.transport, device=local, inplace=inplace)
        if not args.no_parity:                                   # cross-engine checksum of the timed lattice, rough initial state
            dl.lat.InitializeArrays(**weak_scene(nx, ny, perturb=True))
            dl.run(CHECK_STEPS)
            s_mine = ring_sum(dl)
            other = DistributedLattice(nx, ny, prec, transport=args.transport, device=local, inplace=not inplace)
            other.lat.InitializeArrays(**weak_scene(nx, ny, perturb=True))
            other.run(CHECK_STEPS)
            same = ring_sum(other) == s_mine
            other.close()
            parity.append(f"{prec} in-place+TB ring {'==' if same else '!= (MISMATCH)'} A/B ring checksum @{nx}x{ny} ({CHECK_STEPS} steps, rough initial state)")
        dl.lat.InitializeArrays(**weak_scene(nx, ny))
        dl.lat.sync()
        total_ms, launches, wall, ck = timed(dl, T, W, K, ClockSampler(local) if (prec == "dp" and rank == 0) else None)
        clocks = ck or clocks
        res = {"value": round(nx * ny * T * K / (total_ms * 1e-3) / 1e6, 1), "unit": "MLUPS", "ms_per_step": round(total_ms / K, 3),
               "steps": K, "warmup": W, "launches": launches, "wall_ms_per_step": round(1e3 * wall / K, 3),
               "roofline": roofline(prec, args.engine, rows * ny, T * K, max(K, launches - 3 * K), total_ms,
                                    traffic_key=f"{prec}_{args.engine}_8192x32768"),
               "reynolds": dl.gather_reynolds()}

        # e2e: every rank uploads its slab from PINNED host memory, runs T timesteps, downloads it again; >= 3 steps
        need = 9 * rows * ny * (8 if prec == "dp" else 4)
        fits = min_over_ranks(1.0 if mem_available() > 1.6 * need * world else 0.0) > 0
        if fits and not args.no_e2e and elapsed() < args.budget_s:
            tdt = torch.float64 if prec == "dp" else torch.float32
            slab = pinned((9, rows, ny), tdt)                    # host array of this rank's rows (lbm_upload_slab / lbm_download_slab)
            dl.lat.sync()
            dl.lat.download_slab(slab)
            walls = dl.lat.download_slab(walls=True)
            reps, e2e_s, up_s = (3 if prec == "dp" else 2), 0.0, 0.0
            for k in range(1 + reps):
                dist.barrier()
                t0 = time.perf_counter()
                dl.lat.upload_slab(slab, walls)
                t1 = time.perf_counter()
                dl.run(T)
                dl.lat.download_slab(slab)
                dt, du = max_over_ranks(time.perf_counter() - t0), max_over_ranks(t1 - t0)
                if k:
                    e2e_s, up_s = e2e_s + dt, up_s + du
            res["e2e"] = {"value": round(nx * ny * T * reps / e2e_s / 1e6, 1), "unit": "MLUPS", "steps": reps,
                          "h2d_bytes_per_step": int(slab.nbytes + walls.nbytes) * world, "d2h_bytes_per_step": int(slab.nbytes) * world,
                          "ms_per_step": round(1e3 * e2e_s / reps, 3),
                          "note": f"pinned host slabs ({slab.nbytes / 1e9:.1f} GB per rank), all ranks copy concurrently; "
                                  f"upload {(slab.nbytes + walls.nbytes) / (up_s / reps) / 1e9:.1f} GB/s H2D per GPU (max-over-ranks time)"}
            del slab
        else:
            res["e2e"] = {"value": None, "unit": "MLUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                          "note": "skipped (--no-e2e)" if args.no_e2e else "skipped: host memory too small for pinned per-rank slabs, or time budget"}
        out[prec] = res
        dl.close()
        dist.barrier()

    # ---- BASELINE configs[3]: strong scaling, 16384x65536 DP over the same ranks ----
    strong = None
    if elapsed() < args.budget_s and not args.no_strong:
        sx, sy, st = STRONG["nx"], STRONG["ny"], STRONG["timesteps"]
        dl = DistributedLattice(sx, sy, "dp", transport=args.transport, device=local, inplace=inplace)
        dl.lat.InitializeArrays(**weak_scene(sx, sy))
        dl.lat.sync()
        total_ms, launches, wall, _ = timed(dl, st, 1, 3)
        strong = {"value": round(sx * sy * st * 3 / (total_ms * 1e-3) / 1e6, 1), "unit": "MLUPS", "ms_per_step": round(total_ms / 3, 3),
                  "lattice": f"{sx}x{sy}", "timesteps_per_step": st, "steps": 3, "warmup": 1, "scaling": "strong", "n_gpus": world,
                  "rows_per_gpu": sx // world,
                  "roofline": roofline("dp", args.engine, (sx // world) * sy, st * 3, max(3, launches - 9), total_ms, traffic_key=f"dp_{args.engine}_8192x32768"),
                  "reynolds": dl.gather_reynolds(),
                  "reference_shape": "mpi-runtimes.dat:62-79 (the reference's MPI strong scaling: 8.37x on 13 nodes)"}
        dl.close()
        dist.barrier()
    else:
        skipped.append("strong_16384x65536")

    if rank == 0:
        dp = out["dp"]
        ok = all("MISMATCH" not in p for p in parity) and bool(parity)
        line = {"metric": "MLUPS", "value": dp["value"], "unit": "MLUPS", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": dp["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": headline_workload(world), "lattice": f"{nx}x{ny}", "timesteps_per_step": T,
                           "scene": "channel walls + disc obstacle (InitializeArrays closed form), tau 0.7, accel 0.005",
                           "parallelism": f"row slabs x{world}, halo rows stored peer-to-peer over NVLink ({args.transport}), seam chain on a second stream",
                           "engine": ("in-place single buffer per slab (AA kernels) + 2-D temporal blocking in L2 across the slab seams, two concurrent tile chains per slab (library defaults)"
                                      if inplace else "A/B double buffer per slab, one pull kernel per timestep (HBM-bound)"),
                           "l2": "19.3 GB per GPU, far larger than L2; nothing is flushed: L2 residency of two ~30 MB tiles for 16 timesteps is the engine",
                           "parity": ("bit-exact: " if ok else "NOT CLEAN: ") + ("; ".join(parity) if parity else "skipped (--no-parity)"),
                           "skipped": skipped},
                "roofline": dp["roofline"], "e2e": dp["e2e"], "gpu_launches": dp["launches"], "clocks": clocks,
                "reynolds": dp["reynolds"], "wall_ms_per_step": dp["wall_ms_per_step"],
                "sp": {k: out["sp"][k] for k in ("value", "ms_per_step", "steps", "warmup", "roofline", "e2e", "reynolds")},
                "strong_16384x65536": strong, "elapsed_s": round(elapsed(), 1)}
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--transport", default="p2p", choices=["p2p", "staged"])
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the ring-parity pass and the cross-engine checksum")
    ap.add_argument("--no-strong", action="store_true", help="N > 1: skip the strong-scaling sub-record")
    ap.add_argument("--budget-s", type=float, default=540.0,
                    help="optional sub-records (configs[2], configs[3], e2e at N > 1) are skipped, and named in config.skipped, once the run is this old")
    ap.add_argument("--reference-lattice", default="auto", choices=["auto", "small"],
                    help="--impl reference: auto = the headline lattice 8192x32768 (needs ~41 GB of host memory), small = 400x4000")
    ap.add_argument("--engine", default="inplace", choices=["ab", "inplace"],
                    help="N > 1: slab engine -- inplace (default) = one buffer per slab, in-place kernels, 2-D temporal blocking in L2 "
                         "across the slab seams; ab = the reference's fA/fB pair, one pull kernel per timestep, HBM-bound")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        return bench_multi(args)
    if args.gpus > 1:
        raise SystemExit("bench.py: launch N > 1 with torchrun (one rank per GPU), e.g. python -m torch.distributed.run "
                         "--nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29500 bench.py --gpus N")
    return bench_single(args)


if __name__ == "__main__":
    sys.exit(main())
