"""ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --cache-control none --csv)
of tools/profile_target.py  ->  one entry of profiles/traffic.json:

    python tools/traffic_from_ncu.py launches.csv <key> <prec> <nx> <ny> <nsteps> [kernel-substring] >> (merged into profiles/traffic.json)
    python tools/traffic_from_ncu.py --summary launches.csv out.summary.csv        (per-kernel table)

dram_bytes_per_algorithmic_byte = sum of DRAM bytes over ALL step-kernel launches of the run / (18 reals x sites x timesteps):
the capture must cover whole macro-steps (profile_target's nsteps a multiple of tb_depth, -c large enough)."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def parse(path, needle):
    rows = {}
    with open(path, newline="") as fh:
        lines = [ln for ln in fh if not ln.startswith("==")]
    for r in csv.DictReader(lines):
        if needle not in r.get("Kernel Name", ""):
            continue
        d = rows.setdefault(r["ID"], {"name": r["Kernel Name"]})
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "")
        scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1, "us": 1e3, "ms": 1e6, "nsecond": 1, "usecond": 1e3, "msecond": 1e6, "second": 1e9}.get(unit, 1)
        d[r["Metric Name"]] = v * scale
    return list(rows.values())


def summary(path, out):
    """per-kernel table of a launch list: launches, time, DRAM bytes (what profiles/*/launches_*.summary.csv hold)"""
    ls = parse(path, "")
    tab = {}
    for d in ls:
        name = d["name"].split("(")[0]
        t = tab.setdefault(name, [0, 0.0, 0.0, 0.0])
        t[0] += 1
        t[1] += d.get("gpu__time_duration.sum", 0) / 1e3
        t[2] += d.get("dram__bytes_read.sum", 0) / 1e6
        t[3] += d.get("dram__bytes_write.sum", 0) / 1e6
    total = sum(t[1] for t in tab.values()) or 1.0
    with open(out, "w") as fh:
        fh.write("kernel,launches,total_time_us,avg_time_us,dram_read_MB,dram_write_MB,share_of_time\n")
        for name, t in tab.items():
            fh.write(f'"{name}",{t[0]},{t[1]:.1f},{t[1] / t[0]:.2f},{t[2]:.1f},{t[3]:.1f},{t[1] / total:.4f}\n')


def main():
    if sys.argv[1] == "--summary":
        return summary(sys.argv[2], sys.argv[3])
    path, key, prec, nx, ny, nsteps = sys.argv[1], sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6])
    needle = sys.argv[7] if len(sys.argv) > 7 else "stream_collide"
    ls = parse(path, needle)
    dram = sum(d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0) for d in ls)
    dur = sum(d.get("gpu__time_duration.sum", 0) for d in ls)
    algo = (144 if prec == "dp" else 72) * nx * ny * nsteps
    rec = {"dram_bytes_per_launch": int(dram / len(ls)), "algorithmic_bytes_per_launch": int(algo / len(ls)),
           "dram_bytes_per_algorithmic_byte": round(dram / algo, 4), "launches": len(ls), "timesteps": nsteps, "lattice": f"{nx}x{ny}",
           "sum_kernel_time_ms_under_ncu": round(dur / 1e6, 3),
           "source": f"ncu launch list, single pass, --cache-control none ({os.path.basename(path)}); all {len(ls)} step-kernel launches of {nsteps} timesteps"}
    tj = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        d = json.load(open(tj))
    except Exception:
        d = {}
    d[key] = rec
    json.dump(d, open(tj, "w"), indent=1)
    print(key, json.dumps(rec))


if __name__ == "__main__":
    main()
