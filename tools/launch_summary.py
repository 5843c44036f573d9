#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv --log-file X`
launch list: per-kernel count, total time, share and (when collected) DRAM bytes moved, GB/s, time-weighted SM throughput % of
peak, active warps % of peak and tensor-pipe instructions (non-zero only for the tcgen05 kernels).

    python tools/launch_summary.py gpurun_out/launches.csv [--steps N] [--list] > profiles/launches_summary.txt

The durations are per-launch device times under ncu (serialised, cold caches): compare SHARES, not absolutes.
--list appends every launch in issue order (id, kernel, grid, block, us, MB read, MB written)."""
import csv
import re
import sys
from collections import defaultdict

UNIT = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}
BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def short(name):
    name = re.sub(r"\(.*$", "", name)
    name = re.sub(r"^void ", "", name)
    return re.sub(r"\(anonymous namespace\)::", "", name)


def main():
    path = sys.argv[1]
    steps = 1
    if "--steps" in sys.argv:
        steps = int(sys.argv[sys.argv.index("--steps") + 1])
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    header = None
    launches = {}
    order = []
    for r in csv.reader(lines):
        if header is None:
            if "Kernel Name" in r:
                header = r
            continue
        if len(r) != len(header):
            continue
        d = dict(zip(header, r))
        lid = d.get("ID")
        if lid not in launches:
            launches[lid] = {"name": short(d["Kernel Name"]), "grid": d.get("Grid Size", ""), "block": d.get("Block Size", ""),
                             "us": 0.0, "rd": 0.0, "wr": 0.0, "smpct": None, "tensor": 0.0, "warps": None}
            order.append(lid)
        m = d.get("Metric Name")
        try:
            v = float(d["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        u = d.get("Metric Unit", "")
        if m == "gpu__time_duration.sum":
            launches[lid]["us"] = v * UNIT.get(u, 1e-3)
        elif m == "dram__bytes_read.sum":
            launches[lid]["rd"] = v * BYTES.get(u, 1.0)
        elif m == "dram__bytes_write.sum":
            launches[lid]["wr"] = v * BYTES.get(u, 1.0)
        elif m == "sm__throughput.avg.pct_of_peak_sustained_elapsed":
            launches[lid]["smpct"] = v
        elif m == "sm__inst_executed_pipe_tensor.sum":
            launches[lid]["tensor"] = v
        elif m == "sm__warps_active.avg.pct_of_peak_sustained_active":
            launches[lid]["warps"] = v
    tot, cnt, byt = defaultdict(float), defaultdict(int), defaultdict(float)
    smw, wrw, tens = defaultdict(float), defaultdict(float), defaultdict(float)
    for l in launches.values():
        tot[l["name"]] += l["us"]
        cnt[l["name"]] += 1
        byt[l["name"]] += l["rd"] + l["wr"]
        smw[l["name"]] += (l["smpct"] or 0.0) * l["us"]   # time-weighted
        wrw[l["name"]] += (l["warps"] or 0.0) * l["us"]
        tens[l["name"]] += l["tensor"]
    total = sum(tot.values())
    n = sum(cnt.values())
    allb = sum(byt.values())
    print("# window: %d launches, %.1f us total (%d step(s): %.1f us and %.0f launches per step), DRAM %.1f MB per step" % (
        n, total, steps, total / max(steps, 1), n / max(steps, 1), allb / max(steps, 1) / 1e6))
    print("%-70s %6s %10s %7s %9s %9s %8s %7s %8s %10s" % ("kernel", "count", "total_us", "share", "us/launch", "DRAM_MB", "GB/s", "SM%", "warps%", "tensor_ops"))
    for k in sorted(tot, key=lambda k: -tot[k]):
        print("%-70s %6d %10.1f %6.1f%% %9.1f %9.1f %8.0f %7.1f %8.1f %10.0f" % (
            k[:70], cnt[k], tot[k], 100.0 * tot[k] / max(total, 1e-9), tot[k] / cnt[k], byt[k] / 1e6, byt[k] / max(tot[k], 1e-9) / 1e3,
            smw[k] / max(tot[k], 1e-9), wrw[k] / max(tot[k], 1e-9), tens[k]))
    if "--list" in sys.argv:
        print("\n# launches in issue order")
        for lid in order:
            l = launches[lid]
            print("%5s %-70s %-16s %-14s %8.1f us %8.2f MB rd %8.2f MB wr" % (lid, l["name"][:70], l["grid"], l["block"], l["us"],
                                                                             l["rd"] / 1e6, l["wr"] / 1e6))


if __name__ == "__main__":
    main()
