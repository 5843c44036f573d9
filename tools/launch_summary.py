#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv --log-file X` launch list: per-kernel count, total time, share.

    python tools/launch_summary.py gpurun_out/launches.csv [--steps N] > profiles/launches_summary.txt

The durations are per-launch device times under ncu (serialised, cold caches): compare SHARES, not absolutes."""
import csv
import re
import sys
from collections import defaultdict


def main():
    path = sys.argv[1]
    steps = 1
    if "--steps" in sys.argv:
        steps = int(sys.argv[sys.argv.index("--steps") + 1])
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    rd = csv.reader(lines)
    header = None
    for r in rd:
        if header is None:
            if "Kernel Name" in r:
                header = r
            continue
        if len(r) != len(header):
            continue
        rows.append(dict(zip(header, r)))
    tot = defaultdict(float)
    cnt = defaultdict(int)
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*$", "", r["Kernel Name"])
        name = re.sub(r"^void ", "", name)
        name = re.sub(r"\(anonymous namespace\)::", "", name)
        v = float(r["Metric Value"].replace(",", ""))
        unit = r.get("Metric Unit", "ns")
        us = v / 1000.0 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1000.0)
        tot[name] += us
        cnt[name] += 1
    total = sum(tot.values())
    n = sum(cnt.values())
    print("# window: %d launches, %.1f us total (%d step(s): %.1f us and %.0f launches per step)" % (n, total, steps, total / steps, n / steps))
    print("%-86s %7s %10s %7s %8s" % ("kernel", "count", "total_us", "share", "us/launch"))
    for k in sorted(tot, key=lambda k: -tot[k]):
        print("%-86s %7d %10.1f %6.1f%% %8.1f" % (k[:86], cnt[k], tot[k], 100.0 * tot[k] / total, tot[k] / cnt[k]))


if __name__ == "__main__":
    main()
