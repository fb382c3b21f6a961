#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel launches, total
and mean device time, share.  usage: summarize_launches.py launches.csv [--skip N] [--last N]"""
import csv
import collections
import re
import sys


def short(name: str) -> str:
    name = re.sub(r"\(anonymous namespace\)::", "", name)
    m = re.match(r"(?:void )?([\w:]+)", name)
    base = m.group(1) if m else name[:60]
    t = re.search(r"<(.{0,70})", name)
    return base + ("<" + t.group(1) if t and not base.startswith("at::") else "")


def main():
    path = sys.argv[1]
    skip = int(sys.argv[sys.argv.index("--skip") + 1]) if "--skip" in sys.argv else 0
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else None
    with open(path) as fh:
        lines = [l for l in fh if not l.startswith("==")]
    rows = list(csv.DictReader(lines))
    rows = rows[skip:]
    if last:
        rows = rows[-last:]
    # one row per (launch ID, metric): fold the metrics of a launch together
    launches = collections.OrderedDict()
    for r in rows:
        d = launches.setdefault(r["ID"], {"name": short(r["Kernel Name"])})
        d[r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
    agg = collections.OrderedDict()
    total = 0.0
    has_dram = any("dram__bytes_read.sum" in d for d in launches.values())
    for d in launches.values():
        ns = d.get("gpu__time_duration.sum", 0.0)
        a = agg.setdefault(d["name"], [0, 0.0, 0.0])
        a[0] += 1
        a[1] += ns
        a[2] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        total += ns
    print(f"# {len(launches)} launches, {total / 1e6:.3f} ms total device time (cold-cache, serialised under ncu)")
    hdr = f"{'kernel':<100} {'n':>5} {'total_us':>10} {'mean_us':>9} {'share':>6}"
    print(hdr + (f" {'dram_MB/launch':>15} {'dram_GB/s':>10}" if has_dram else ""))
    for k, (n, ns, by) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        line = f"{k[:100]:<100} {n:>5} {ns / 1e3:>10.1f} {ns / 1e3 / n:>9.1f} {ns / total:>6.1%}"
        if has_dram:
            line += f" {by / n / 1e6:>15.3f} {by / max(ns, 1.0):>10.1f}"
        print(line)


if __name__ == "__main__":
    main()
