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
    agg = collections.OrderedDict()
    total = 0.0
    for r in rows:
        k = short(r["Kernel Name"])
        ns = float(r["Metric Value"].replace(",", ""))
        d = agg.setdefault(k, [0, 0.0])
        d[0] += 1
        d[1] += ns
        total += ns
    print(f"# {len(rows)} launches, {total / 1e6:.3f} ms total device time (cold-cache, serialised under ncu)")
    print(f"{'kernel':<100} {'n':>5} {'total_us':>10} {'mean_us':>9} {'share':>6}")
    for k, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k[:100]:<100} {n:>5} {ns / 1e3:>10.1f} {ns / 1e3 / n:>9.1f} {ns / total:>6.1%}")


if __name__ == "__main__":
    main()
