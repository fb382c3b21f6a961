#!/usr/bin/env python
"""Trim `ncu -i rep --page raw --csv` to the columns the roofline discussion uses.
usage: ncu_raw_summary.py raw.csv out.csv"""
import csv
import re
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [k for k in KEEP if k in idx]
    with open(sys.argv[2], "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["kernel"] + [f"{c} [{units[idx[c]]}]" for c in cols])
        for r in data:
            nm = r[idx["Kernel Name"]]
            m = re.search(r"k_\w+(<[^>]*>)?", nm)
            w.writerow([m.group(0) if m else nm[:60]] + [r[idx[c]] for c in cols])
    print(f"{len(data)} kernels -> {sys.argv[2]}")


if __name__ == "__main__":
    main()
