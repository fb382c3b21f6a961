"""Readable digest of a bench.py JSON line: python tools/print_bench_line.py <file>  (or stdin)."""
import json
import sys

src = open(sys.argv[1]).read() if len(sys.argv) > 1 else sys.stdin.read()
line = [ln for ln in src.splitlines() if ln.strip().startswith("{")][-1]
d = json.loads(line)
e = d.get("e2e") or {}
print("value", round(d["value"]), "cells/s", round(d["ms_per_step"], 3), "ms/step | e2e", round(e.get("value", 0)),
      round(e.get("ms_per_step", 0), 3), "ms | n_gpus", d["n_gpus"], "| launches", d.get("gpu_launches"))
print("clocks", d.get("clocks"))
r = d.get("roofline") or {}
print("roofline", {k: r.get(k) for k in ("kernel", "achieved", "peak", "frac", "traffic")})
for r2 in d.get("rooflines", []):
    print("roofline+", {k: r2.get(k) for k in ("kernel", "achieved", "peak", "frac", "traffic")})
if "parity" in d:
    print("parity", d["parity"])
for k, v in (d.get("kernels") or {}).items():
    print(f"  {k:28s} {v['launches_per_step']:5.1f}/step {v['ms_per_step']:.4f} ms {v['share']:.3f} "
          f"{v['achieved']:9.1f} {v['unit']} {v['frac_of_peak']:.3f}")
if "cpu_baseline" in d:
    print("cpu_baseline", d["cpu_baseline"])
