"""How much of a training step is the GPU actually busy?  Runs C2-shaped steps under torch.profiler (CUDA
activity only), merges the kernel / memcpy intervals of all streams and reports busy time per step, the
idle gaps and which kernel precedes the longest gaps.  The unprofiled step time is measured first with
CUDA events for comparison (CUPTI tracing slows the host side down, not the kernels).
usage: gpu_busy.py [cells] [steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
from torch.profiler import ProfilerActivity, profile   # noqa: E402
import bench                                    # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 40
sys.argv = [sys.argv[0], "--cells", str(cells)]
args = bench.parse_args()
torch.cuda.set_device(0)
wl = dict(bench.WORKLOADS["c2"])
wl["n_cells"] = cells
adata = bench.build_adata(wl, 0)
model, tr = bench.make_trainer(adata, wl, args)
it = bench.batch_stream(tr)
bench.run_steps(tr, it, 10)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
bench.run_steps(tr, it, STEPS)
e1.record()
torch.cuda.synchronize()
plain_ms = e0.elapsed_time(e1) / STEPS

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    e0.record()
    bench.run_steps(tr, it, STEPS)
    e1.record()
    torch.cuda.synchronize()
prof_ms = e0.elapsed_time(e1) / STEPS
iv = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start:
        iv.append((e.time_range.start, e.time_range.end, e.name))
iv.sort()
busy, gaps, cur_s, cur_e, last_name = 0.0, [], None, None, None
ksum = {}
for s, e, n in iv:
    ksum[n] = ksum.get(n, 0.0) + (e - s)
    if cur_e is None:
        cur_s, cur_e, last_name = s, e, n
        continue
    if s > cur_e:
        busy += cur_e - cur_s
        gaps.append((s - cur_e, last_name, n))
        cur_s, cur_e, last_name = s, e, n
    elif e > cur_e:
        cur_e, last_name = e, n
busy += cur_e - cur_s
span = iv[-1][1] - iv[0][0]
gaps.sort(reverse=True)
by_prev = {}
for g, a, b in gaps:
    k = a[:60]
    d = by_prev.setdefault(k, [0, 0.0])
    d[0] += 1
    d[1] += g
out = {"cells": cells, "steps": STEPS, "step_ms_unprofiled": plain_ms, "step_ms_under_profiler": prof_ms,
       "span_ms_per_step": span / 1e3 / STEPS, "busy_ms_per_step": busy / 1e3 / STEPS,
       "idle_ms_per_step": (span - busy) / 1e3 / STEPS, "kernels_per_step": len(iv) / STEPS,
       "kernel_time_sum_ms_per_step": sum(ksum.values()) / 1e3 / STEPS,
       "gaps_per_step": len(gaps) / STEPS,
       "gaps_over_20us_per_step": sum(1 for g in gaps if g[0] > 20) / STEPS,
       "idle_in_gaps_over_20us_ms_per_step": sum(g[0] for g in gaps if g[0] > 20) / 1e3 / STEPS}
print(json.dumps(out))
print("# idle time by preceding kernel (us per step, count per step)")
for k, (c, t) in sorted(by_prev.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"{t / STEPS:9.1f} us {c / STEPS:6.1f}  {k}")
print("# kernel time (us per step)")
for k, t in sorted(ksum.items(), key=lambda kv: -kv[1])[:25]:
    print(f"{t / STEPS:9.1f} us  {k[:90]}")
