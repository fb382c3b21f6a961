"""Which torch (aten) ops still launch kernels inside one training step?  Runs a few C2-shaped steps
under torch.profiler and prints the aten ops with device time, per step.  usage: glue_probe.py [cells]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
from torch.profiler import ProfilerActivity, profile   # noqa: E402
import bench                                    # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
sys.argv = [sys.argv[0], "--cells", str(cells)]
args = bench.parse_args()
torch.cuda.set_device(0)
wl = dict(bench.WORKLOADS["c2"])
wl["n_cells"] = cells
adata = bench.build_adata(wl, 0)
model, tr = bench.make_trainer(adata, wl, args)
it = bench.batch_stream(tr)
bench.run_steps(tr, it, 5)
torch.cuda.synchronize()
STEPS = 4
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True,
             experimental_config=torch._C._profiler._ExperimentalConfig(verbose=True)) as prof:
    bench.run_steps(tr, it, STEPS)
    torch.cuda.synchronize()
rows = {}
for e in prof.events():
    if not e.name.startswith("aten::") or e.cpu_parent is not None and e.cpu_parent.name.startswith("aten::"):
        continue                                    # top-level aten ops only
    site = next((f for f in (e.stack or []) if "nichecompass_b200/" in f or "bench.py" in f), "?")
    site = site.split("nichecompass_b200/")[-1][:70]
    k = (e.name, site)
    r = rows.setdefault(k, [0.0, 0.0, 0])
    r[0] += getattr(e, "device_time_total", 0) / STEPS
    r[1] += e.cpu_time_total / STEPS
    r[2] += 1 / STEPS
tot_d = sum(r[0] for r in rows.values())
tot_c = sum(r[1] for r in rows.values())
print(f"top-level aten ops: device {tot_d:.1f} us / step, host {tot_c:.1f} us / step, {sum(r[2] for r in rows.values()):.0f} ops / step")
for (k, site), (d, c, n) in sorted(rows.items(), key=lambda kv: -kv[1][1])[:90]:
    print(f"dev {d:7.1f} us  cpu {c:7.1f} us  x{n:5.1f}  {k:24s} {site}")
