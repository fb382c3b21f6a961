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
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    bench.run_steps(tr, it, STEPS)
    torch.cuda.synchronize()
rows = []
for e in prof.key_averages(group_by_stack_n=4):
    dt = getattr(e, "self_device_time_total", 0) or getattr(e, "self_cuda_time_total", 0)
    if e.key.startswith("aten::") and dt > 0:
        rows.append((dt / STEPS, e.count / STEPS, e.key, [s for s in e.stack if "nichecompass_b200" in s or "bench.py" in s][:2]))
rows.sort(reverse=True)
tot = sum(r[0] for r in rows)
print(f"aten ops with device time: {tot:.1f} us / step")
for dt, n, k, st in rows[:45]:
    print(f"{dt:8.1f} us  x{n:5.1f}  {k:28s} {' | '.join(s.split('/')[-1] for s in st)}")
