"""torch.profiler view of training iterations of a bench workload (which torch ops / library kernels take the time)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from torch.profiler import ProfilerActivity, profile
import bench

args = bench.parse_args()
wl = dict(bench.WORKLOADS[args.workload])
if args.cells:
    wl["n_cells"] = args.cells
adata = bench.build_adata(wl, 0)
model, tr = bench.make_trainer(adata, wl, args)
it = bench.batch_stream(tr)
bench.run_steps(tr, it, 8)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    bench.run_steps(tr, it, 5)
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=22, max_name_column_width=60))
print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=18, max_name_column_width=60))
