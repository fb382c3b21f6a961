"""One training step of the C2 model (reduced cell count, same per-step shapes) bracketed by
cudaProfilerStart/Stop, for `ncu --profile-from-start off`.  usage: profile_step.py [cells] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
import bench                                    # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
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
torch.cuda.profiler.start()
bench.run_steps(tr, it, steps)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("profiled", steps, "step(s)")
