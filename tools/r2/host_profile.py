"""Host-side profile (cProfile) of the training loop in CUDA-graph mode on the C2 workload."""
import cProfile, io, os, pstats, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
import bench

sys.argv = [sys.argv[0]] + sys.argv[1:]
args = bench.parse_args()
wl = dict(bench.WORKLOADS[args.workload])
if args.cells:
    wl["n_cells"] = args.cells
adata = bench.build_adata(wl, 0)
model, tr = bench.make_trainer(adata, wl, args)
it = bench.batch_stream(tr)
bench.run_steps(tr, it, 60)
torch.cuda.synchronize()
t0 = time.perf_counter()
bench.run_steps(tr, it, 100)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"enqueue {1e3 * (t1 - t0) / 100:.3f} ms/step, with final sync {1e3 * (t2 - t0) / 100:.3f} ms/step, graphs {tr.graph_stats_}")
pr = cProfile.Profile()
pr.enable()
bench.run_steps(tr, it, 100)
pr.disable()
torch.cuda.synchronize()
s = io.StringIO()
pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
print("\n".join(l[:150] for l in s.getvalue().splitlines()[:80]))
