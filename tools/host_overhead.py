"""Host-side cost of one training step: enqueue time (no sync) vs device time, + cProfile of 30 steps."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
import bench                                    # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
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
for rep in range(3):
    t0 = time.perf_counter()
    bench.run_steps(tr, it, 30)
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"rep {rep}: host enqueue {1e3 * (t1 - t0) / 30:.3f} ms/step, with final sync {1e3 * (t2 - t0) / 30:.3f} ms/step")
# pure host cost of train_iteration: batches sampled beforehand, no read-backs inside the timed region
batches = [next(it) for _ in range(40)]
torch.cuda.synchronize()
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for eb, nb in batches[rep * 10:rep * 10 + 10]:
        tr.train_iteration(eb, nb)
    e1.record()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"presampled rep {rep}: host enqueue {1e2 * (t1 - t0):.3f} ms/step, wall {1e2 * (t2 - t0):.3f} ms/step, "
          f"device {e0.elapsed_time(e1) / 10:.3f} ms/step")
pr = cProfile.Profile()
pr.enable()
bench.run_steps(tr, it, 30)
pr.disable()
torch.cuda.synchronize()
st = pstats.Stats(pr)
st.sort_stats("cumulative").print_stats(45)
st.sort_stats("tottime").print_stats(30)

# host time spent INSIDE the C-ABI calls (ctypes conversion + library host code + cudaLaunch), per entry point
from nichecompass_b200 import _lib                                   # noqa: E402


class _Hook:
    def __init__(self):
        self.t, self.n, self.t0 = {}, {}, 0.0

    def before(self, name, args):
        self.t0 = time.perf_counter()

    def after(self, name, args):
        dt = time.perf_counter() - self.t0
        self.t[name] = self.t.get(name, 0.0) + dt
        self.n[name] = self.n.get(name, 0) + 1


hk = _Hook()
_lib.set_call_hook(hk)
N = 30
t0 = time.perf_counter()
bench.run_steps(tr, it, N)
t1 = time.perf_counter()
torch.cuda.synchronize()
_lib.set_call_hook(None)
tot = sum(hk.t.values())
print(f"\nC-ABI calls: {sum(hk.n.values()) / N:.1f} per step, {1e3 * tot / N:.3f} ms/step inside the calls "
      f"(of {1e3 * (t1 - t0) / N:.3f} ms/step host enqueue)")
for k, v in sorted(hk.t.items(), key=lambda kv: -kv[1]):
    print(f"  {k:28s} x{hk.n[k] / N:5.1f}  {1e6 * v / hk.n[k]:7.1f} us/call  {1e3 * v / N:.3f} ms/step")

# where does the link loader's host time go?  (statements of LinkLoader.__iter__, timed one by one while a
# training step is in flight on the compute stream, as in the real loop)
import math                                                           # noqa: E402
from nichecompass_b200.data.sampler import sample_subgraph, unique_sorted   # noqa: E402
ld = tr.edge_train_loader
g = ld.data.device_graph()
dev = g.device
side = torch.cuda.Stream()
order = torch.randperm(ld.pos.size(1)).to(dev)
pos_dev = ld.pos.to(dev, torch.int32)
acc = {}
batches = [next(it) for _ in range(12)]
torch.cuda.synchronize()


def tick(name, t0):
    t1 = time.perf_counter()
    acc[name] = acc.get(name, 0.0) + (t1 - t0)
    return t1


for b in range(10):
    tr.train_iteration(*batches[b])                                   # keeps the compute stream busy
    with torch.cuda.stream(side):
        t = time.perf_counter()
        idx = order[b * ld.batch_size:(b + 1) * ld.batch_size]
        t = tick("slice", t)
        pos = pos_dev[:, idx]
        t = tick("index pos[:, idx]", t)
        pos2 = torch.index_select(pos_dev, 1, idx)
        t = tick("index_select", t)
        src_neg = torch.randint(g.num_nodes, (ld.batch_size,))
        dst_neg = torch.randint(g.num_nodes, (ld.batch_size,))
        neg = torch.stack((src_neg, dst_neg)).to(torch.int32).pin_memory().to(dev, non_blocking=True)
        t = tick("negatives", t)
        ids = torch.cat((pos[0], neg[0], pos[1], neg[1])).contiguous()
        t = tick("cat", t)
        seeds, inv, n_seed = unique_sorted(ids)
        t = tick("unique_sorted (sync)", t)
        n_id, graph = sample_subgraph(g, seeds.contiguous(), n_seed, ld.num_neighbors, 0, b)
        t = tick("sample_subgraph (sync)", t)
torch.cuda.synchronize()
print("\nlink loader statements, host us per batch:")
for k, v in acc.items():
    print(f"  {k:26s} {1e5 * v:8.1f}")
