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
