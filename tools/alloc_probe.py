"""Host time of torch.empty / torch.zeros inside C2 training steps (is the caching allocator calling cudaMalloc
in steady state?).  usage: alloc_probe.py [cells] [steps]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
import bench                                    # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 60
sys.argv = [sys.argv[0], "--cells", str(cells)]
args = bench.parse_args()
torch.cuda.set_device(0)
wl = dict(bench.WORKLOADS["c2"])
wl["n_cells"] = cells
adata = bench.build_adata(wl, 0)
model, tr = bench.make_trainer(adata, wl, args)
it = bench.batch_stream(tr)
bench.run_steps(tr, it, 20)
torch.cuda.synchronize()
log = []
real_empty, real_zeros = torch.empty, torch.zeros


def wrap(fn, name):
    def f(*a, **k):
        t0 = time.perf_counter()
        r = fn(*a, **k)
        dt = time.perf_counter() - t0
        log.append((dt, name, r.numel() * r.element_size(), bool(k.get("pin_memory", False)), str(r.device)))
        return r
    return f


torch.empty, torch.zeros = wrap(real_empty, "empty"), wrap(real_zeros, "zeros")
s0 = torch.cuda.memory_stats()
t0 = time.perf_counter()
bench.run_steps(tr, it, STEPS)
host = time.perf_counter() - t0
torch.cuda.synchronize()
s1 = torch.cuda.memory_stats()
torch.empty, torch.zeros = real_empty, real_zeros
tot = sum(l[0] for l in log)
print(f"{len(log) / STEPS:.1f} allocations per step, {tot / STEPS * 1e3:.3f} ms per step inside them "
      f"(host enqueue {host / STEPS * 1e3:.3f} ms per step)")
print("cudaMalloc calls during the probe:", s1["num_alloc_retries"] - s0["num_alloc_retries"], "retries,",
      s1["segment.all.allocated"] - s0["segment.all.allocated"], "new segments,",
      s1["segment.all.freed"] - s0["segment.all.freed"], "freed segments; reserved",
      s1["reserved_bytes.all.current"] / 2**30, "GiB")
log.sort(reverse=True)
print("slowest:")
for dt, name, nbytes, pin, dev in log[:25]:
    print(f"  {dt * 1e6:9.1f} us  {name:5s} {nbytes / 2**20:10.3f} MiB pin={pin} {dev}")
buckets = {}
for dt, name, nbytes, pin, dev in log:
    k = ("pinned" if pin else dev.split(":")[0], "<1MiB" if nbytes < 2**20 else "<32MiB" if nbytes < 2**25 else ">=32MiB")
    b = buckets.setdefault(k, [0, 0.0])
    b[0] += 1
    b[1] += dt
for k, (n, t) in sorted(buckets.items()):
    print(f"  {k}: {n / STEPS:.1f} per step, {t / n * 1e6:.1f} us each, {t / STEPS * 1e3:.3f} ms per step")
