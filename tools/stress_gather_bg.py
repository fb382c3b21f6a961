import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np, scipy.sparse as sp, torch
from nichecompass_b200 import ops
from nichecompass_b200.data import CSRFeatureStore
DEV = "cuda"
rng = np.random.default_rng(0)
bad = 0
side = torch.cuda.Stream()
for trial in range(120):
    n, F = int(rng.integers(50, 4000)), int(rng.choice([77, 40, 500, 2000, 5000]))
    X = sp.random(n, F, density=float(rng.choice([0.02, 0.06, 0.3])), format="csr", random_state=int(trial), dtype=np.float32)
    X.data = np.ceil(X.data * 20).astype(np.float32)
    loc = "host" if trial % 2 else "device"
    store = CSRFeatureStore.from_scipy(X, DEV, location=loc)
    dstore = CSRFeatureStore.from_scipy(X, DEV)
    m = int(rng.integers(1, 6000))
    n_id = torch.from_numpy(rng.integers(0, n, m).astype(np.int32)).to(DEV)
    for dtype in (torch.float32, torch.bfloat16):
        ld = ops.pad8(F)
        ref = ops.csr_gather_rows(dstore, n_id, m, True, dtype, ld)
        if trial % 3 == 0:
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                out = ops.csr_gather_rows(store, n_id, m, True, dtype, ld, background=True)
            a = torch.randn(4096, 4096, device=DEV) @ torch.randn(4096, 4096, device=DEV)
            torch.cuda.current_stream().wait_stream(side)
        else:
            out = ops.csr_gather_rows(store, n_id, m, True, dtype, ld, background=True)
        if not torch.equal(out, ref):
            bad += 1
            d = (out.float() - ref.float()).abs()
            print("MISMATCH", trial, loc, dtype, n, F, m, int((d > 0).sum()), float(d.max()), flush=True)
print("trials done, mismatches:", bad)
