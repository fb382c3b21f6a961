"""Where does the end-to-end leg of bench.py lose time against the HBM-resident leg?  Times the C2 step with
the count store on the device / in pinned host memory x loss not read / read one step late / read blocking.
usage: e2e_breakdown.py [cells] [steps]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch                                    # noqa: E402
import bench                                    # noqa: E402
from nichecompass_b200.data import CSRFeatureStore   # noqa: E402

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
STEPS = int(sys.argv[2]) if len(sys.argv) > 2 else 100
sys.argv = [sys.argv[0], "--cells", str(cells)]
args = bench.parse_args()
torch.cuda.set_device(0)
wl = dict(bench.WORKLOADS["c2"])
wl["n_cells"] = cells
adata = bench.build_adata(wl, 0)
model, tr = bench.make_trainer(adata, wl, args)
ds = tr.node_masked_data.dataset
dev_store = ds._store
host_store = CSRFeatureStore.from_scipy(ds._x_rna, ds.device, ds._x_atac, location="host")
out = {}
for loc, store in (("device", dev_store), ("host", host_store)):
    ds._store = store
    for stage in ((None,) if loc == "device" else (False, True)):
        tr.stage_inputs_ = stage
        it = bench.batch_stream(tr)
        bench.run_steps(tr, it, 10)
        for read in (False, "pipelined", True):
            ms, _ = bench.timed(tr, it, STEPS, 1, read_loss=read)
            out[f"{loc}{'' if stage is None else '+staged' if stage else '+unstaged'}/read={read}"] = round(ms / STEPS, 4)
from nichecompass_b200 import ops               # noqa: E402
ds._store = host_store
tr.stage_inputs_ = True
for ctas in (74, 148, 296, 592):
    for lab in (True, False):
        ops.BG_MAX_CTAS = ctas
        type(tr.model).stage_labels = lab
        it = bench.batch_stream(tr)
        bench.run_steps(tr, it, 10)
        ms, _ = bench.timed(tr, it, STEPS, 1, read_loss="pipelined")
        out[f"host+staged/ctas={ctas}/labels_staged={lab}"] = round(ms / STEPS, 4)
print(json.dumps(out, indent=1))
