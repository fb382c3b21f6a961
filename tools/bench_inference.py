"""Throughput of full-graph latent inference (NicheCompass.get_latent_representation) on the C2 model
shape: layer-wise tiled path (models/tiled_inference.py) against the reference-shaped loader walk.
Device time by CUDA events around the encoder passes (graph / partition plumbing excluded and reported
separately as wall time).  Under torchrun the nodes are split over the ranks (spatial tiles + halo exchange).
usage: [torchrun --nproc-per-node N] tools/bench_inference.py [cells] [--walk]"""
import contextlib
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                               # noqa: E402
import torch                                     # noqa: E402
import torch.distributed as dist                 # noqa: E402

import bench                                     # noqa: E402
from nichecompass_b200.models import tiled_inference as TI     # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dev = torch.device("cuda", torch.cuda.current_device())
    cells = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 200_000
    walk = "--walk" in sys.argv
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    sys.argv = [sys.argv[0], "--cells", str(cells)]
    args = bench.parse_args()
    wl = dict(bench.WORKLOADS["c2"])
    wl["n_cells"] = cells
    adata = bench.build_adata(wl, 0)
    model, tr = bench.make_trainer(adata, wl, args)          # C2 module (bf16), count store + graph in HBM
    module = tr.model
    module.eval()
    data = tr.node_masked_data
    n = adata.shape[0]
    t0 = time.time()
    coords = adata.obsm["spatial"] if "spatial" in adata.obsm else None
    perm_np = TI.partition_order(n, coords, adata.obsp["spatial_connectivities"], "auto" if world > 1 else "none")
    bounds = TI.tile_bounds(n, world)
    perm = torch.from_numpy(perm_np).to(dev)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(n, device=dev)
    g = data.device_graph()
    lo, hi = bounds[rank], bounds[rank + 1]
    lg = TI.build_local_graph(g.rowptr, g.col, perm, inv, lo, hi)
    torch.cuda.synchronize()
    t_plumb = time.time() - t0
    res = {}
    for want_std in (False, True):
        for rep in range(3):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ml = TI.encode_partition(module, data.store, lg, perm[lo:hi], bounds, data.cats_device(), want_logstd=want_std)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res["mu_std" if want_std else "mu"] = float(t.item())
        del ml
    out = {"cells": n, "n_gpus": world, "n_own": lg.n_own, "n_halo": lg.n_halo, "dtype": "bf16",
           "graph_plumbing_s": round(t_plumb, 3),
           "layerwise_mu_ms": res["mu"], "layerwise_mu_cells_per_s": n / (res["mu"] * 1e-3),
           "layerwise_mu_std_ms": res["mu_std"], "layerwise_mu_std_cells_per_s": n / (res["mu_std"] * 1e-3)}
    if walk and world == 1:
        from nichecompass_b200.data import initialize_dataloaders
        loader = initialize_dataloaders(node_masked_data=data, edge_train_data=None, edge_val_data=None,
                                        edge_batch_size=None, node_batch_size=8192, shuffle=False)["node_train_loader"]
        n_walk = min(n, 8192 * 6)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        done = 0
        with torch.no_grad():
            for b in loader:
                module.get_latent_representation(node_batch=b, only_active_gps=False, return_mu_std=False)
                done += b.batch_size
                if done >= n_walk:
                    break
        e1.record()
        torch.cuda.synchronize()
        out["loader_walk_cells_per_s"] = done / (e0.elapsed_time(e1) * 1e-3)
        out["loader_walk_sample"] = f"{done} cells in tiles of 8192 (mu | logstd of every sampled sub-graph)"
    if rank == 0:
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
