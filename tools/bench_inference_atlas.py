"""
BASELINE.json config 5: full-graph latent inference of an atlas-scale sample (default 8.5 M cells x 5 k genes, kNN
k = 8, C2 model shape, bf16) over spatial tiles, one tile per GPU, halo rows read straight from the owners' HBM.

    torchrun --nproc-per-node 8 tools/bench_inference_atlas.py [--cells 8500000] [--halo peer|nccl] [--check]

Rank 0 draws the dataset once and publishes the raw arrays under /dev/shm; every rank maps them (no per-rank copy
of the 18 GB count matrix in host memory), uploads the CSR count store + graph and runs
``models/tiled_inference.encode_partition`` on its tile.  Timed with CUDA events around the encoder passes, max over
ranks; the one-off graph plumbing is reported separately.  ``--check``: rows of a sample of this rank's nodes are
recomputed from scratch through the reference-shaped loader path (sampled 1-hop sub-graph, same kernels) and compared.
"""
import argparse
import contextlib
import io
import json
import os
import pickle
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                               # noqa: E402
import scipy.sparse as sp                        # noqa: E402
import torch                                     # noqa: E402
import torch.distributed as dist                 # noqa: E402

from nichecompass_b200.data.store import CSRFeatureStore          # noqa: E402
from nichecompass_b200.models import tiled_inference as TI         # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cells", type=int, default=8_500_000)
    ap.add_argument("--genes", type=int, default=5000)
    ap.add_argument("--halo", default="peer", choices=["peer", "nccl"])
    ap.add_argument("--check", action="store_true")
    a = ap.parse_args()
    os.environ["NC_HALO_EXCHANGE"] = a.halo
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
    dev = torch.device("cuda", torch.cuda.current_device())
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    base = f"/dev/shm/nc_atlas_{os.getuid()}_{a.cells}_{a.genes}"
    t0 = time.time()
    if rank == 0:
        from nichecompass_b200.data import make_synthetic_adata_large
        adata = make_synthetic_adata_large(a.cells, a.genes, 8, 1500, seed=0, lib_mean=300.)
        x = sp.csr_matrix(adata.X)
        adj = sp.csr_matrix(adata.obsp["spatial_connectivities"])
        for name, arr in (("x_indptr", x.indptr.astype(np.int64)), ("x_indices", x.indices.astype(np.int32)),
                          ("x_data", x.data.astype(np.float32)), ("a_indptr", adj.indptr.astype(np.int64)),
                          ("a_indices", adj.indices.astype(np.int32)),
                          ("coords", np.asarray(adata.obsm["spatial"], dtype=np.float64))):
            np.save(f"{base}_{name}.npy", arr)
        with open(f"{base}_meta.pkl", "wb") as fh:
            pickle.dump({"varm": adata.varm, "uns": adata.uns}, fh)
        del adata, x, adj
    if world > 1:
        dist.barrier()
    t_gen = time.time() - t0
    ld = lambda name: np.load(f"{base}_{name}.npy", mmap_mode="r")              # noqa: E731
    n = a.cells
    t0 = time.time()
    store = CSRFeatureStore(torch.from_numpy(np.asarray(ld("x_indptr"))).to(dev),
                            torch.from_numpy(np.asarray(ld("x_indices"))).to(dev),
                            torch.from_numpy(np.asarray(ld("x_data"))).to(dev), a.genes)
    # the kNN-union adjacency is symmetric: its CSR is the CSC the encoder needs (self entries are replaced by the
    # attention kernel anyway)
    g_rowptr = torch.from_numpy(np.asarray(ld("a_indptr"))).to(dev)
    g_col = torch.from_numpy(np.asarray(ld("a_indices"))).to(dev)
    coords = np.asarray(ld("coords"))
    with open(f"{base}_meta.pkl", "rb") as fh:
        meta = pickle.load(fh)
    torch.cuda.synchronize()
    t_upload = time.time() - t0
    # ---- the C2 module (random init, same seed on every rank), bf16
    from nichecompass_b200.modules import VGPGAE
    tm = torch.tensor(np.asarray(meta["varm"]["nichecompass_gp_targets"]).T, dtype=torch.bool)
    sm = torch.tensor(np.asarray(meta["varm"]["nichecompass_gp_sources"]).T, dtype=torch.bool)
    fd = {"masked_rna_idx": list(meta["uns"]["nichecompass_genes_idx"])}
    for ent in ("target", "source"):
        masked = [int(i) for i in meta["uns"][f"nichecompass_{ent}_genes_idx"]]
        keep = np.ones(a.genes, dtype=bool)
        keep[masked] = False
        fd[f"{ent}_masked_rna_idx"] = masked
        fd[f"{ent}_unmasked_rna_idx"] = np.nonzero(keep)[0].tolist()
    torch.manual_seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        module = VGPGAE(n_input=a.genes, n_fc_layers_encoder=1, n_layers_encoder=1, n_hidden_encoder=1600,
                        n_prior_gp=1500, n_addon_gp=100, cat_covariates_embeds_nums=[], n_output_genes=a.genes,
                        target_rna_decoder_mask=tm, source_rna_decoder_mask=sm, features_idx_dict=fd,
                        features_scale_factors=torch.ones(2 * a.genes), conv_layer_encoder="gatv2conv",
                        encoder_n_attention_heads=4, node_label_method="one-hop-norm",
                        include_cat_covariates_contrastive_loss=False).to(dev)
    module.set_compute_dtype(torch.bfloat16)
    module.eval()
    # ---- partition + rank-local graph
    t0 = time.time()
    perm_np = TI.partition_order(n, coords, None, "spatial" if world > 1 else "none")
    bounds = TI.tile_bounds(n, world)
    perm = torch.from_numpy(perm_np).to(dev)
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(n, device=dev)
    lo, hi = bounds[rank], bounds[rank + 1]
    lg = TI.build_local_graph(g_rowptr, g_col, perm, inv, lo, hi)
    torch.cuda.synchronize()
    t_plumb = time.time() - t0
    res = {}
    ml = None
    for want_std in (False, True):
        for rep in range(2):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            ml = TI.encode_partition(module, store, lg, perm[lo:hi], bounds, None, want_logstd=want_std)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
        t = torch.tensor([ms], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        res["mu_std" if want_std else "mu"] = float(t.item())
    halo = torch.tensor([lg.n_halo, lg.n_own], device=dev, dtype=torch.int64)
    halos = [torch.zeros_like(halo) for _ in range(world)]
    if world > 1:
        dist.all_gather(halos, halo)
    else:
        halos = [halo]
    out = {"config": "C5 full-graph latent inference", "cells": n, "genes": a.genes, "n_gpus": world, "dtype": "bf16",
           "halo_exchange": a.halo if world > 1 else "none",
           "n_own_per_rank": [int(h[1]) for h in halos], "n_halo_per_rank": [int(h[0]) for h in halos],
           "dataset_generation_s": round(t_gen, 1), "upload_s": round(t_upload, 1), "graph_plumbing_s": round(t_plumb, 2),
           "mu_ms": res["mu"], "mu_cells_per_s": n / (res["mu"] * 1e-3),
           "mu_std_ms": res["mu_std"], "mu_std_cells_per_s": n / (res["mu_std"] * 1e-3),
           "peak_hbm_gb": round(torch.cuda.max_memory_allocated() / 2 ** 30, 1)}
    if a.check:
        # a sample of this rank's nodes through the loader-shaped path: 1-hop sub-graph of the sample, same kernels
        from nichecompass_b200.data.sampler import DeviceGraph  # noqa: F401  (documentation of the equivalent path)
        from nichecompass_b200.functional import SubgraphCSR
        P = module.n_prior_gp_ + module.n_addon_gp_
        pick = torch.arange(lo, hi, max((hi - lo) // 512, 1), device=dev)[:512]
        seeds = perm[pick]                                                   # original ids
        rp = g_rowptr
        deg = (rp[seeds + 1] - rp[seeds])
        starts = rp[seeds]
        within = torch.arange(int(deg.sum()), device=dev) - torch.repeat_interleave(torch.cumsum(deg, 0) - deg, deg)
        nbr = g_col[(torch.repeat_interleave(starts, deg) + within)].long()
        ids, inv_ids = torch.unique(torch.cat((seeds, nbr)), return_inverse=True)
        # local order: seeds first
        order = torch.cat((inv_ids[:seeds.numel()], torch.tensor([], dtype=torch.long, device=dev)))
        is_seed = torch.zeros(ids.numel(), dtype=torch.bool, device=dev)
        is_seed[order] = True
        rest = (~is_seed).nonzero().flatten()
        local_of = torch.empty(ids.numel(), dtype=torch.long, device=dev)
        local_of[order] = torch.arange(seeds.numel(), device=dev)
        local_of[rest] = torch.arange(rest.numel(), device=dev) + seeds.numel()
        n_id = torch.empty(ids.numel(), dtype=torch.int32, device=dev)
        n_id[local_of] = ids.to(torch.int32)
        rowptr = torch.zeros(seeds.numel() + 1, dtype=torch.int32, device=dev)
        rowptr[1:] = torch.cumsum(deg, 0)
        col = local_of[inv_ids[seeds.numel():]].to(torch.int32).contiguous()
        gsub = SubgraphCSR(rowptr, col, ids.numel(), seeds.numel(), int(deg.max()))
        batch = type("B", (), {})()
        batch.store, batch.n_id, batch.graph, batch.batch_size = store, n_id, gsub, seeds.numel()
        batch.cat_covariates_cats = None
        with torch.no_grad():
            mu_ref, std_ref = module.get_latent_representation(batch, only_active_gps=False, return_mu_std=True)
        mine = ml[pick - lo]
        err_mu = float((mine[:, :P] - mu_ref).abs().max() / mu_ref.abs().max())
        err_std = float((torch.exp(mine[:, P:]) - std_ref).abs().max() / std_ref.abs().max())
        errs = torch.tensor([err_mu, err_std], device=dev)
        if world > 1:
            dist.all_reduce(errs, op=dist.ReduceOp.MAX)
        out["check_vs_loader_path"] = {"rows_per_rank": int(pick.numel()), "max_rel_err_mu": float(errs[0]),
                                       "max_rel_err_std": float(errs[1])}
    if world > 1:
        dist.barrier()
    if rank == 0:
        for f in os.listdir("/dev/shm"):
            if f.startswith(os.path.basename(base)):
                with contextlib.suppress(OSError):
                    os.remove(os.path.join("/dev/shm", f))
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
