"""Shared builders for tests: tiny synthetic problems, oracle cfg, batches."""
import math
from types import SimpleNamespace

import numpy as np
import torch

from nichecompass_b200.data import make_synthetic_adata


def features_idx_dict_from_adata(adata):
    """Mirror of models/nichecompass.py:387-402 (vectorised)."""
    n = adata.n_vars
    d = {}
    d["masked_rna_idx"] = adata.uns["nichecompass_genes_idx"]
    for ent in ("target", "source"):
        masked = [int(i) for i in adata.uns[f"nichecompass_{ent}_genes_idx"]]
        ms = set(masked)
        d[f"{ent}_masked_rna_idx"] = masked
        d[f"{ent}_unmasked_rna_idx"] = [i for i in range(n) if i not in ms]
    return d


def oracle_cfg_from_adata(adata, n_addon_gp=4, conv_layer="gatv2conv", heads=4,
                          node_label_method="one-hop-norm", n_hidden=None, **extra):
    tm = torch.tensor(adata.varm["nichecompass_gp_targets"].T, dtype=torch.bool)
    sm = torch.tensor(adata.varm["nichecompass_gp_sources"].T, dtype=torch.bool)
    Pp = tm.shape[0]
    F = adata.n_vars
    if n_hidden is None:
        n_hidden = min(F, Pp + n_addon_gp)
    cfg = dict(n_input=F, n_output_genes=F, n_output_peaks=0, n_hidden=n_hidden,
               n_prior_gp=Pp, n_addon_gp=n_addon_gp, conv_layer=conv_layer, heads=heads,
               n_layers=1, node_label_method=node_label_method, log_variational=True,
               active_gp_thresh_ratio=0.01, active_gp_type="separate",
               include_edge_kl_loss=True, include_edge_recon_loss=True,
               include_gene_expr_recon_loss=True, include_chrom_access_recon_loss=True,
               include_cat_covariates_contrastive_loss=False,
               cat_covariates_embeds_injection=["gene_expr_decoder", "chrom_access_decoder"],
               cat_covariates_embeds_nums=[], cat_covariates_no_edges=[],
               target_rna_mask=tm, source_rna_mask=sm,
               features_idx_dict=features_idx_dict_from_adata(adata))
    cfg.update(extra)
    return cfg


def tiny_problem(n_cells=120, n_genes=40, k=4, n_prior_gp=10, seed=0, lib_mean=60., n_samples=1):
    return make_synthetic_adata(n_cells, n_genes, k, n_prior_gp, seed=seed, lib_mean=lib_mean,
                                n_samples=n_samples)


def dense_x(adata, dtype=torch.float32):
    return torch.tensor(np.asarray(adata.X.todense()), dtype=dtype)


def batch_to(batch, dtype=None, device=None):
    out = SimpleNamespace()
    for k, v in batch.__dict__.items():
        if torch.is_tensor(v):
            if v.is_floating_point() and dtype is not None:
                v = v.to(dtype)
            if device is not None:
                v = v.to(device)
        setattr(out, k, v)
    return out


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))
