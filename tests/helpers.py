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


def multimodal_covariate_problem(n_cells=140, n_genes=36, n_peaks=52, n_prior_gp=10, n_addon_gp=3, seed=0,
                                 conv_layer="gatv2conv", node_label_method="one-hop-norm",
                                 injection=("encoder", "gene_expr_decoder", "chrom_access_decoder"),
                                 with_atac=True, with_covariates=True):
    """Tiny C3/C4-shaped problem: two samples (covariate 0, no cross-category edges) + a random
    'condition' covariate (covariate 1), RNA + ATAC counts, gene-peak mapping, peak masks derived
    from the gene masks as in models/nichecompass.py:404-447.  Returns (adata, cfg, x [N, F+Fp]
    fp64, cats long [N, 2] or None)."""
    rng = np.random.default_rng(seed + 100)
    adata = tiny_problem(n_cells=n_cells, n_genes=n_genes, n_prior_gp=n_prior_gp, seed=seed, n_samples=2)
    extra = {}
    x = dense_x(adata, torch.float64)
    if with_atac:
        gp = (rng.random((n_genes, n_peaks)) < 0.08)
        gp[rng.integers(0, n_genes, n_peaks), np.arange(n_peaks)] = True       # every peak maps to a gene
        gpm = torch.tensor(gp)
        tm = torch.tensor(adata.varm["nichecompass_gp_targets"].T, dtype=torch.int64)
        sm = torch.tensor(adata.varm["nichecompass_gp_sources"].T, dtype=torch.int64)
        tam = (tm @ gpm.to(torch.int64)) > 0                                    # [Pp, Fp]
        sam = (sm @ gpm.to(torch.int64)) > 0
        xa = torch.tensor(rng.poisson(0.6, size=(n_cells, n_peaks)), dtype=torch.float64)
        xa[:, 0] += 1.                                                          # no empty rows
        x = torch.cat((x, xa), dim=1)
        extra.update(n_input=n_genes + n_peaks, n_output_peaks=n_peaks, gene_peaks_mask=gpm,
                     target_atac_mask=tam, source_atac_mask=sam)
    cats = None
    if with_covariates:
        sample = torch.tensor([int(s[-1]) for s in adata.obs["batch"]], dtype=torch.long)
        cond = torch.tensor(rng.integers(0, 3, n_cells), dtype=torch.long)
        cats = torch.stack((sample, cond), dim=1)
        extra.update(cat_covariates_embeds_nums=[3, 2], cat_covariates_num_cats=[2, 3],
                     cat_covariates_no_edges=[True, False],
                     cat_covariates_embeds_injection=list(injection),
                     include_cat_covariates_contrastive_loss=True)
    cfg = oracle_cfg_from_adata(adata, n_addon_gp=n_addon_gp, conv_layer=conv_layer,
                                node_label_method=node_label_method,
                                n_hidden=min(n_genes, n_prior_gp + n_addon_gp), **extra)
    if with_atac:
        fd = cfg["features_idx_dict"]
        for ent, m in (("target", tam), ("source", sam)):
            masked = torch.nonzero(m.sum(0) > 0).flatten().tolist()
            ms = set(masked)
            fd[f"{ent}_masked_atac_idx"] = masked
            fd[f"{ent}_unmasked_atac_idx"] = [i for i in range(n_peaks) if i not in ms]
        both = sorted(set(fd["target_masked_atac_idx"]) | {i + n_peaks for i in fd["source_masked_atac_idx"]})
        fd["masked_atac_idx"] = both
    return adata, cfg, x, cats


def oracle_cfg_multimodal(adata, atac, n_addon_gp=3, **extra):
    """Oracle configuration of an RNA + ATAC problem built by ``add_synthetic_atac`` (the masks the reference
    reads at models/nichecompass.py:404-447, densified)."""
    import scipy.sparse as sp
    n_genes, n_peaks = adata.n_vars, atac.n_vars
    gpm = torch.from_numpy(np.asarray(sp.csr_matrix(adata.varm["nichecompass_gene_peaks"]).astype(bool).todense()))
    tam = torch.from_numpy(np.asarray(sp.csr_matrix(atac.varm["nichecompass_ca_targets"]).astype(bool).todense()).T.copy())
    sam = torch.from_numpy(np.asarray(sp.csr_matrix(atac.varm["nichecompass_ca_sources"]).astype(bool).todense()).T.copy())
    cfg = oracle_cfg_from_adata(adata, n_addon_gp=n_addon_gp, n_input=n_genes + n_peaks, n_output_peaks=n_peaks,
                                gene_peaks_mask=gpm, target_atac_mask=tam, source_atac_mask=sam, **extra)
    cfg["n_hidden"] = min(n_genes, cfg["n_prior_gp"] + n_addon_gp)
    fd = cfg["features_idx_dict"]
    keep = np.ones(n_peaks, dtype=bool)
    for ent in ("target", "source"):
        masked = [int(i) for i in atac.uns[f"nichecompass_{ent}_peaks_idx"]]
        k = keep.copy()
        k[masked] = False
        fd[f"{ent}_masked_atac_idx"] = masked
        fd[f"{ent}_unmasked_atac_idx"] = np.nonzero(k)[0].tolist()
    fd["masked_atac_idx"] = [int(i) for i in atac.uns["nichecompass_peaks_idx"]]
    return cfg
