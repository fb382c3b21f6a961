"""
Deterministic synthetic spatial-omics data of the shapes BASELINE.json names.

Coordinates: uniform points (per-sample unit squares; block-diagonal graph across samples).
Graph: exact kNN (self excluded) symmetrised with ``A.maximum(A.T)`` as in the reference
tutorials.  Counts: multinomial draws of a log-normal library size over the softmax profile of
one of 8 spatial clusters (sparse, float32, every cell >= 1 count).  Gene programs: random
target / source gene sets per program, written with the key schema of the reference's
``add_gps_from_gp_dict_to_adata`` (utils/gene_programs.py:252-273): ``varm`` int masks
[n_genes, n_gps], ``uns`` gp names and (target|source|all) gene index lists where the combined
index is targets ++ (sources + n_vars).
"""
from typing import Optional

import numpy as np
import pandas as pd
import scipy.sparse as sp
from scipy.spatial import cKDTree

from .anndata_lite import AnnDataLite


def knn_adjacency(coords: np.ndarray, k: int) -> sp.csr_matrix:
    n = coords.shape[0]
    _, nbr = cKDTree(coords).query(coords, k=k + 1, workers=-1)
    rows = np.repeat(np.arange(n), k)
    cols = nbr[:, 1:].reshape(-1)
    a = sp.csr_matrix((np.ones(rows.size, dtype=np.float32), (rows, cols)), shape=(n, n))
    a = a.maximum(a.T).tocsr()
    a.sort_indices()
    return a


def _multinomial_counts(rng, n_cells, n_genes, cluster, profiles_cdf, lib):
    """CSR counts: cell i draws lib[i] genes from categorical(profile[cluster[i]])."""
    total = int(lib.sum())
    cell_of_draw = np.repeat(np.arange(n_cells, dtype=np.int64), lib)
    u = rng.random(total)
    gene = np.empty(total, dtype=np.int64)
    cl = cluster[cell_of_draw]
    for c in range(profiles_cdf.shape[0]):
        m = cl == c
        gene[m] = np.searchsorted(profiles_cdf[c], u[m], side="right")
    np.minimum(gene, n_genes - 1, out=gene)
    key = cell_of_draw * n_genes + gene
    uk, cnt = np.unique(key, return_counts=True)
    rows, cols = uk // n_genes, uk % n_genes
    x = sp.csr_matrix((cnt.astype(np.float32), (rows, cols)), shape=(n_cells, n_genes))
    x.sort_indices()
    return x


def random_gp_masks(rng, n_genes: int, n_gps: int, max_targets=50, max_sources=5,
                    min_targets=5):
    tm = np.zeros((n_genes, n_gps), dtype=np.int32)
    sm = np.zeros((n_genes, n_gps), dtype=np.int32)
    hi_t = min(max_targets, n_genes)
    lo_t = min(min_targets, hi_t)
    hi_s = min(max_sources, n_genes)
    for g in range(n_gps):
        nt = rng.integers(lo_t, hi_t + 1)
        ns = rng.integers(1, hi_s + 1)
        tm[rng.choice(n_genes, nt, replace=False), g] = 1
        sm[rng.choice(n_genes, ns, replace=False), g] = 1
    return tm, sm


def add_gp_masks(adata: AnnDataLite, targets_mask: np.ndarray, sources_mask: np.ndarray,
                 gp_targets_mask_key="nichecompass_gp_targets",
                 gp_sources_mask_key="nichecompass_gp_sources",
                 gp_names_key="nichecompass_gp_names",
                 genes_idx_key="nichecompass_genes_idx",
                 target_genes_idx_key="nichecompass_target_genes_idx",
                 source_genes_idx_key="nichecompass_source_genes_idx"):
    n_vars = adata.n_vars
    adata.varm[gp_targets_mask_key] = targets_mask
    adata.varm[gp_sources_mask_key] = sources_mask
    adata.uns[gp_names_key] = np.array([f"GP_{i}" for i in range(targets_mask.shape[1])])
    t_idx = np.nonzero(targets_mask.sum(1))[0]
    s_idx = np.nonzero(sources_mask.sum(1))[0]
    adata.uns[target_genes_idx_key] = t_idx
    adata.uns[source_genes_idx_key] = s_idx
    adata.uns[genes_idx_key] = np.concatenate((t_idx, s_idx + n_vars), axis=0)


def make_synthetic_adata(n_cells: int, n_genes: int, k: int, n_prior_gp: int, seed: int = 0,
                         n_samples: int = 1, lib_mean: float = 2000., lib_sigma: float = 0.5,
                         n_clusters: int = 8, counts_key: Optional[str] = "counts",
                         adj_key: str = "spatial_connectivities",
                         sample_key: str = "batch") -> AnnDataLite:
    rng = np.random.default_rng(seed)
    per = n_cells // n_samples
    sizes = [per] * n_samples
    sizes[-1] += n_cells - per * n_samples
    coords, sample = [], []
    blocks = []
    for s, ns in enumerate(sizes):
        c = rng.random((ns, 2))
        coords.append(c + np.array([2.0 * s, 0.0]))
        sample.append(np.full(ns, s))
        blocks.append(knn_adjacency(c, k))
    coords = np.concatenate(coords)
    adj = sp.block_diag(blocks, format="csr").astype(np.float32)
    adj.sort_indices()
    # 8 spatial clusters: 4 x 2 grid cells of each unit square
    local = coords - np.floor(coords / 2.0) * 2.0
    cluster = (np.minimum((local[:, 0] * 4).astype(int), 3)
               + 4 * np.minimum((local[:, 1] * 2).astype(int), 1)) % n_clusters
    logits = rng.normal(0., 1.5, size=(n_clusters, n_genes))
    prof = np.exp(logits - logits.max(1, keepdims=True))
    prof /= prof.sum(1, keepdims=True)
    cdf = np.cumsum(prof, axis=1)
    lib = np.maximum(rng.lognormal(np.log(lib_mean), lib_sigma, n_cells).astype(np.int64), 1)
    x = _multinomial_counts(rng, n_cells, n_genes, cluster, cdf, lib)
    obs = pd.DataFrame({sample_key: [f"sample{s}" for s in np.concatenate(sample)]},
                       index=[f"cell{i}" for i in range(n_cells)])
    adata = AnnDataLite(x, obs=obs)
    if counts_key is not None:
        adata.layers[counts_key] = x
    adata.obsp[adj_key] = adj
    adata.obsm["spatial"] = coords
    tm, sm = random_gp_masks(rng, n_genes, n_prior_gp)
    add_gp_masks(adata, tm, sm)
    return adata


def _poisson_counts_chunked(rng, n_cells, n_genes, cluster, prof, lib, device=None, chunk=16384):
    """CSR counts with x[i, g] ~ Poisson(lib[i] * prof[cluster[i], g]) drawn chunk by chunk (on the
    GPU when one is visible: 1M x 5k takes seconds), every cell guaranteed >= 1 count."""
    import torch
    dev = torch.device(device) if device is not None else torch.device(
        "cuda" if torch.cuda.is_available() else "cpu")
    g = torch.Generator(device=dev)
    g.manual_seed(int(rng.integers(0, 2 ** 31 - 1)))
    prof_t = torch.from_numpy(prof.astype(np.float32)).to(dev)
    indptr = [np.zeros(1, dtype=np.int64)]
    indices, data = [], []
    base = 0
    for lo in range(0, n_cells, chunk):
        hi = min(lo + chunk, n_cells)
        cl = torch.from_numpy(cluster[lo:hi]).to(dev)
        lb = torch.from_numpy(lib[lo:hi].astype(np.float32)).to(dev)
        rate = prof_t[cl] * lb[:, None]
        x = torch.poisson(rate, generator=g)
        empty = x.sum(1) == 0
        if bool(empty.any()):
            rows = empty.nonzero().flatten()
            x[rows, (rows + lo) % n_genes] = 1.
        nz = x.nonzero()
        vals = x[nz[:, 0], nz[:, 1]]
        counts = torch.bincount(nz[:, 0], minlength=hi - lo)
        ptr = torch.cumsum(counts, 0).cpu().numpy().astype(np.int64) + base
        base = int(ptr[-1])
        indptr.append(ptr)
        indices.append(nz[:, 1].to(torch.int32).cpu().numpy())
        data.append(vals.cpu().numpy().astype(np.float32))
    return sp.csr_matrix((np.concatenate(data), np.concatenate(indices), np.concatenate(indptr)),
                         shape=(n_cells, n_genes))


def make_synthetic_adata_large(n_cells: int, n_genes: int, k: int, n_prior_gp: int, seed: int = 0,
                               n_samples: int = 1, lib_mean: float = 300., lib_sigma: float = 0.5,
                               n_clusters: int = 8, counts_key: Optional[str] = "counts",
                               adj_key: str = "spatial_connectivities", sample_key: str = "batch",
                               device=None) -> AnnDataLite:
    """Same construction as ``make_synthetic_adata`` at atlas scale (BASELINE.json configs 2-5):
    Poisson counts around lib_i * profile[cluster_i] instead of explicit multinomial draws, obs
    frame built without per-cell Python work."""
    rng = np.random.default_rng(seed)
    per = n_cells // n_samples
    sizes = [per] * n_samples
    sizes[-1] += n_cells - per * n_samples
    coords, sample, blocks = [], [], []
    for s, ns in enumerate(sizes):
        c = rng.random((ns, 2))
        coords.append(c + np.array([2.0 * s, 0.0]))
        sample.append(np.full(ns, s))
        blocks.append(knn_adjacency(c, k))
    coords = np.concatenate(coords)
    adj = (sp.block_diag(blocks, format="csr") if n_samples > 1 else blocks[0]).astype(np.float32)
    adj.sort_indices()
    local = coords - np.floor(coords / 2.0) * 2.0
    cluster = ((np.minimum((local[:, 0] * 4).astype(int), 3)
                + 4 * np.minimum((local[:, 1] * 2).astype(int), 1)) % n_clusters).astype(np.int64)
    logits = rng.normal(0., 1.5, size=(n_clusters, n_genes))
    prof = np.exp(logits - logits.max(1, keepdims=True))
    prof /= prof.sum(1, keepdims=True)
    lib = np.maximum(rng.lognormal(np.log(lib_mean), lib_sigma, n_cells), 1.0)
    x = _poisson_counts_chunked(rng, n_cells, n_genes, cluster, prof, lib, device)
    sample_ids = np.concatenate(sample)
    obs = pd.DataFrame({sample_key: pd.Categorical.from_codes(sample_ids, [f"sample{s}" for s in range(n_samples)])},
                       index=pd.RangeIndex(n_cells).astype(str))
    adata = AnnDataLite(x, obs=obs)
    if counts_key is not None:
        adata.layers[counts_key] = x
    adata.obsp[adj_key] = adj
    adata.obsm["spatial"] = coords
    tm, sm = random_gp_masks(rng, n_genes, n_prior_gp)
    add_gp_masks(adata, tm, sm)
    return adata


def add_synthetic_atac(adata: AnnDataLite, n_peaks: int, seed: int = 0, lib_mean: float = 600., lib_sigma: float = 0.5,
                       n_clusters: int = 8, extra_links: float = 0.2, device=None) -> AnnDataLite:
    """Chromatin-accessibility modality for a synthetic ``adata`` (BASELINE.json config 4: RNA + ATAC): returns
    ``adata_atac`` [n_obs, n_peaks] and writes the gene-peak mapping into ``adata.varm['nichecompass_gene_peaks']``.
    Every peak is linked to one random gene (+ ``extra_links`` more on average); the peak masks of the gene programs
    follow from the gene masks through that mapping, as the reference derives them
    (utils/multimodal_mapping.py -> ``adata_atac.varm['nichecompass_ca_targets' / '_sources']``, consumed at
    models/nichecompass.py:404-447)."""
    rng = np.random.default_rng(seed + 7)
    n_cells, n_genes = adata.shape
    n_links = n_peaks + int(extra_links * n_peaks)
    rows = np.concatenate((rng.integers(0, n_genes, n_peaks), rng.integers(0, n_genes, n_links - n_peaks)))
    cols = np.concatenate((np.arange(n_peaks), rng.integers(0, n_peaks, n_links - n_peaks)))
    gene_peaks = sp.csr_matrix((np.ones(n_links, dtype=np.float32), (rows, cols)), shape=(n_genes, n_peaks))
    gene_peaks.data[:] = 1.
    adata.varm["nichecompass_gene_peaks"] = gene_peaks.astype(bool)
    coords = adata.obsm["spatial"]
    local = coords - np.floor(coords / 2.0) * 2.0
    cluster = ((np.minimum((local[:, 0] * 4).astype(int), 3)
                + 4 * np.minimum((local[:, 1] * 2).astype(int), 1)) % n_clusters).astype(np.int64)
    logits = rng.normal(0., 1.5, size=(n_clusters, n_peaks))
    prof = np.exp(logits - logits.max(1, keepdims=True))
    prof /= prof.sum(1, keepdims=True)
    lib = np.maximum(rng.lognormal(np.log(lib_mean), lib_sigma, n_cells), 1.0)
    x = _poisson_counts_chunked(rng, n_cells, n_peaks, cluster, prof, lib, device, chunk=4096)
    atac = AnnDataLite(x, obs=adata.obs.copy(),
                       var=pd.DataFrame(index=[f"peak{i}" for i in range(n_peaks)]))
    atac.layers["counts"] = x
    gpt = gene_peaks.T.tocsr().astype(np.float32)                                  # [n_peaks, n_genes]
    for name, key in (("targets", "nichecompass_gp_targets"), ("sources", "nichecompass_gp_sources")):
        gene_mask = sp.csr_matrix(np.asarray(adata.varm[key], dtype=np.float32))   # [n_genes, P]
        atac.varm[f"nichecompass_ca_{name}"] = (gpt @ gene_mask).astype(bool).tocsr()   # [n_peaks, P]
    t_idx = np.nonzero(np.asarray(atac.varm["nichecompass_ca_targets"].sum(1)).reshape(-1))[0]
    s_idx = np.nonzero(np.asarray(atac.varm["nichecompass_ca_sources"].sum(1)).reshape(-1))[0]
    atac.uns["nichecompass_target_peaks_idx"] = t_idx
    atac.uns["nichecompass_source_peaks_idx"] = s_idx
    atac.uns["nichecompass_peaks_idx"] = np.concatenate((t_idx, s_idx + n_peaks))
    return atac
