"""
Generates the golden vectors under tests/golden/ from the UNMODIFIED reference sources.

Run in the builder container only (needs /root/reference):

    python tests/golden/make_golden.py

For every case it instantiates the reference ``nichecompass.modules.VGPGAE`` (loaded from
/root/reference/src/nichecompass; the missing third-party imports torch_geometric / torch_sparse /
mlflow are served by the pure-torch stand-ins in oracle/pyg_shim), runs one node batch through the
omics decoders and one edge batch through the graph decoder in fp64 with injected
reparameterisation noise, evaluates ``loss`` + ``backward``, the active gene-program mask and the
latent representation, and writes inputs (weights, batches, noise) and outputs to
``tests/golden/<case>.npz``.  tests/test_golden.py replays the inputs through the CPU oracle
(not gpu) and through the CUDA module (gpu) -- neither needs /root/reference at test time.
"""
import copy
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import sampler_ref as R                      # noqa: E402
from tests import helpers as H                           # noqa: E402
from tests.ref_harness import load_reference, quiet      # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
LAM = dict(lambda_l1_masked=3., lambda_l1_addon=30., lambda_group_lasso=2., lambda_gene_expr_recon=300.,
           lambda_edge_recon=500000.)
CASES = {
    # name: (conv_layer, node_label_method, n_addon_gp, n_genes, use_only_active_gps)
    "gatv2_norm_addon3": ("gatv2conv", "one-hop-norm", 3, 40, False),
    "gcn_sum_addon3": ("gcnconv", "one-hop-sum", 3, 40, False),
    "gatv2_attention_addon0": ("gatv2conv", "one-hop-attention", 0, 40, False),
    "gatv2_norm_active_gps": ("gatv2conv", "one-hop-norm", 3, 41, True),
    # (+ n_fc_layers_encoder, n_layers_encoder): fc_l2 + BatchNorm1d, conv_l1
    "gatv2_fc2_bn_two_gnn_layers": ("gatv2conv", "one-hop-norm", 3, 40, False, 2, 2),
    "gcn_fc2_bn": ("gcnconv", "one-hop-norm", 3, 40, False, 2, 1),
}


def batch_arrays(prefix, b, out):
    out[f"{prefix}/x"] = b.x.numpy()
    out[f"{prefix}/edge_index"] = b.edge_index.numpy()
    if hasattr(b, "edge_label_index"):
        out[f"{prefix}/edge_label_index"] = b.edge_label_index.numpy()
        out[f"{prefix}/edge_label"] = b.edge_label.numpy()
    else:
        out[f"{prefix}/batch_size"] = np.int64(b.batch_size)


def make_case(name, conv_layer, node_label_method, n_addon, n_genes, only_active, n_fc=1, n_gnn=1):
    ref = load_reference()
    import torch_sparse
    torch_sparse.DEFAULT_DTYPE = torch.float64
    adata = H.tiny_problem(n_cells=120, n_genes=n_genes, n_prior_gp=10, seed=len(name))
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=n_addon, conv_layer=conv_layer,
                                  node_label_method=node_label_method)
    torch.manual_seed(1)
    with quiet():
        m = ref.modules.VGPGAE(
            n_input=cfg["n_input"], n_fc_layers_encoder=n_fc, n_layers_encoder=n_gnn,
            n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"], n_addon_gp=n_addon,
            cat_covariates_embeds_nums=[], n_output_genes=cfg["n_output_genes"],
            target_rna_decoder_mask=cfg["target_rna_mask"], source_rna_decoder_mask=cfg["source_rna_mask"],
            features_idx_dict=copy.deepcopy(cfg["features_idx_dict"]),
            features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
            conv_layer_encoder=conv_layer, encoder_n_attention_heads=cfg["heads"],
            node_label_method=node_label_method, active_gp_thresh_ratio=0.01,
            include_cat_covariates_contrastive_loss=False).double()
    m.train()
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith("bias"):
                p.uniform_(-0.1, 0.1)           # PyG zero-initialises biases: make them count
        if n_fc == 2:
            m.encoder.fc_l2_bn.weight.uniform_(0.8, 1.2)
        if only_active:
            m.running_mean_abs_mu.copy_(torch.rand(cfg["n_prior_gp"] + n_addon, dtype=torch.float64) ** 6)
    out = {}
    for k, v in m.state_dict().items():
        out[f"state/{k}"] = v.detach().numpy().copy()
    out["cfg/target_rna_mask"] = cfg["target_rna_mask"].numpy()
    out["cfg/source_rna_mask"] = cfg["source_rna_mask"].numpy()
    for k, v in cfg["features_idx_dict"].items():
        out[f"cfg/features_idx/{k}"] = np.asarray(v, dtype=np.int64)
    out["cfg/meta"] = np.array([conv_layer, node_label_method, str(n_addon), str(cfg["n_hidden"]),
                                str(cfg["heads"]), str(int(only_active)), str(n_fc), str(n_gnn)])

    torch.manual_seed(0)
    x = H.dense_x(adata, torch.float64)
    n = adata.n_obs
    ei = R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    splits = R.edge_level_split(ei, n, 0.1)
    train_mask, _ = R.node_level_split_mask(n, 0.1)
    node_b = next(iter(R.NodeLoaderRef(x, ei, n, train_mask, 8)))
    edge_b = next(iter(R.LinkLoaderRef(x, splits["train"].edge_index, n, splits["train"].edge_label_index, 16)))
    batch_arrays("node_batch", node_b, out)
    batch_arrays("edge_batch", edge_b, out)
    P = cfg["n_prior_gp"] + n_addon
    g = torch.Generator().manual_seed(5)
    eps_n = torch.randn(node_b.batch_size, P, generator=g, dtype=torch.float64)
    eps_e = torch.randn(2 * edge_b.edge_label_index.size(1), P, generator=g, dtype=torch.float64)
    out["eps/node"], out["eps/edge"] = eps_n.numpy(), eps_e.numpy()

    draws = [eps_n, eps_e]
    real = torch.randn_like
    torch.randn_like = lambda t, *a, **k: draws.pop(0)
    try:
        rn = m(data_batch=node_b, decoder="omics", use_only_active_gps=only_active)
        re_ = m(data_batch=edge_b, decoder="graph", use_only_active_gps=only_active)
    finally:
        torch.randn_like = real
    rl = m.loss(edge_model_output=re_, node_model_output=rn, l1_targets_mask=None, l1_sources_mask=None, **LAM)
    rl["optim_loss"].backward()
    for k in ("mu", "logstd", "target_rna_nb_means", "source_rna_nb_means"):
        out[f"node_out/{k}"] = rn[k].detach().numpy()
    for k in ("target_rna", "source_rna"):
        out[f"node_out/node_labels/{k}"] = rn["node_labels"][k].detach().numpy()
    for k in ("mu", "logstd", "edge_recon_logits"):
        out[f"edge_out/{k}"] = re_[k].detach().numpy()
    for k, v in rl.items():
        out[f"loss/{k}"] = np.float64(float(v))
    for k, p in m.named_parameters():
        if p.grad is not None:
            out[f"grad/{k}"] = p.grad.numpy().copy()
    out["after/running_mean_abs_mu"] = m.running_mean_abs_mu.numpy().copy()
    if n_fc == 2:
        out["after/fc_l2_bn.running_mean"] = m.encoder.fc_l2_bn.running_mean.numpy().copy()
        out["after/fc_l2_bn.running_var"] = m.encoder.fc_l2_bn.running_var.numpy().copy()
    m.eval()
    out["after/active_gp_mask"] = m.get_active_gp_mask().numpy()
    with torch.no_grad():
        mu, std = m.get_latent_representation(node_b, only_active_gps=True, return_mu_std=True)
    out["latent/mu"], out["latent/std"] = mu.numpy(), std.numpy()
    path = os.path.join(OUT, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB, "
          f"global_loss {float(rl['global_loss']):.10g}")


# ----------------------------------------------------------------------------- C3 / C4 shapes
LAM_MM = dict(lambda_l1_masked=3., lambda_l1_addon=30., lambda_group_lasso=2., lambda_gene_expr_recon=300.,
              lambda_chrom_access_recon=100., lambda_edge_recon=500000., lambda_cat_covariates_contrastive=1000.,
              contrastive_logits_pos_ratio=0.25, contrastive_logits_neg_ratio=0.25)
CASES_MM = {
    # name: (with_atac, with_covariates, conv_layer, embedding injection)
    "mm_covariates_gatv2": (False, True, "gatv2conv", ("encoder", "gene_expr_decoder")),
    "mm_atac_gatv2": (True, False, "gatv2conv", ()),
    "mm_atac_covariates_gcn": (True, True, "gcnconv", ("gene_expr_decoder", "chrom_access_decoder")),
}


def make_case_mm(name, with_atac, with_cov, conv_layer, injection):
    """Multi-sample (categorical covariates: embeddings, edge inclusion, contrastive loss) and
    multimodal (RNA + ATAC decoders) cases; same protocol as ``make_case``."""
    ref = load_reference()
    import torch_sparse
    torch_sparse.DEFAULT_DTYPE = torch.float64
    adata, cfg, x, cats = H.multimodal_covariate_problem(conv_layer=conv_layer, injection=injection,
                                                         with_atac=with_atac, with_covariates=with_cov,
                                                         seed=len(name))
    kw = {}
    if with_atac:
        kw.update(n_output_peaks=cfg["n_output_peaks"], target_atac_decoder_mask=cfg["target_atac_mask"],
                  source_atac_decoder_mask=cfg["source_atac_mask"], gene_peaks_mask=cfg["gene_peaks_mask"])
    if with_cov:
        kw.update(cat_covariates_cats=[["sample0", "sample1"], ["a", "b", "c"]],
                  cat_covariates_no_edges=cfg["cat_covariates_no_edges"],
                  cat_covariates_embeds_injection=cfg["cat_covariates_embeds_injection"])
    torch.manual_seed(4)
    with quiet():
        m = ref.modules.VGPGAE(
            n_input=cfg["n_input"], n_fc_layers_encoder=1, n_layers_encoder=1,
            n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"], n_addon_gp=cfg["n_addon_gp"],
            cat_covariates_embeds_nums=cfg["cat_covariates_embeds_nums"], n_output_genes=cfg["n_output_genes"],
            target_rna_decoder_mask=cfg["target_rna_mask"], source_rna_decoder_mask=cfg["source_rna_mask"],
            features_idx_dict=copy.deepcopy(cfg["features_idx_dict"]),
            features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
            conv_layer_encoder=conv_layer, encoder_n_attention_heads=cfg["heads"],
            node_label_method=cfg["node_label_method"], active_gp_thresh_ratio=0.01,
            include_cat_covariates_contrastive_loss=with_cov, **kw).double()
    m.train()
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    out = {}
    for k, v in m.state_dict().items():
        out[f"state/{k}"] = v.detach().numpy().copy()
    for k in ("target_rna_mask", "source_rna_mask", "target_atac_mask", "source_atac_mask", "gene_peaks_mask"):
        if k in cfg:
            out[f"cfg/{k}"] = cfg[k].numpy()
    for k, v in cfg["features_idx_dict"].items():
        out[f"cfg/features_idx/{k}"] = np.asarray(v, dtype=np.int64)
    out["cfg/meta"] = np.array([conv_layer, cfg["node_label_method"], str(cfg["n_addon_gp"]), str(cfg["n_hidden"]),
                                str(cfg["heads"]), str(int(with_atac)), str(int(with_cov)), ",".join(injection)])


    torch.manual_seed(0)
    n = adata.n_obs
    ei = R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    splits = R.edge_level_split(ei, n, 0.1)
    train_mask, _ = R.node_level_split_mask(n, 0.1)
    node_b = next(iter(R.NodeLoaderRef(x, ei, n, train_mask, 8, cat_covariates_cats=cats)))
    edge_b = next(iter(R.LinkLoaderRef(x, splits["train"].edge_index, n, splits["train"].edge_label_index, 16,
                                       cat_covariates_cats=cats)))
    edge_b.edge_label = edge_b.edge_label.double()
    batch_arrays("node_batch", node_b, out)
    batch_arrays("edge_batch", edge_b, out)
    if with_cov:
        out["node_batch/cat_covariates_cats"] = node_b.cat_covariates_cats.numpy()
        out["edge_batch/cat_covariates_cats"] = edge_b.cat_covariates_cats.numpy()
    P = cfg["n_prior_gp"] + cfg["n_addon_gp"]
    g = torch.Generator().manual_seed(5)
    eps_n = torch.randn(node_b.batch_size, P, generator=g, dtype=torch.float64)
    eps_e = torch.randn(2 * edge_b.edge_label_index.size(1), P, generator=g, dtype=torch.float64)
    out["eps/node"], out["eps/edge"] = eps_n.numpy(), eps_e.numpy()
    draws = [eps_n, eps_e]
    real = torch.randn_like
    torch.randn_like = lambda t, *a, **k: draws.pop(0)
    try:
        rn = m(data_batch=node_b, decoder="omics")
        re_ = m(data_batch=copy.deepcopy(edge_b), decoder="graph")      # the contrastive loss relabels in place
    finally:
        torch.randn_like = real
    lam = dict(LAM_MM)
    if not with_cov:
        lam["lambda_cat_covariates_contrastive"] = 0.
    rl = m.loss(edge_model_output=re_, node_model_output=rn, l1_targets_mask=None, l1_sources_mask=None, **lam)
    rl["optim_loss"].backward()
    mods = ["rna"] + (["atac"] if with_atac else [])
    for k in ("mu", "logstd"):
        out[f"node_out/{k}"] = rn[k].detach().numpy()
        out[f"edge_out/{k}"] = re_[k].detach().numpy()
    for mod in mods:
        for ent in ("target", "source"):
            out[f"node_out/{ent}_{mod}_nb_means"] = rn[f"{ent}_{mod}_nb_means"].detach().numpy()
            out[f"node_out/node_labels/{ent}_{mod}"] = rn["node_labels"][f"{ent}_{mod}"].detach().numpy()
    out["edge_out/edge_recon_logits"] = re_["edge_recon_logits"].detach().numpy()
    if with_cov:
        out["edge_out/edge_incl"] = re_["edge_incl"].numpy()
    for k, v in rl.items():
        out[f"loss/{k}"] = np.float64(float(v))
    for k, p in m.named_parameters():
        if p.grad is not None:
            out[f"grad/{k}"] = p.grad.numpy().copy()
    out["after/running_mean_abs_mu"] = m.running_mean_abs_mu.numpy().copy()
    m.eval()
    with torch.no_grad():
        mu, std = m.get_latent_representation(node_b, only_active_gps=False, return_mu_std=True)
    out["latent/mu"], out["latent/std"] = mu.numpy(), std.numpy()
    path = os.path.join(OUT, f"{name}.npz")
    np.savez_compressed(path, **out)
    print(f"{name}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB, "
          f"global_loss {float(rl['global_loss']):.10g}")


if __name__ == "__main__":
    only_mm = "--mm-only" in sys.argv
    only = sys.argv[sys.argv.index("--only") + 1:] if "--only" in sys.argv else None     # --only NAME [NAME ...]
    if not only_mm:
        for name, args in CASES.items():
            if only is None or name in only:
                make_case(name, *args)
    for name, args in CASES_MM.items():
        if only is None or name in only:
            make_case_mm(name, *args)
