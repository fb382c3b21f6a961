"""
Golden-vector parity.  tests/golden/*.npz were written by tests/golden/make_golden.py from the
UNMODIFIED reference modules (fp64).  Here the stored inputs (weights, sampled batches, injected
noise) are replayed
  * through the CPU oracle           (not gpu; tolerance 1e-10 -- pins the oracle), and
  * through the CUDA module, fp32    (gpu; 1e-5 relative on outputs / losses, 2e-5 on gradients).
Neither needs /root/reference at test time.
"""
import copy
import glob
import os
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle.vgpgae_ref import VGPGAERef
from tests import helpers as H

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_ALL = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
CASES = [c for c in _ALL if not c.startswith("mm_")]
CASES_MM = [c for c in _ALL if c.startswith("mm_")]       # multi-sample covariates / RNA + ATAC (C3 / C4 shapes)
LAM_MM = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
              lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_chrom_access_recon=100.,
              lambda_edge_recon=500000., lambda_cat_covariates_contrastive=1000.,
              contrastive_logits_pos_ratio=0.25, contrastive_logits_neg_ratio=0.25)
NON_PARAM = ("mask", "abs_mu", "running_mean", "running_var", "num_batches_tracked")
LAM = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
           lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_edge_recon=500000.)


def load_case(name):
    z = np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))
    meta = [str(s) for s in z["cfg/meta"]]
    conv_layer, method, n_addon, n_hidden, heads, only_active = meta[:6]
    n_fc, n_gnn = (int(meta[6]), int(meta[7])) if len(meta) > 6 else (1, 1)
    tm, sm = torch.from_numpy(z["cfg/target_rna_mask"]), torch.from_numpy(z["cfg/source_rna_mask"])
    feat = {k[len("cfg/features_idx/"):]: z[k].tolist() for k in z.files if k.startswith("cfg/features_idx/")}
    F = tm.shape[1]
    cfg = dict(n_input=F, n_output_genes=F, n_output_peaks=0, n_hidden=int(n_hidden), n_prior_gp=tm.shape[0],
               n_addon_gp=int(n_addon), conv_layer=conv_layer, heads=int(heads), n_layers=n_gnn, n_fc_layers=n_fc,
               node_label_method=method, log_variational=True, active_gp_thresh_ratio=0.01,
               active_gp_type="separate", include_edge_kl_loss=True, include_edge_recon_loss=True,
               include_gene_expr_recon_loss=True, include_chrom_access_recon_loss=True,
               include_cat_covariates_contrastive_loss=False,
               cat_covariates_embeds_injection=["gene_expr_decoder", "chrom_access_decoder"],
               cat_covariates_embeds_nums=[], cat_covariates_no_edges=[], target_rna_mask=tm, source_rna_mask=sm,
               features_idx_dict=feat)
    state = {k[len("state/"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("state/")}
    node_b = SimpleNamespace(x=torch.from_numpy(z["node_batch/x"]),
                             edge_index=torch.from_numpy(z["node_batch/edge_index"]),
                             batch_size=int(z["node_batch/batch_size"]))
    edge_b = SimpleNamespace(x=torch.from_numpy(z["edge_batch/x"]),
                             edge_index=torch.from_numpy(z["edge_batch/edge_index"]),
                             edge_label_index=torch.from_numpy(z["edge_batch/edge_label_index"]),
                             edge_label=torch.from_numpy(z["edge_batch/edge_label"]))
    return z, cfg, state, node_b, edge_b, bool(int(only_active))


def test_golden_files_present():
    assert len(CASES) >= 4, CASES


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_reference_golden(name):
    z, cfg, state, node_b, edge_b, only_active = load_case(name)
    params = {k: v.clone().requires_grad_(v.dtype.is_floating_point and not k.endswith(NON_PARAM))
              for k, v in state.items()}
    orc = VGPGAERef(cfg, params)
    on = orc.forward(node_b, "omics", use_only_active_gps=only_active, eps=torch.from_numpy(z["eps/node"]))
    oe = orc.forward(edge_b, "graph", use_only_active_gps=only_active, eps=torch.from_numpy(z["eps/edge"]))
    ol = orc.loss(oe, on, **LAM)
    ol["optim_loss"].backward()
    tol = 1e-10
    for k in ("mu", "logstd", "target_rna_nb_means", "source_rna_nb_means"):
        assert H.rel_err(on[k], torch.from_numpy(z[f"node_out/{k}"])) < tol, k
    for k in ("target_rna", "source_rna"):
        assert H.rel_err(on["node_labels"][k], torch.from_numpy(z[f"node_out/node_labels/{k}"])) < 1e-12, k
    for k in ("mu", "logstd", "edge_recon_logits"):
        assert H.rel_err(oe[k], torch.from_numpy(z[f"edge_out/{k}"])) < tol, k
    loss_keys = [k for k in z.files if k.startswith("loss/")]
    assert len(loss_keys) >= 7
    for k in loss_keys:
        assert H.rel_err(torch.as_tensor(ol[k[5:]]), torch.tensor(float(z[k]), dtype=torch.float64)) < tol, k
    n_grads = 0
    for k in z.files:
        if k.startswith("grad/"):
            g = params[k[5:]].grad
            assert g is not None, k
            assert H.rel_err(g, torch.from_numpy(z[k])) < 1e-9, k
            n_grads += 1
    assert n_grads >= 10
    assert H.rel_err(orc.running_mean_abs_mu, torch.from_numpy(z["after/running_mean_abs_mu"])) < 1e-12
    for k in ("running_mean", "running_var"):
        if f"after/fc_l2_bn.{k}" in z.files:
            assert H.rel_err(params[f"encoder.fc_l2_bn.{k}"], torch.from_numpy(z[f"after/fc_l2_bn.{k}"])) < 1e-12, k
    orc.training = False
    assert np.array_equal(orc.get_active_gp_mask().numpy(), z["after/active_gp_mask"])
    with torch.no_grad():
        mu, std = orc.get_latent_representation(node_b, only_active_gps=True, return_mu_std=True)
    assert H.rel_err(mu, torch.from_numpy(z["latent/mu"])) < tol
    assert H.rel_err(std, torch.from_numpy(z["latent/std"])) < tol


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_reproduces_reference_golden(name):
    from tests.test_vgpgae_gpu import build_module
    z, cfg, state, node_b, edge_b, only_active = load_case(name)
    mod = build_module(cfg, state, cfg["conv_layer"], cfg["node_label_method"]).set_compute_dtype(torch.float32)
    mod.train()
    mod._eps_queue = [torch.from_numpy(z["eps/node"]), torch.from_numpy(z["eps/edge"])]
    cn = mod(H.batch_to(node_b, torch.float32, "cuda"), "omics", use_only_active_gps=only_active)
    ce = mod(H.batch_to(edge_b, torch.float32, "cuda"), "graph", use_only_active_gps=only_active)
    cl = mod.loss(ce, cn, **LAM)
    cl["optim_loss"].backward()
    tol = 1e-5
    for k in ("mu", "logstd", "target_rna_nb_means", "source_rna_nb_means"):
        assert H.rel_err(cn[k], torch.from_numpy(z[f"node_out/{k}"])) < tol, k
    for k in ("mu", "logstd", "edge_recon_logits"):
        assert H.rel_err(ce[k], torch.from_numpy(z[f"edge_out/{k}"])) < tol, k
    for k in z.files:
        if k.startswith("loss/"):
            assert H.rel_err(torch.as_tensor(cl[k[5:]]), torch.tensor(float(z[k]), dtype=torch.float64)) < tol, k
    checked = 0
    grads = dict(mod.named_parameters())
    for k in z.files:
        if k.startswith("grad/"):
            ref = torch.from_numpy(z[k])
            g = grads[k[5:]].grad
            if float(ref.abs().max()) == 0.0:
                assert g is None or float(g.abs().max()) == 0.0, k
                continue
            assert g is not None, k
            assert H.rel_err(g, ref) < 2e-5, k
            checked += 1
    assert checked >= 8
    assert H.rel_err(mod.running_mean_abs_mu, torch.from_numpy(z["after/running_mean_abs_mu"])) < tol
    for k in ("running_mean", "running_var"):
        if f"after/fc_l2_bn.{k}" in z.files:
            assert H.rel_err(getattr(mod.encoder.fc_l2_bn, k), torch.from_numpy(z[f"after/fc_l2_bn.{k}"])) < tol, k
    mod.eval()
    assert np.array_equal(mod.get_active_gp_mask().cpu().numpy(), z["after/active_gp_mask"])
    with torch.no_grad():
        mu, std = mod.get_latent_representation(H.batch_to(node_b, torch.float32, "cuda"), only_active_gps=True,
                                                return_mu_std=True)
    assert H.rel_err(mu, torch.from_numpy(z["latent/mu"])) < tol
    assert H.rel_err(std, torch.from_numpy(z["latent/std"])) < tol


# ----------------------------------------------------------------------------- C3 / C4 shapes
def load_case_mm(name):
    z = np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))
    conv_layer, method, n_addon, n_hidden, heads, with_atac, with_cov, inj = [str(s) for s in z["cfg/meta"]]
    with_atac, with_cov = bool(int(with_atac)), bool(int(with_cov))
    tm, sm = torch.from_numpy(z["cfg/target_rna_mask"]), torch.from_numpy(z["cfg/source_rna_mask"])
    feat = {k[len("cfg/features_idx/"):]: z[k].tolist() for k in z.files if k.startswith("cfg/features_idx/")}
    F = tm.shape[1]
    cfg = dict(n_input=F, n_output_genes=F, n_output_peaks=0, n_hidden=int(n_hidden), n_prior_gp=tm.shape[0],
               n_addon_gp=int(n_addon), conv_layer=conv_layer, heads=int(heads), n_layers=1,
               node_label_method=method, log_variational=True, active_gp_thresh_ratio=0.01,
               active_gp_type="separate", include_edge_kl_loss=True, include_edge_recon_loss=True,
               include_gene_expr_recon_loss=True, include_chrom_access_recon_loss=True,
               include_cat_covariates_contrastive_loss=with_cov,
               cat_covariates_embeds_injection=[s_ for s_ in inj.split(",") if s_],
               cat_covariates_embeds_nums=[], cat_covariates_no_edges=[], target_rna_mask=tm, source_rna_mask=sm,
               features_idx_dict=feat)
    if with_atac:
        gpm = torch.from_numpy(z["cfg/gene_peaks_mask"])
        cfg.update(n_input=F + gpm.shape[1], n_output_peaks=gpm.shape[1], gene_peaks_mask=gpm,
                   target_atac_mask=torch.from_numpy(z["cfg/target_atac_mask"]),
                   source_atac_mask=torch.from_numpy(z["cfg/source_atac_mask"]))
    if with_cov:
        cfg.update(cat_covariates_embeds_nums=[3, 2], cat_covariates_num_cats=[2, 3],
                   cat_covariates_no_edges=[True, False])
    state = {k[len("state/"):]: torch.from_numpy(z[k]) for k in z.files if k.startswith("state/")}
    node_b = SimpleNamespace(x=torch.from_numpy(z["node_batch/x"]),
                             edge_index=torch.from_numpy(z["node_batch/edge_index"]),
                             batch_size=int(z["node_batch/batch_size"]))
    edge_b = SimpleNamespace(x=torch.from_numpy(z["edge_batch/x"]),
                             edge_index=torch.from_numpy(z["edge_batch/edge_index"]),
                             edge_label_index=torch.from_numpy(z["edge_batch/edge_label_index"]),
                             edge_label=torch.from_numpy(z["edge_batch/edge_label"]))
    if with_cov:
        node_b.cat_covariates_cats = torch.from_numpy(z["node_batch/cat_covariates_cats"])
        edge_b.cat_covariates_cats = torch.from_numpy(z["edge_batch/cat_covariates_cats"])
    lam = dict(LAM_MM)
    if not with_cov:
        lam["lambda_cat_covariates_contrastive"] = 0.
    return z, cfg, state, node_b, edge_b, with_atac, with_cov, lam


def _check_mm(z, node_out, edge_out, losses, grad_of, with_atac, with_cov, tol, gtol, loss_tol32):
    mods = ["rna"] + (["atac"] if with_atac else [])
    for k in ("mu", "logstd"):
        assert H.rel_err(node_out[k], torch.from_numpy(z[f"node_out/{k}"])) < tol, k
        assert H.rel_err(edge_out[k], torch.from_numpy(z[f"edge_out/{k}"])) < tol, k
    for mod in mods:
        for ent in ("target", "source"):
            k = f"{ent}_{mod}_nb_means"
            assert H.rel_err(node_out[k], torch.from_numpy(z[f"node_out/{k}"])) < tol, k
            assert H.rel_err(node_out["node_labels"][f"{ent}_{mod}"],
                             torch.from_numpy(z[f"node_out/node_labels/{ent}_{mod}"])) < max(1e-12, tol * 0.1)
    assert H.rel_err(edge_out["edge_recon_logits"], torch.from_numpy(z["edge_out/edge_recon_logits"])) < tol
    if with_cov:
        assert np.array_equal(edge_out["edge_incl"].cpu().numpy(), z["edge_out/edge_incl"])
    loss_keys = [k for k in z.files if k.startswith("loss/")]
    assert ("loss/chrom_access_recon_loss" in loss_keys) == with_atac
    assert ("loss/cat_covariates_contrastive_loss" in loss_keys) == with_cov
    # the reference rounds the contrastive loss (and the sums containing it) to fp32 (modules/losses.py:80,118)
    fp32_keys = ("cat_covariates_contrastive_loss", "global_loss", "optim_loss") if with_cov else ()
    for k in loss_keys:
        t = max(tol, loss_tol32) if k[5:] in fp32_keys else tol
        assert H.rel_err(torch.as_tensor(losses[k[5:]]), torch.tensor(float(z[k]), dtype=torch.float64)) < t, k
    n_grads = 0
    for k in z.files:
        if k.startswith("grad/"):
            ref = torch.from_numpy(z[k])
            g = grad_of(k[5:])
            if float(ref.abs().max()) == 0.0:
                assert g is None or float(g.abs().max()) == 0.0, k
                continue
            assert g is not None, k
            assert H.rel_err(g, ref) < gtol, k
            n_grads += 1
    assert n_grads >= 10


def test_golden_mm_files_present():
    assert len(CASES_MM) >= 3, CASES_MM


@pytest.mark.parametrize("name", CASES_MM)
def test_oracle_reproduces_reference_golden_mm(name):
    z, cfg, state, node_b, edge_b, with_atac, with_cov, lam = load_case_mm(name)
    params = {k: v.clone().requires_grad_(v.dtype.is_floating_point and not k.endswith(("mask", "abs_mu")))
              for k, v in state.items()}
    orc = VGPGAERef(cfg, params)
    on = orc.forward(node_b, "omics", eps=torch.from_numpy(z["eps/node"]))
    oe = orc.forward(copy.deepcopy(edge_b), "graph", eps=torch.from_numpy(z["eps/edge"]))
    ol = orc.loss(oe, on, **lam)
    ol["optim_loss"].backward()
    _check_mm(z, on, oe, ol, lambda k: params[k].grad, with_atac, with_cov, 1e-10, 1e-9, 2e-7)
    assert H.rel_err(orc.running_mean_abs_mu, torch.from_numpy(z["after/running_mean_abs_mu"])) < 1e-12
    orc.training = False
    with torch.no_grad():
        mu, std = orc.get_latent_representation(node_b, only_active_gps=False, return_mu_std=True)
    assert H.rel_err(mu, torch.from_numpy(z["latent/mu"])) < 1e-10
    assert H.rel_err(std, torch.from_numpy(z["latent/std"])) < 1e-10


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES_MM)
def test_cuda_reproduces_reference_golden_mm(name):
    from tests.test_vgpgae_gpu import build_module_mm
    z, cfg, state, node_b, edge_b, with_atac, with_cov, lam = load_case_mm(name)
    mod = build_module_mm(cfg, state, with_atac, with_cov).set_compute_dtype(torch.float32)
    mod.train()
    mod._eps_queue = [torch.from_numpy(z["eps/node"]), torch.from_numpy(z["eps/edge"])]
    cn = mod(H.batch_to(node_b, torch.float32, "cuda"), "omics")
    ce = mod(H.batch_to(copy.deepcopy(edge_b), torch.float32, "cuda"), "graph")
    cl = mod.loss(ce, cn, **lam)
    cl["optim_loss"].backward()
    grads = dict(mod.named_parameters())
    _check_mm(z, cn, ce, cl, lambda k: grads[k].grad, with_atac, with_cov, 1e-5, 2e-5, 1e-5)
    assert H.rel_err(mod.running_mean_abs_mu, torch.from_numpy(z["after/running_mean_abs_mu"])) < 1e-5
    mod.eval()
    with torch.no_grad():
        mu, std = mod.get_latent_representation(H.batch_to(node_b, torch.float32, "cuda"), only_active_gps=False,
                                                return_mu_std=True)
    assert H.rel_err(mu, torch.from_numpy(z["latent/mu"])) < 1e-5
    assert H.rel_err(std, torch.from_numpy(z["latent/std"])) < 1e-5
