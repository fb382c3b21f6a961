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
CASES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
LAM = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
           lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_edge_recon=500000.)


def load_case(name):
    z = np.load(os.path.join(GOLDEN_DIR, f"{name}.npz"))
    conv_layer, method, n_addon, n_hidden, heads, only_active = [str(s) for s in z["cfg/meta"]]
    tm, sm = torch.from_numpy(z["cfg/target_rna_mask"]), torch.from_numpy(z["cfg/source_rna_mask"])
    feat = {k[len("cfg/features_idx/"):]: z[k].tolist() for k in z.files if k.startswith("cfg/features_idx/")}
    F = tm.shape[1]
    cfg = dict(n_input=F, n_output_genes=F, n_output_peaks=0, n_hidden=int(n_hidden), n_prior_gp=tm.shape[0],
               n_addon_gp=int(n_addon), conv_layer=conv_layer, heads=int(heads), n_layers=1,
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
    params = {k: v.clone().requires_grad_(v.dtype.is_floating_point and not k.endswith(("mask", "abs_mu")))
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
    mod.eval()
    assert np.array_equal(mod.get_active_gp_mask().cpu().numpy(), z["after/active_gp_mask"])
    with torch.no_grad():
        mu, std = mod.get_latent_representation(H.batch_to(node_b, torch.float32, "cuda"), only_active_gps=True,
                                                return_mu_std=True)
    assert H.rel_err(mu, torch.from_numpy(z["latent/mu"])) < tol
    assert H.rel_err(std, torch.from_numpy(z["latent/std"])) < tol
