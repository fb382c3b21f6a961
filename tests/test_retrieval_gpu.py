"""
SURVEY.md section 8 (f2): ``get_omics_decoder_outputs`` / ``get_neighbor_importances`` / ``get_recon_edge_probs``
(models/nichecompass.py:1383-1686) and the exported ``compute_omics_recon_nb_loss`` (modules/losses.py:169-212),
``get_active_gp_mask`` aggregation modes (modules/vgpgae.py:1289-1330) on the CUDA kernels against the CPU oracle /
plain fp64 torch.
"""
import contextlib
import io
import math

import numpy as np
import pytest
import torch

from oracle import pyg_semantics as S
from oracle import train_ref as TR
from oracle import vgpgae_ref as V
from tests import helpers as H

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def trained():
    from nichecompass_b200.models import NicheCompass
    adata = H.tiny_problem(n_cells=300, n_genes=48, n_prior_gp=12, lib_mean=80.)
    with contextlib.redirect_stdout(io.StringIO()):
        model = NicheCompass(adata, n_addon_gp=4, active_gp_thresh_ratio=0.01)
        model.train(n_epochs=2, n_epochs_all_gps=1, edge_batch_size=64, verbose=False)
    sd = {k: v.detach().double().cpu() for k, v in model.model.state_dict().items()}
    orc = V.VGPGAERef(H.oracle_cfg_from_adata(adata, n_addon_gp=4), sd)
    orc.training = False
    x = H.dense_x(adata, torch.float64)
    ei = TR.R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    return adata, model, orc, x, ei


def test_omics_decoder_outputs_match_oracle(trained):
    adata, model, orc, x, ei = trained
    out = model.get_omics_decoder_outputs(node_batch_size=64)
    loader = TR.R.NodeLoaderRef(x, ei, 300, torch.ones(300, dtype=torch.bool), 64, shuffle=False)
    ref = [orc.get_omics_decoder_outputs(b, only_active_gps=True) for b in loader]
    for key in ("target_rna_nb_means", "source_rna_nb_means"):
        want = torch.cat([r[key] for r in ref])
        assert out[key].shape == (300, 48) and out[key].dtype == np.float64
        assert H.rel_err(torch.from_numpy(out[key]), want) < 2e-5, key
    # App. B-1: no column selection here -> every row sums to its library size
    lib = torch.cat([r["node_labels"]["target_rna"].sum(1) for r in ref])
    assert np.allclose(out["target_rna_nb_means"].sum(1), lib.numpy(), rtol=1e-5)
    out_all = model.get_omics_decoder_outputs(only_active_gps=False, node_batch_size=100)
    loader = TR.R.NodeLoaderRef(x, ei, 300, torch.ones(300, dtype=torch.bool), 100, shuffle=False)
    want = torch.cat([orc.get_omics_decoder_outputs(b, only_active_gps=False)["source_rna_nb_means"] for b in loader])
    assert H.rel_err(torch.from_numpy(out_all["source_rna_nb_means"]), want) < 2e-5


def test_neighbor_importances_match_oracle(trained):
    adata, model, orc, x, ei = trained
    w = model.get_neighbor_importances(node_batch_size=50)
    assert w.shape == (300, 300)
    loader = TR.R.NodeLoaderRef(x, ei, 300, torch.ones(300, dtype=torch.bool), 50, shuffle=False)
    dense = np.zeros((300, 300))
    for b in loader:
        # adj_norm[r, c] of the sampled block (App. A.3); x_nbr[c] = sum_r adj_norm[r, c] x[r]: the weight target c
        # gives to neighbour r
        e, val = S.gcn_norm_sparse(b.edge_index, b.x.size(0), torch.float64)
        n_id = b.n_id.numpy()
        keep = e[1] < b.batch_size
        dense[n_id[e[1][keep].numpy()], n_id[e[0][keep].numpy()]] = val[keep].numpy()
    got = np.asarray(w.todense())
    assert (got != 0).sum() == (dense != 0).sum()
    assert np.allclose(got, dense, rtol=1e-6, atol=1e-9)
    # the weights reproduce the aggregated labels: W @ x == source node labels of a full-graph block
    full = model.get_neighbor_importances(node_batch_size=300)
    xs, _ = S.one_hop_gcn_norm_aggregate(x, ei)
    assert np.allclose(np.asarray(full @ x.numpy()), xs.numpy(), rtol=1e-5, atol=1e-6)


def test_recon_edge_probs_match_fp64(trained):
    adata, model, orc, x, ei = trained
    z = torch.from_numpy(np.asarray(adata.obsm["nichecompass_latent"]))
    active = orc.get_active_gp_mask()
    zf = torch.zeros(300, active.numel(), dtype=torch.float64)
    zf[:, active] = z
    zn = zf / zf.norm(dim=1, keepdim=True).clamp_min(1e-8)
    probs = torch.sigmoid(zn @ zn.t())
    got = model.get_recon_edge_probs(return_edge_probs=True, node_batch_size=128)
    assert got.shape == (300, 300) and H.rel_err(got, probs) < 1e-5
    # fixed threshold
    thr = float(probs.flatten().kthvalue(int(0.9 * probs.numel())).values)
    adj = model.get_recon_edge_probs(edge_thresh=thr + 1e-4, node_batch_size=128)
    want = (probs >= thr + 1e-4).numpy()
    assert (np.asarray(adj.todense()).astype(bool) != want).mean() < 1e-4
    # k-th largest per row, symmetrised
    adj = np.asarray(model.get_recon_edge_probs(n_neighbors=5, node_batch_size=77).todense()).astype(bool)
    kth = probs.sort(dim=1, descending=True).values[:, 4:5]
    want = (probs >= kth).numpy()
    want = want | want.T
    assert (adj != want).mean() < 1e-4 and (adj == adj.T).all()
    # neighbours of the input graph per row, not symmetrised
    adj = np.asarray(model.get_recon_edge_probs(node_batch_size=300).todense()).astype(bool)
    deg = np.asarray(adata.obsp["spatial_connectivities"].sum(axis=1)).flatten().astype(int)
    kth = probs.sort(dim=1, descending=True).values[torch.arange(300), torch.from_numpy(deg) - 1][:, None]
    assert (adj != (probs >= kth).numpy()).mean() < 1e-4
    # tensor-core path: bf16 operands, looser
    got16 = model.get_recon_edge_probs(return_edge_probs=True, compute_dtype=torch.bfloat16)
    assert H.rel_err(got16, probs) < 5e-3


def test_exported_nb_loss_matches_oracle_with_gradients():
    from nichecompass_b200.modules.losses import compute_omics_recon_nb_loss
    g = torch.Generator().manual_seed(0)
    x = torch.poisson(torch.full((9, 33), 2.5), generator=g)
    mu = (torch.rand(9, 33, generator=g) * 6 + 0.1).requires_grad_(True)
    theta = (torch.rand(33, generator=g) * 4 + 0.2).requires_grad_(True)
    want = V.nb_recon_loss(x.double(), mu.double(), theta.double())
    gm, gt = torch.autograd.grad(want, (mu, theta))
    mu_c = mu.detach().cuda().requires_grad_(True)
    th_c = theta.detach().cuda().requires_grad_(True)
    got = compute_omics_recon_nb_loss(x.cuda(), mu_c, th_c)
    assert abs(float(got) - float(want)) < 2e-6 * abs(float(want))
    (3.0 * got).backward()
    assert H.rel_err(mu_c.grad, 3.0 * gm) < 1e-5 and H.rel_err(th_c.grad, 3.0 * gt) < 1e-5


@pytest.mark.parametrize("mode", ["sum", "nzmeans", "sum+nzmeans", "nzmedians", "sum+nzmedians"])
@pytest.mark.parametrize("normalize", [False, True])
def test_active_gp_mask_modes(trained, mode, normalize):
    """modules/vgpgae.py:1254-1330 statement by statement in fp64 torch."""
    adata, model, orc, x, ei = trained
    m = model.model
    got = m.get_active_gp_mask(abs_gp_weights_agg_mode=mode, normalize_gp_weights_with_features_scale_factors=normalize)
    Pp = orc.cfg["n_prior_gp"]
    scale = m.features_scale_factors_.double().cpu()
    ema = m.running_mean_abs_mu.double().cpu()
    want = torch.zeros(Pp + 4, dtype=torch.bool)
    for gp_type, sl in (("masked", slice(None, Pp)), ("addon", slice(Pp, None))):
        w = orc.get_gp_weights(gp_type)
        if normalize:
            w = w / scale[:, None]
        w = ema[sl] * w
        sums = w.abs().sum(0)
        if mode in ("sum", "sum+nzmeans", "sum+nzmedians"):
            want[sl] |= sums >= 0.01 * sums.max()
        if mode in ("nzmeans", "sum+nzmeans"):
            nz = torch.nan_to_num(sums / torch.count_nonzero(w, dim=0))
            want[sl] |= nz >= 0.01 * nz.max()
        if "nzmedians" in mode:
            wn = torch.where(w == 0, torch.tensor(float("nan"), dtype=w.dtype), w.abs())
            med = torch.nan_to_num(torch.nanmedian(wn, dim=0).values)
            want[sl] |= med >= 0.01 * med.max()
    assert torch.equal(got.cpu(), want), (mode, normalize)
