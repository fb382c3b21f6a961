"""
Tier-1 pin (builder container only): the oracle restatement must reproduce the UNMODIFIED
reference modules (loaded from /root/reference with oracle/pyg_shim) on identical weights,
batches and injected reparameterisation noise -- fp64, tolerance 1e-10.
"""
import copy

import pytest
import torch

from oracle import sampler_ref as R
from oracle.vgpgae_ref import VGPGAERef
from tests import helpers as H
from tests.ref_harness import load_reference, quiet

pytestmark = pytest.mark.reference


@pytest.fixture(autouse=True)
def _fp64_sparse_values():
    load_reference()
    import torch_sparse
    old = torch_sparse.DEFAULT_DTYPE
    torch_sparse.DEFAULT_DTYPE = torch.float64
    yield
    torch_sparse.DEFAULT_DTYPE = old


def build_reference_module(ref, adata, cfg, conv_layer, node_label_method, n_addon_gp):
    feat = copy.deepcopy(cfg["features_idx_dict"])
    with quiet():
        m = ref.modules.VGPGAE(
            n_input=cfg["n_input"], n_fc_layers_encoder=1, n_layers_encoder=1,
            n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"],
            n_addon_gp=n_addon_gp, cat_covariates_embeds_nums=[],
            n_output_genes=cfg["n_output_genes"],
            target_rna_decoder_mask=cfg["target_rna_mask"],
            source_rna_decoder_mask=cfg["source_rna_mask"],
            features_idx_dict=feat,
            features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
            conv_layer_encoder=conv_layer, encoder_n_attention_heads=cfg["heads"],
            node_label_method=node_label_method, active_gp_thresh_ratio=0.01,
            include_cat_covariates_contrastive_loss=False)
    return m.double()


def make_batches(adata, dtype=torch.float64, edge_bs=16, node_bs=8):
    torch.manual_seed(0)
    x = H.dense_x(adata, dtype)
    n = adata.n_obs
    ei = R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    splits = R.edge_level_split(ei, n, 0.1)
    train_mask, _ = R.node_level_split_mask(n, 0.1)
    nl = R.NodeLoaderRef(x, ei, n, train_mask, node_bs)
    ll = R.LinkLoaderRef(x, splits["train"].edge_index, n, splits["train"].edge_label_index, edge_bs)
    return next(iter(nl)), next(iter(ll))


@pytest.mark.parametrize("conv_layer,node_label_method,n_addon", [
    ("gatv2conv", "one-hop-norm", 3),
    ("gcnconv", "one-hop-sum", 3),
    ("gatv2conv", "one-hop-attention", 0),
])
def test_forward_loss_grads_match_reference(conv_layer, node_label_method, n_addon):
    ref = load_reference()
    adata = H.tiny_problem()
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=n_addon, conv_layer=conv_layer,
                                  node_label_method=node_label_method)
    torch.manual_seed(1)
    m = build_reference_module(ref, adata, cfg, conv_layer, node_label_method, n_addon)
    m.train()
    # non-trivial biases (PyG zero-initialises them)
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    params = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point and
                                                   not k.endswith(("mask", "running_mean_abs_mu")))
              for k, v in m.state_dict().items()}
    orc = VGPGAERef(cfg, params)
    node_b, edge_b = make_batches(adata)
    P = cfg["n_prior_gp"] + n_addon
    g = torch.Generator().manual_seed(5)
    eps_n = torch.randn(node_b.batch_size, P, generator=g, dtype=torch.float64)
    eps_e = torch.randn(2 * edge_b.edge_label_index.size(1), P, generator=g, dtype=torch.float64)

    draws = [eps_n, eps_e]
    real_randn_like = torch.randn_like

    def fake_randn_like(t, *a, **k):
        return draws.pop(0)

    torch.randn_like = fake_randn_like
    try:
        rn = m(data_batch=node_b, decoder="omics")
        re_ = m(data_batch=edge_b, decoder="graph")
    finally:
        torch.randn_like = real_randn_like
    lam = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
               lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_edge_recon=500000.)
    # reference indexes params with a None mask -> param[None] (all entries)
    rl = m.loss(edge_model_output=re_, node_model_output=rn, **lam)
    rl["optim_loss"].backward()

    on = orc.forward(node_b, "omics", eps=eps_n)
    oe = orc.forward(edge_b, "graph", eps=eps_e)
    ol = orc.loss(oe, on, **lam)
    ol["optim_loss"].backward()

    for k in ("mu", "logstd", "target_rna_nb_means", "source_rna_nb_means"):
        assert H.rel_err(on[k], rn[k]) < 1e-10, k
    for k in ("target_rna", "source_rna"):
        assert H.rel_err(on["node_labels"][k], rn["node_labels"][k]) < 1e-12, k
    assert H.rel_err(oe["edge_recon_logits"], re_["edge_recon_logits"]) < 1e-10
    for k, v in rl.items():
        assert H.rel_err(torch.as_tensor(ol[k]), torch.as_tensor(v)) < 1e-10, k
    assert H.rel_err(orc.running_mean_abs_mu, m.running_mean_abs_mu) < 1e-12
    n_checked = 0
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        assert params[k].grad is not None, k
        assert H.rel_err(params[k].grad, p.grad) < 1e-9, k
        n_checked += 1
    assert n_checked >= 10
    # active-GP mask + latent (eval mode)
    m.eval()
    orc.training = False
    assert torch.equal(orc.get_active_gp_mask(), m.get_active_gp_mask())
    mu_r, std_r = m.get_latent_representation(node_b, only_active_gps=True, return_mu_std=True)
    mu_o, std_o = orc.get_latent_representation(node_b, only_active_gps=True, return_mu_std=True)
    assert H.rel_err(mu_o, mu_r) < 1e-10 and H.rel_err(std_o, std_r) < 1e-10


def test_only_active_gps_path_matches_reference():
    ref = load_reference()
    adata = H.tiny_problem(seed=3)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=3)
    torch.manual_seed(2)
    m = build_reference_module(ref, adata, cfg, "gatv2conv", "one-hop-norm", 3)
    m.train()
    with torch.no_grad():
        m.running_mean_abs_mu.copy_(torch.rand(cfg["n_prior_gp"] + 3, dtype=torch.float64) ** 6)
    params = {k: v.detach().clone() for k, v in m.state_dict().items()}
    orc = VGPGAERef(cfg, params)
    node_b, _ = make_batches(adata)
    P = cfg["n_prior_gp"] + 3
    eps_n = torch.randn(node_b.batch_size, P, dtype=torch.float64)
    real = torch.randn_like
    torch.randn_like = lambda t, *a, **k: eps_n
    try:
        with torch.no_grad():
            rn = m(data_batch=node_b, decoder="omics", use_only_active_gps=True)
    finally:
        torch.randn_like = real
    with torch.no_grad():
        on = orc.forward(node_b, "omics", use_only_active_gps=True, eps=eps_n)
    assert on["target_rna_nb_means"].shape == rn["target_rna_nb_means"].shape
    for k in ("target_rna_nb_means", "source_rna_nb_means"):
        assert H.rel_err(on[k], rn[k]) < 1e-10, k
    assert H.rel_err(orc.running_mean_abs_mu, m.running_mean_abs_mu) < 1e-12


# ----------------------------------------------------------------------------- C3 / C4 shapes
def build_reference_module_mm(ref, cfg, with_atac, with_cov):
    feat = copy.deepcopy(cfg["features_idx_dict"])
    kw = {}
    if with_atac:
        kw.update(n_output_peaks=cfg["n_output_peaks"], target_atac_decoder_mask=cfg["target_atac_mask"],
                  source_atac_decoder_mask=cfg["source_atac_mask"], gene_peaks_mask=cfg["gene_peaks_mask"])
    if with_cov:
        kw.update(cat_covariates_cats=[["sample0", "sample1"], ["a", "b", "c"]],
                  cat_covariates_no_edges=cfg["cat_covariates_no_edges"],
                  cat_covariates_embeds_injection=cfg["cat_covariates_embeds_injection"])
    with quiet():
        m = ref.modules.VGPGAE(
            n_input=cfg["n_input"], n_fc_layers_encoder=1, n_layers_encoder=1,
            n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"], n_addon_gp=cfg["n_addon_gp"],
            cat_covariates_embeds_nums=cfg["cat_covariates_embeds_nums"],
            n_output_genes=cfg["n_output_genes"],
            target_rna_decoder_mask=cfg["target_rna_mask"], source_rna_decoder_mask=cfg["source_rna_mask"],
            features_idx_dict=feat,
            features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
            conv_layer_encoder=cfg["conv_layer"], encoder_n_attention_heads=cfg["heads"],
            node_label_method=cfg["node_label_method"], active_gp_thresh_ratio=0.01,
            include_cat_covariates_contrastive_loss=with_cov, **kw)
    return m.double()


def make_batches_mm(adata, x, cats, edge_bs=16, node_bs=8):
    torch.manual_seed(0)
    n = adata.n_obs
    ei = R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    splits = R.edge_level_split(ei, n, 0.1)
    train_mask, _ = R.node_level_split_mask(n, 0.1)
    nl = R.NodeLoaderRef(x, ei, n, train_mask, node_bs, cat_covariates_cats=cats)
    ll = R.LinkLoaderRef(x, splits["train"].edge_index, n, splits["train"].edge_label_index, edge_bs,
                         cat_covariates_cats=cats)
    return next(iter(nl)), next(iter(ll))


LAM_MM = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
              lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_chrom_access_recon=100.,
              lambda_edge_recon=500000., lambda_cat_covariates_contrastive=1000.,
              contrastive_logits_pos_ratio=0.25, contrastive_logits_neg_ratio=0.25)


@pytest.mark.parametrize("with_atac,with_cov,conv_layer,injection", [
    (False, True, "gatv2conv", ("encoder", "gene_expr_decoder")),          # C3: multi-sample atlas
    (True, False, "gatv2conv", ()),                                        # C4: RNA + ATAC
    (True, True, "gcnconv", ("gene_expr_decoder", "chrom_access_decoder")),
])
def test_covariates_and_atac_match_reference(with_atac, with_cov, conv_layer, injection):
    ref = load_reference()
    adata, cfg, x, cats = H.multimodal_covariate_problem(conv_layer=conv_layer, injection=injection,
                                                         with_atac=with_atac, with_covariates=with_cov)
    torch.manual_seed(4)
    m = build_reference_module_mm(ref, cfg, with_atac, with_cov)
    m.train()
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    params = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point and
                                                   not k.endswith(("mask", "running_mean_abs_mu")))
              for k, v in m.state_dict().items()}
    orc = VGPGAERef(cfg, params)
    node_b, edge_b = make_batches_mm(adata, x, cats)
    edge_b.edge_label = edge_b.edge_label.double()      # keep the reference's contrastive BCE in fp64 too
    P = cfg["n_prior_gp"] + cfg["n_addon_gp"]
    g = torch.Generator().manual_seed(5)
    eps_n = torch.randn(node_b.batch_size, P, generator=g, dtype=torch.float64)
    eps_e = torch.randn(2 * edge_b.edge_label_index.size(1), P, generator=g, dtype=torch.float64)
    draws = [eps_n, eps_e]
    real_randn_like = torch.randn_like
    torch.randn_like = lambda t, *a, **k: draws.pop(0)
    try:
        rn = m(data_batch=node_b, decoder="omics")
        re_ = m(data_batch=copy.deepcopy(edge_b), decoder="graph")
    finally:
        torch.randn_like = real_randn_like
    lam = dict(LAM_MM)
    if not with_cov:
        lam["lambda_cat_covariates_contrastive"] = 0.
    rl = m.loss(edge_model_output=re_, node_model_output=rn, **lam)
    rl["optim_loss"].backward()
    on = orc.forward(node_b, "omics", eps=eps_n)
    oe = orc.forward(copy.deepcopy(edge_b), "graph", eps=eps_e)
    ol = orc.loss(oe, on, **lam)
    ol["optim_loss"].backward()

    mods = ["rna"] + (["atac"] if with_atac else [])
    for mod in mods:
        for ent in ("target", "source"):
            assert H.rel_err(on[f"{ent}_{mod}_nb_means"], rn[f"{ent}_{mod}_nb_means"]) < 1e-10, (ent, mod)
            assert H.rel_err(on["node_labels"][f"{ent}_{mod}"], rn["node_labels"][f"{ent}_{mod}"]) < 1e-12
    assert H.rel_err(oe["edge_recon_logits"], re_["edge_recon_logits"]) < 1e-10
    if with_cov:
        assert torch.equal(oe["edge_incl"], re_["edge_incl"])
        assert "cat_covariates_contrastive_loss" in rl
    if with_atac:
        assert "chrom_access_recon_loss" in rl
    assert set(ol.keys()) == set(rl.keys())
    # modules/losses.py:80-81,118: the reference accumulates the contrastive loss into a float32
    # ``torch.tensor(0.)`` in place, so its value (and the sums containing it) is rounded to fp32
    fp32_keys = ("cat_covariates_contrastive_loss", "global_loss", "optim_loss") if with_cov else ()
    for k, v in rl.items():
        assert H.rel_err(torch.as_tensor(ol[k]), torch.as_tensor(v)) < (2e-7 if k in fp32_keys else 1e-10), k
    n_checked = 0
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        assert params[k].grad is not None, k
        assert H.rel_err(params[k].grad, p.grad) < 1e-9, k
        n_checked += 1
    assert n_checked >= 10
    m.eval()
    orc.training = False
    mu_r = m.get_latent_representation(node_b, only_active_gps=False)
    mu_o = orc.get_latent_representation(node_b, only_active_gps=False)
    assert H.rel_err(mu_o, mu_r) < 1e-10


@pytest.mark.parametrize("conv_layer", ["gatv2conv", "gcnconv"])
def test_two_layer_encoder_matches_reference(conv_layer):
    """n_layers_encoder=2 (nn/encoders.py:106-108,123-127,191-194): conv_l1 over the sampled sub-graph."""
    ref = load_reference()
    adata = H.tiny_problem(seed=5)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=3, conv_layer=conv_layer, n_layers=2)
    torch.manual_seed(6)
    feat = copy.deepcopy(cfg["features_idx_dict"])
    with quiet():
        m = ref.modules.VGPGAE(
            n_input=cfg["n_input"], n_fc_layers_encoder=1, n_layers_encoder=2,
            n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"], n_addon_gp=3,
            cat_covariates_embeds_nums=[], n_output_genes=cfg["n_output_genes"],
            target_rna_decoder_mask=cfg["target_rna_mask"], source_rna_decoder_mask=cfg["source_rna_mask"],
            features_idx_dict=feat, features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
            conv_layer_encoder=conv_layer, encoder_n_attention_heads=cfg["heads"],
            node_label_method="one-hop-norm", active_gp_thresh_ratio=0.01,
            include_cat_covariates_contrastive_loss=False).double()
    m.train()
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith("bias"):
                p.uniform_(-0.1, 0.1)
    assert any(k.startswith("encoder.conv_l1.") for k in m.state_dict())
    params = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point and
                                                   not k.endswith(("mask", "running_mean_abs_mu")))
              for k, v in m.state_dict().items()}
    orc = VGPGAERef(cfg, params)
    node_b, edge_b = make_batches(adata)
    P = cfg["n_prior_gp"] + 3
    g = torch.Generator().manual_seed(5)
    eps_n = torch.randn(node_b.batch_size, P, generator=g, dtype=torch.float64)
    eps_e = torch.randn(2 * edge_b.edge_label_index.size(1), P, generator=g, dtype=torch.float64)
    draws = [eps_n, eps_e]
    real = torch.randn_like
    torch.randn_like = lambda t, *a, **k: draws.pop(0)
    try:
        rn = m(data_batch=node_b, decoder="omics")
        re_ = m(data_batch=edge_b, decoder="graph")
    finally:
        torch.randn_like = real
    lam = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
               lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_edge_recon=500000.)
    rl = m.loss(edge_model_output=re_, node_model_output=rn, **lam)
    rl["optim_loss"].backward()
    on = orc.forward(node_b, "omics", eps=eps_n)
    oe = orc.forward(edge_b, "graph", eps=eps_e)
    ol = orc.loss(oe, on, **lam)
    ol["optim_loss"].backward()
    for k in ("mu", "logstd", "target_rna_nb_means", "source_rna_nb_means"):
        assert H.rel_err(on[k], rn[k]) < 1e-10, k
    assert H.rel_err(oe["edge_recon_logits"], re_["edge_recon_logits"]) < 1e-10
    for k, v in rl.items():
        assert H.rel_err(torch.as_tensor(ol[k]), torch.as_tensor(v)) < 1e-10, k
    n_l1 = 0
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        assert H.rel_err(params[k].grad, p.grad) < 1e-9, k
        n_l1 += k.startswith("encoder.conv_l1.")
    assert n_l1 >= 2


@pytest.mark.parametrize("conv_layer", ["gatv2conv", "gcnconv"])
def test_two_fc_layers_with_batchnorm_match_reference(conv_layer):
    """n_fc_layers_encoder=2 (nn/encoders.py:93-96,180-182): fc_l1 [F/2, F] -> fc_l2 -> ReLU -> BatchNorm1d.
    Training mode (batch statistics + update of the running buffers), then eval mode (running statistics)."""
    ref = load_reference()
    adata = H.tiny_problem(seed=7)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=3, conv_layer=conv_layer, n_fc_layers=2)
    torch.manual_seed(8)
    feat = copy.deepcopy(cfg["features_idx_dict"])
    with quiet():
        m = ref.modules.VGPGAE(
            n_input=cfg["n_input"], n_fc_layers_encoder=2, n_layers_encoder=1,
            n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"], n_addon_gp=3,
            cat_covariates_embeds_nums=[], n_output_genes=cfg["n_output_genes"],
            target_rna_decoder_mask=cfg["target_rna_mask"], source_rna_decoder_mask=cfg["source_rna_mask"],
            features_idx_dict=feat, features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
            conv_layer_encoder=conv_layer, encoder_n_attention_heads=cfg["heads"],
            node_label_method="one-hop-norm", active_gp_thresh_ratio=0.01,
            include_cat_covariates_contrastive_loss=False).double()
    m.train()
    with torch.no_grad():
        for k, p in m.named_parameters():
            if k.endswith("bias"):
                p.uniform_(-0.1, 0.1)
        m.encoder.fc_l2_bn.weight.uniform_(0.8, 1.2)
    sd = m.state_dict()
    assert sd["encoder.fc_l1.weight"].shape == (cfg["n_input"] // 2, cfg["n_input"])
    assert "encoder.fc_l2_bn.running_var" in sd
    non_param = ("mask", "running_mean_abs_mu", "running_mean", "running_var", "num_batches_tracked")
    params = {k: v.detach().clone().requires_grad_(v.dtype.is_floating_point and not k.endswith(non_param))
              for k, v in sd.items()}
    orc = VGPGAERef(cfg, params)
    node_b, edge_b = make_batches(adata)
    P = cfg["n_prior_gp"] + 3
    g = torch.Generator().manual_seed(5)
    eps_n = torch.randn(node_b.batch_size, P, generator=g, dtype=torch.float64)
    eps_e = torch.randn(2 * edge_b.edge_label_index.size(1), P, generator=g, dtype=torch.float64)
    draws = [eps_n, eps_e]
    real = torch.randn_like
    torch.randn_like = lambda t, *a, **k: draws.pop(0)
    try:
        rn = m(data_batch=node_b, decoder="omics")
        re_ = m(data_batch=edge_b, decoder="graph")
    finally:
        torch.randn_like = real
    lam = dict(lambda_l1_masked=3., l1_targets_mask=None, l1_sources_mask=None, lambda_l1_addon=30.,
               lambda_group_lasso=2., lambda_gene_expr_recon=300., lambda_edge_recon=500000.)
    rl = m.loss(edge_model_output=re_, node_model_output=rn, **lam)
    rl["optim_loss"].backward()
    on = orc.forward(node_b, "omics", eps=eps_n)
    oe = orc.forward(edge_b, "graph", eps=eps_e)
    ol = orc.loss(oe, on, **lam)
    ol["optim_loss"].backward()
    for k in ("mu", "logstd", "target_rna_nb_means", "source_rna_nb_means"):
        assert H.rel_err(on[k], rn[k]) < 1e-10, k
    assert H.rel_err(oe["edge_recon_logits"], re_["edge_recon_logits"]) < 1e-10
    for k, v in rl.items():
        assert H.rel_err(torch.as_tensor(ol[k]), torch.as_tensor(v)) < 1e-10, k
    n_fc = 0
    for k, p in m.named_parameters():
        if p.grad is None:
            continue
        assert H.rel_err(params[k].grad, p.grad) < 1e-9, k
        n_fc += k.startswith(("encoder.fc_l2", "encoder.fc_l1"))
    assert n_fc == 6                                        # fc_l1, fc_l2, fc_l2_bn: weight + bias each
    # the running buffers after two training passes (node batch, edge batch)
    sd2 = m.state_dict()
    for k in ("running_mean", "running_var"):
        assert H.rel_err(params[f"encoder.fc_l2_bn.{k}"], sd2[f"encoder.fc_l2_bn.{k}"]) < 1e-12, k
    assert int(params["encoder.fc_l2_bn.num_batches_tracked"]) == int(sd2["encoder.fc_l2_bn.num_batches_tracked"]) == 2
    # eval mode: running statistics
    m.eval()
    orc.training = False
    with torch.no_grad():
        mu_r = m.get_latent_representation(node_batch=node_b, only_active_gps=False)
    mu_o = orc.get_latent_representation(node_b, only_active_gps=False)
    assert H.rel_err(mu_o, mu_r) < 1e-10


@pytest.mark.parametrize("conv_layer", ["gatv2conv", "gcnconv"])
@pytest.mark.parametrize("n_fc,n_gnn,n_embed,dropout", [(1, 1, 0, 0.), (2, 1, 0, 0.), (1, 2, 3, 0.), (2, 2, 3, 0.2)])
def test_encoder_state_dict_keys_and_shapes_match_reference(conv_layer, n_fc, n_gnn, n_embed, dropout):
    """Checkpoint compatibility of the encoder: every constructor option gives the reference's
    state_dict keys, shapes and dtypes (nn/encoders.py:84-140).  Construction only, no compute."""
    ref = load_reference()
    from nichecompass_b200.nn.encoders import Encoder
    kw = dict(n_input=40, n_cat_covariates_embed_input=n_embed, n_hidden=16, n_latent=12, n_addon_latent=4,
              n_fc_layers=n_fc, conv_layer=conv_layer, n_layers=n_gnn, n_attention_heads=4, dropout_rate=dropout)
    with quiet():
        theirs = ref.nn.Encoder(**kw).state_dict()
        ours = Encoder(**kw).state_dict()
    assert list(ours.keys()) == list(theirs.keys())
    for k, v in theirs.items():
        assert ours[k].shape == v.shape and ours[k].dtype == v.dtype, k


def test_load_and_expand_state_dict_matches_reference():
    """modules/basemodulemixin.py:50-95 (the unmodified reference mixin) against nichecompass_b200's on a module
    that grew a covariate category (embedding rows) and add-on gene programs (weight columns, mask columns)."""
    import torch.nn as nn
    from nichecompass_b200.modules.basemodulemixin import BaseModuleMixin as Ours
    ref = load_reference()
    Theirs = ref.modules.basemodulemixin.BaseModuleMixin

    def make(mixin, n_cat, n_addon, seed):
        class M(nn.Module, mixin):
            def __init__(self):
                super().__init__()
                torch.manual_seed(seed)
                self.embedder = nn.Embedding(n_cat, 3)
                self.addon_l = nn.Linear(n_addon, 5, bias=False)          # weight [5, n_addon]: grows along dim 1
                self.conv = nn.Linear(4, 2 * (6 + n_addon))               # weight + bias grow along dim 0
                self.register_buffer("mask", torch.ones(5, n_addon))
                self.extra = nn.Parameter(torch.randn(2))
        return M()
    small = make(Ours, 2, 3, seed=0).state_dict()
    a, b = make(Ours, 3, 5, seed=1), make(Theirs, 3, 5, seed=1)
    a.load_and_expand_state_dict({k: v.clone() for k, v in small.items()})
    b.load_and_expand_state_dict({k: v.clone() for k, v in small.items()})
    for (ka, va), (kb, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        assert ka == kb and torch.equal(va, vb), ka
    assert torch.equal(a.state_dict()["embedder.weight"][:2], small["embedder.weight"])
    assert torch.equal(a.state_dict()["addon_l.weight"][:, :3], small["addon_l.weight"])
