"""
GPU parity of the training loop and of the public model API: ``Trainer`` (CUDA) against
``oracle.train_ref.TrainerRef`` (CPU restatement of train/trainer.py:339-454) on identical seeds,
batches and injected reparameterisation noise; ``NicheCompass.train`` / ``get_latent_representation``
/ ``save`` / ``load`` end to end.
"""
import contextlib
import copy
import io
import os

import numpy as np
import pytest
import torch

from oracle import train_ref as TR
from oracle.vgpgae_ref import VGPGAERef, init_params
from tests import helpers as H

pytestmark = pytest.mark.gpu
LAM = dict(lambda_l1_masked=0., lambda_l1_addon=30., lambda_group_lasso=1.5, lambda_gene_expr_recon=300.,
           lambda_edge_recon=500000.)


def _module(cfg, p32, conv="gatv2conv"):
    from nichecompass_b200.modules import VGPGAE
    with contextlib.redirect_stdout(io.StringIO()):
        m = VGPGAE(n_input=cfg["n_input"], n_fc_layers_encoder=cfg.get("n_fc_layers", 1), n_layers_encoder=1,
                   n_hidden_encoder=cfg["n_hidden"], n_prior_gp=cfg["n_prior_gp"], n_addon_gp=cfg["n_addon_gp"],
                   cat_covariates_embeds_nums=[], n_output_genes=cfg["n_output_genes"],
                   target_rna_decoder_mask=cfg["target_rna_mask"], source_rna_decoder_mask=cfg["source_rna_mask"],
                   features_idx_dict=copy.deepcopy(cfg["features_idx_dict"]),
                   features_scale_factors=torch.ones(2 * cfg["n_output_genes"]),
                   conv_layer_encoder=conv, encoder_n_attention_heads=cfg["heads"],
                   node_label_method="one-hop-norm", active_gp_thresh_ratio=0.01,
                   include_cat_covariates_contrastive_loss=False)
    m.load_state_dict({k: v.float() for k, v in p32.items()}, strict=True)
    return m


def _cpu_run(adata, cfg, p32, n_iters, edge_bs, **kw):
    torch.manual_seed(0)
    p64 = {k: v.double().clone() for k, v in p32.items()}
    ld = TR.build_loaders(H.dense_x(adata, torch.float64), adata.obsp["spatial_connectivities"], edge_bs)
    ref = TR.TrainerRef(cfg, p64, ld.edge_loader, ld.node_loader, lr=1e-3, loss_kwargs=LAM, **kw)
    return ref, ref.run(max_iters=n_iters), p64


def _gpu_run(adata, cfg, p32, eps, n_iters, edge_bs, conv="gatv2conv", dtype=torch.float32, **tkw):
    from nichecompass_b200.train import Trainer
    with contextlib.redirect_stdout(io.StringIO()):
        # the CPU loop draws from the global host generator strictly on demand: compare without the
        # sampler's software pipelining (same draws per loader, but interleaved one / two batches earlier)
        tkw.setdefault("pipeline_sampler", False)
        tr = Trainer(adata, _module(cfg, p32, conv), edge_batch_size=edge_bs, seed=0, monitor=False,
                     compute_dtype=dtype, **tkw)
        tr.setup_training(lr=1e-3, **LAM)
    tr.model._eps_queue = [e.clone() for e in eps]
    logs = []
    for i, (eb, nb) in enumerate(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)):
        L = tr.train_iteration(eb, nb)
        logs.append({k: float(v.detach()) for k, v in L.items()})
        if i + 1 == n_iters:
            break
    return tr, logs


def test_training_iterations_with_two_fc_layers_match_cpu_loop():
    """n_fc_layers_encoder=2: three Adam iterations through the Trainer (fc_l2 / BatchNorm parameters in the
    flat arena, running statistics updated by both encoder passes of every iteration)."""
    adata = H.tiny_problem(n_cells=220, n_genes=44, n_prior_gp=12)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4, n_fc_layers=2)
    p32 = init_params(cfg, seed=3)
    ref, ref_logs, p64 = _cpu_run(adata, cfg, p32, 3, 24)
    tr, logs = _gpu_run(adata, cfg, p32, ref.noise_log, 3, 24)
    for it, (a, b) in enumerate(zip(logs, ref_logs)):
        for k in b:
            assert abs(a[k] - b[k]) <= 5e-5 * abs(b[k]) * (1 + it) + 1e-12, (it, k, a[k], b[k])
    sd = tr.model.state_dict()
    assert int(sd["encoder.fc_l2_bn.num_batches_tracked"]) == 6
    for k in ("running_mean", "running_var"):
        # Adam moves an element whose gradient sits at fp32 rounding level by +-lr per step (see below)
        assert H.rel_err(sd[f"encoder.fc_l2_bn.{k}"], p64[f"encoder.fc_l2_bn.{k}"]) < 2e-2, k
    for k in ("encoder.fc_l2.weight", "encoder.fc_l2_bn.weight", "encoder.fc_l1.bias"):
        d = (sd[k].double().cpu() - p64[k].detach()).abs()
        assert float(d.max()) <= 2 * 1e-3 * 3 + 1e-9, k
        assert float((sd[k].double().cpu() - p32[k].double()).abs().max()) > 0, k      # the optimiser moved it


@pytest.mark.parametrize("conv", ["gatv2conv", "gcnconv"])
def test_training_iterations_match_cpu_loop(conv):
    adata = H.tiny_problem(n_cells=220, n_genes=44, n_prior_gp=12)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4, conv_layer=conv)
    p32 = init_params(cfg, seed=3)
    ref, ref_logs, p64 = _cpu_run(adata, cfg, p32, 4, 24)
    tr, logs = _gpu_run(adata, cfg, p32, ref.noise_log, 4, 24, conv)
    assert tr.node_batch_size_ == ref.node_loader.batch_size
    for it, (a, b) in enumerate(zip(logs, ref_logs)):
        for k in b:
            assert abs(a[k] - b[k]) <= 2e-5 * abs(b[k]) * (1 + it) + 1e-12, (it, k, a[k], b[k])
    # parameters after 4 Adam steps (Adam normalises gradients: allow lr-sized slack on near-zero gradients)
    # Adam divides by sqrt(v): an element whose gradient is at fp32 rounding level gets a +-lr step of
    # either sign, so elements are bounded by 2 * lr * steps and the bulk must agree tightly.
    sd = tr.model.state_dict()
    close, total = 0, 0
    for k in TR.trainable_keys(p64):
        d = (sd[k].double().cpu() - p64[k].detach()).abs()
        assert float(d.max()) <= 2 * 1e-3 * 4 + 1e-9, k
        close += int((d <= 1e-5 + 1e-4 * p64[k].detach().abs()).sum())
        total += d.numel()
    assert close / total > 0.93, close / total
    assert H.rel_err(tr.model.running_mean_abs_mu, ref.model.running_mean_abs_mu) < 1e-4


def test_host_resident_store_gives_identical_losses():
    adata = H.tiny_problem(n_cells=220, n_genes=44, n_prior_gp=12)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
    p32 = init_params(cfg, seed=3)
    ref, _, _ = _cpu_run(adata, cfg, p32, 2, 24)
    _, dev_logs = _gpu_run(adata, cfg, p32, ref.noise_log, 2, 24)
    tr_h, host_logs = _gpu_run(adata, cfg, p32, ref.noise_log, 2, 24, store_location="host")
    # host store: the joint encoder input of iteration k + 1 is staged on a side stream during iteration k
    assert len(tr_h._side_streams) == 3
    _, plain_logs = _gpu_run(adata, cfg, p32, ref.noise_log, 2, 24, store_location="host", stage_inputs=False)
    assert plain_logs[0] == host_logs[0]
    for a, b in zip(plain_logs, host_logs):
        for k in a:
            assert abs(a[k] - b[k]) <= 1e-6 * abs(b[k]), (k, a[k], b[k])
    # same kernels, same inputs: the first iteration is bit-identical; later ones may differ in the last
    # ulp because a few gradients (att / bias / scatter-add) are accumulated with float atomics
    assert dev_logs[0] == host_logs[0]
    for a, b in zip(dev_logs, host_logs):
        for k in a:
            assert abs(a[k] - b[k]) <= 1e-6 * abs(b[k]), (k, a[k], b[k])


def test_nichecompass_train_latent_save_load(tmp_path):
    from nichecompass_b200.models import NicheCompass
    adata = H.tiny_problem(n_cells=300, n_genes=48, n_prior_gp=12, lib_mean=80.)
    with contextlib.redirect_stdout(io.StringIO()):
        model = NicheCompass(adata, n_addon_gp=4, active_gp_thresh_ratio=0.01)
        model.train(n_epochs=3, n_epochs_all_gps=1, edge_batch_size=64, lambda_group_lasso=1.0, verbose=True)
    tr = model.trainer
    assert len(tr.epoch_logs["train_global_loss"]) == 3 and len(tr.epoch_logs["val_global_loss"]) == 3
    assert tr.epoch_logs["train_global_loss"][-1] < tr.epoch_logs["train_global_loss"][0]
    assert 0.0 <= tr.val_eval_dict_["auroc_score"] <= 1.0
    lat = adata.obsm["nichecompass_latent"]
    n_active = len(adata.uns["nichecompass_active_gp_names"])
    assert lat.shape == (300, n_active) and lat.dtype == np.float64 and np.isfinite(lat).all()

    # latent inference: tiles of 8192 (product) == reference-sized batches evaluated by the CPU oracle
    sd = {k: v.detach().double().cpu() for k, v in model.model.state_dict().items()}
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
    orc = VGPGAERef(cfg, sd)
    orc.training = False
    x = H.dense_x(adata, torch.float64)
    ei = TR.R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    loader = TR.R.NodeLoaderRef(x, ei, 300, torch.ones(300, dtype=torch.bool), 64, shuffle=False)
    mu_ref = torch.cat([orc.get_latent_representation(b, only_active_gps=True, return_mu_std=True)[0] for b in loader])
    assert H.rel_err(torch.from_numpy(lat), mu_ref) < 2e-5

    # save / load round trip with the reference's file layout; frozen weights after load
    d = os.path.join(tmp_path, "model")
    model.save(d, overwrite=True)
    assert sorted(os.listdir(d)) == ["attr.pkl", "model_params.pt", "var_names.csv"]
    with contextlib.redirect_stdout(io.StringIO()):
        loaded = NicheCompass.load(d, adata=adata, use_cuda=True, unfreeze_addon_gp_weights=True)
    assert loaded.is_trained_
    for k, v in model.model.state_dict().items():
        assert torch.equal(v.cpu(), loaded.model.state_dict()[k].cpu()), k
    flags = {n: p.requires_grad for n, p in loaded.model.named_parameters()}
    assert all(v == ("addon" in n or "theta" in n or "aggregator" in n) for n, v in flags.items())
    lat2 = loaded.get_latent_representation(adata, only_active_gps=True, node_batch_size=64)
    assert np.allclose(lat2, lat, rtol=1e-6, atol=1e-7)


def test_layerwise_inference_equals_loader_walk():
    """models/tiled_inference.py (every node's hidden / x_l rows computed once, spatially ordered
    tiles) against the reference-shaped loader walk, and against the CPU oracle."""
    from nichecompass_b200.models import NicheCompass
    adata = H.tiny_problem(n_cells=500, n_genes=48, n_prior_gp=12, lib_mean=80., n_samples=2)
    with contextlib.redirect_stdout(io.StringIO()):
        model = NicheCompass(adata, n_addon_gp=4, active_gp_thresh_ratio=0.01, cat_covariates_keys=["batch"],
                             cat_covariates_embeds_injection=["encoder", "gene_expr_decoder"],
                             cat_covariates_embeds_nums=[3], cat_covariates_no_edges=[True])
        model.train(n_epochs=2, n_epochs_all_gps=1, edge_batch_size=64, verbose=False)
    kw = dict(adata=adata, cat_covariates_keys=["batch"], only_active_gps=False, dtype=np.float32)
    walk_mu, walk_std = model.get_latent_representation(layerwise=False, return_mu_std=True, **kw)
    for part in ("none", "spatial", "rcm"):
        mu, std = model.get_latent_representation(layerwise=True, partition=part, return_mu_std=True, **kw)
        assert mu.shape == walk_mu.shape == (500, 16)
        assert np.allclose(mu, walk_mu, rtol=1e-5, atol=1e-6), part
        assert np.allclose(std, walk_std, rtol=1e-5, atol=1e-6), part
    only_mu = model.get_latent_representation(layerwise=True, partition="spatial", **kw)
    assert np.allclose(only_mu, walk_mu, rtol=1e-5, atol=1e-6)
    act = model.get_latent_representation(adata=adata, cat_covariates_keys=["batch"], only_active_gps=True,
                                          dtype=np.float16)
    assert act.dtype == np.float16 and act.shape[1] == len(model.get_active_gps())
    assert model.last_inference_stats_["n_halo"] == 0 and model.last_inference_stats_["n_own"] == 500
    # bf16 compute mode: same path, bf16 rounding of the activations
    model.model.set_compute_dtype(torch.bfloat16)
    mu_bf = model.get_latent_representation(layerwise=True, partition="spatial", **kw)
    assert np.abs(mu_bf - walk_mu).max() < 3e-2 * np.abs(walk_mu).max()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_joint_encoder_pass_equals_separate_passes(dtype):
    """Trainer.train_iteration with the node + edge sub-graphs merged into one block-diagonal encoder
    pass (VGPGAE.forward_joint, nc_merge_subgraphs) against the reference's two forwards."""
    adata = H.tiny_problem(n_cells=220, n_genes=44, n_prior_gp=12)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
    p32 = init_params(cfg, seed=3)
    ref, _, _ = _cpu_run(adata, cfg, p32, 3, 24)
    tr_j, joint = _gpu_run(adata, cfg, p32, ref.noise_log, 3, 24, dtype=dtype, joint_encoder=True)
    tr_s, sep = _gpu_run(adata, cfg, p32, ref.noise_log, 3, 24, dtype=dtype, joint_encoder=False)
    tol = 2e-6 if dtype == torch.float32 else 2e-3
    for it, (a, b) in enumerate(zip(joint, sep)):
        for k in b:
            assert abs(a[k] - b[k]) <= tol * abs(b[k]) * (1 + it) + 1e-12, (it, k, a[k], b[k])
    if dtype == torch.float32:
        sj, ss = tr_j.model.state_dict(), tr_s.model.state_dict()
        for k in ss:
            if ss[k].is_floating_point():
                assert float((sj[k] - ss[k]).abs().max()) <= 2 * 1e-3 * 3 + 1e-9, k


def test_pipelined_sampler_yields_the_same_batches():
    """Loaders with the software-pipelined sampler (kernels of batch k+1 / k+2 in flight while batch k is
    finished) against the synchronous ones: identical batch sequences, each loader iterated on its own."""
    from nichecompass_b200.data import initialize_dataloaders, prepare_data
    adata = H.tiny_problem(n_cells=260, n_genes=30, n_prior_gp=8)
    with contextlib.redirect_stdout(io.StringIO()):
        dd = prepare_data(adata=adata, cat_covariates_label_encoders=[], edge_val_ratio=0.1, node_val_ratio=0.1,
                          device=torch.device("cuda"))
    out = {}
    for pipe in (False, True):
        torch.manual_seed(11)
        lds = initialize_dataloaders(dd["node_masked_data"], dd["edge_train_data"], dd["edge_val_data"],
                                     edge_batch_size=16, node_batch_size=8)
        seqs = {}
        for name in ("edge_train_loader", "node_train_loader"):
            lds[name].set_pipeline(pipe)
            seq = []
            for b in lds[name]:
                seq.append((b.n_id.cpu(), b.graph.rowptr.cpu(), b.graph.col.cpu(), b.n_seed, b.graph.max_deg,
                            None if b.edge_label_index is None else b.edge_label_index.cpu()))
            seqs[name] = seq
        out[pipe] = seqs
    for name in out[False]:
        a, b = out[False][name], out[True][name]
        assert len(a) == len(b) and len(a) > 3
        for x, y in zip(a, b):
            assert torch.equal(x[0], y[0]) and torch.equal(x[1], y[1]) and torch.equal(x[2], y[2])
            assert x[3] == y[3] and x[4] == y[4]
            assert (x[5] is None and y[5] is None) or torch.equal(x[5], y[5])


def test_rna_atac_training_iterations_on_rows_wider_than_shared_memory():
    """Config-4 layout through the product path: RNA ++ ATAC count store with 52 060 columns (wider than one CTA
    can stage: the gather takes the scatter kernel, the label aggregation runs in column chunks), both pairs of
    masked decoders, two full training iterations against the CPU loop (losses incl. chrom_access_recon_loss)."""
    from nichecompass_b200.data import add_synthetic_atac, make_synthetic_adata_large
    from nichecompass_b200.models import NicheCompass
    from nichecompass_b200.train import Trainer
    adata = make_synthetic_adata_large(360, 60, 5, 12, seed=0, lib_mean=120.)
    atac = add_synthetic_atac(adata, 52_000, seed=0, lib_mean=400.)
    cfg = H.oracle_cfg_multimodal(adata, atac, n_addon_gp=3)
    p32 = init_params(cfg, seed=5)
    lam = dict(LAM, lambda_chrom_access_recon=100.)
    # CPU loop on the dense [N, F + Fp] matrix
    torch.manual_seed(0)
    x = torch.cat((H.dense_x(adata, torch.float64), torch.tensor(np.asarray(atac.X.todense()), dtype=torch.float64)), 1)
    p64 = {k: v.double().clone() for k, v in p32.items()}
    ld = TR.build_loaders(x, adata.obsp["spatial_connectivities"], 24)
    ref = TR.TrainerRef(cfg, p64, ld.edge_loader, ld.node_loader, lr=1e-3, loss_kwargs=lam)
    ref_logs = ref.run(max_iters=2)
    assert "chrom_access_recon_loss" in ref_logs[0]
    with contextlib.redirect_stdout(io.StringIO()):
        model = NicheCompass(adata, adata_atac=atac, n_addon_gp=3, active_gp_thresh_ratio=0.01)
        missing = model.model.load_state_dict({k: v.float() for k, v in p32.items()}, strict=True)
        tr = Trainer(adata, model.model, adata_atac=atac, edge_batch_size=24, seed=0, monitor=False,
                     compute_dtype=torch.float32, pipeline_sampler=False)
        tr.setup_training(lr=1e-3, **lam)
    assert tr.node_masked_data.store.n_cols == 52_060
    tr.model._eps_queue = [e.clone() for e in ref.noise_log]
    logs = []
    for i, (eb, nb) in enumerate(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)):
        logs.append({k: float(v.detach()) for k, v in tr.train_iteration(eb, nb).items()})
        if i + 1 == 2:
            break
    for it, (a, b) in enumerate(zip(logs, ref_logs)):
        assert set(a) == set(b)
        for k in b:
            assert abs(a[k] - b[k]) <= 2e-5 * abs(b[k]) * (1 + it) + 1e-12, (it, k, a[k], b[k])


def test_query_fine_tuning_with_a_new_covariate_category(tmp_path):
    """BASELINE.json config 5, second half (reference mapping): a model trained on samples 0 / 1 is loaded on a
    query dataset that brings a third sample; ``load`` grows the covariate embedder through
    ``load_and_expand_state_dict`` (models/basemodelmixin.py:144-303, modules/basemodulemixin.py:50-95), freezes
    everything and unfreezes the embedder; fine-tuning then moves the embedder only."""
    from nichecompass_b200.models import NicheCompass
    ref_adata = H.tiny_problem(n_cells=300, n_genes=48, n_prior_gp=12, lib_mean=80., n_samples=2)
    kw = dict(cat_covariates_keys=["batch"], cat_covariates_embeds_injection=["encoder", "gene_expr_decoder"],
              cat_covariates_embeds_nums=[3], cat_covariates_no_edges=[True])
    with contextlib.redirect_stdout(io.StringIO()):
        model = NicheCompass(ref_adata, n_addon_gp=4, active_gp_thresh_ratio=0.01, **kw)
        model.train(n_epochs=2, n_epochs_all_gps=1, edge_batch_size=64, verbose=False)
    d = os.path.join(tmp_path, "reference_model")
    model.save(d, overwrite=True)
    old = {k: v.detach().cpu().clone() for k, v in model.model.state_dict().items()}
    assert old["cat_covariate0_embedder.weight"].shape == (2, 3)

    query = H.tiny_problem(n_cells=240, n_genes=48, n_prior_gp=12, lib_mean=80., n_samples=3, seed=1)
    for store in ("varm", "uns"):                       # same genes, same gene programs as the reference model
        for k, v in getattr(ref_adata, store).items():
            if k.startswith("nichecompass_") and "latent" not in k and "active" not in k:
                getattr(query, store)[k] = v
    assert sorted(query.obs["batch"].unique().tolist()) == ["sample0", "sample1", "sample2"]
    with contextlib.redirect_stdout(io.StringIO()):
        loaded = NicheCompass.load(d, adata=query, use_cuda=True, unfreeze_cat_covariates_embedder_weights=True)
    new = {k: v.detach().cpu().clone() for k, v in loaded.model.state_dict().items()}
    assert loaded.cat_covariates_cats_ == [["sample0", "sample1", "sample2"]]
    assert new["cat_covariate0_embedder.weight"].shape == (3, 3)
    assert torch.equal(new["cat_covariate0_embedder.weight"][:2], old["cat_covariate0_embedder.weight"])
    for k, v in old.items():
        if k != "cat_covariate0_embedder.weight":
            assert torch.equal(new[k], v), k
    flags = {n: p.requires_grad for n, p in loaded.model.named_parameters()}
    assert [n for n, f in flags.items() if f] == ["cat_covariate0_embedder.weight"]
    with contextlib.redirect_stdout(io.StringIO()):
        loaded.train(n_epochs=2, n_epochs_all_gps=0, edge_batch_size=64, verbose=False)
    tuned = {k: v.detach().cpu() for k, v in loaded.model.state_dict().items()}
    for k, v in new.items():
        if k == "cat_covariate0_embedder.weight":
            assert not torch.equal(tuned[k][2], v[2])               # the new sample's embedding learned something
        elif "running_mean_abs_mu" not in k:
            assert torch.equal(tuned[k], v), k                      # frozen
    lat = loaded.get_latent_representation(adata=query, cat_covariates_keys=["batch"], only_active_gps=False)
    assert lat.shape == (240, 16) and np.isfinite(lat).all()


@pytest.mark.parametrize("dtype,use_active,store", [(torch.float32, False, "device"), (torch.bfloat16, False, "device"),
                                                    (torch.float32, True, "device"), (torch.float32, False, "host")])
def test_cuda_graph_iterations_equal_eager_iterations(dtype, use_active, store):
    """The captured-and-replayed iteration (size buckets with padding rows, device-side Adam step counter, noise drawn
    into static buffers) against the eager iteration on the same batches and the same generator state: same losses
    every iteration and the same parameters after eight updates, up to the order of float atomics."""
    from nichecompass_b200.train import Trainer
    adata = H.tiny_problem(n_cells=600, n_genes=44, n_prior_gp=12)
    cfg = H.oracle_cfg_from_adata(adata, n_addon_gp=4)
    p32 = init_params(cfg, seed=3)

    def run(graphs):
        with contextlib.redirect_stdout(io.StringIO()):
            tr = Trainer(adata, _module(cfg, p32), edge_batch_size=32, seed=0, monitor=False, compute_dtype=dtype,
                         cuda_graphs=graphs, store_location=store)
            tr.setup_training(lr=1e-3, n_epochs_all_gps=0 if use_active else 25, **LAM)
        torch.manual_seed(11)
        torch.cuda.manual_seed(11)
        logs = []
        for i, (eb, nb) in enumerate(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)):
            logs.append({k: float(v) for k, v in tr.train_iteration(eb, nb).items()})
            if i + 1 == 8:
                break
        return tr, logs
    tr_e, eager = run(False)
    tr_g, graph = run(True)
    assert tr_e.graph_stats_["replayed"] == 0
    assert tr_g.graph_stats_["replayed"] >= 5 and tr_g.graph_stats_["captured"] >= 1, tr_g.graph_stats_
    tol = 2e-6 if dtype == torch.float32 else 2e-3
    for it, (a, b) in enumerate(zip(graph, eager)):
        for k in b:
            assert abs(a[k] - b[k]) <= tol * abs(b[k]) * (1 + it) + 1e-10, (it, k, a[k], b[k])
    if dtype == torch.float32:
        sd_e, sd_g = tr_e.model.state_dict(), tr_g.model.state_dict()
        for k in sd_e:
            if sd_e[k].is_floating_point():
                assert float((sd_e[k] - sd_g[k]).abs().max()) <= 2e-3 * 8 * 1e-3 + 1e-6 * float(sd_e[k].abs().max()) + 1e-3 * 2 * 8, k
        assert tr_g.optimizer.step_count == tr_e.optimizer.step_count == 8
        assert abs(float(tr_g.optimizer.dyn[1]) - 9.0) < 1e-6


def test_cuda_graph_iterations_with_covariate_embeddings_equal_eager_iterations():
    """Config-3 shape (sample covariate, embedding injected into the decoders, no cross-sample edges) in the captured
    iteration: the categories of the sampled cells are static inputs of the graph."""
    from nichecompass_b200.models import NicheCompass
    from nichecompass_b200.train import Trainer
    adata = H.tiny_problem(n_cells=600, n_genes=44, n_prior_gp=12, n_samples=2)

    def run(graphs):
        with contextlib.redirect_stdout(io.StringIO()):
            model = NicheCompass(adata, n_addon_gp=4, cat_covariates_keys=["batch"], seed=0)
            tr = Trainer(adata, model.model, cat_covariates_keys=["batch"], edge_batch_size=32, seed=0, monitor=False,
                         compute_dtype=torch.float32, cuda_graphs=graphs)
            tr.setup_training(lr=1e-3, **LAM)
        torch.manual_seed(11)
        torch.cuda.manual_seed(11)
        logs = []
        for i, (eb, nb) in enumerate(tr.joint_batches(tr.edge_train_loader, tr.node_train_loader)):
            logs.append({k: float(v) for k, v in tr.train_iteration(eb, nb).items()})
            if i + 1 == 8:
                break
        return tr, logs
    tr_e, eager = run(False)
    tr_g, graph = run(True)
    assert tr_g.model.n_cat_covariates_ == 1 and tr_g.graph_stats_["replayed"] >= 5, tr_g.graph_stats_
    for it, (a, b) in enumerate(zip(graph, eager)):
        for k in b:
            assert abs(a[k] - b[k]) <= 2e-6 * abs(b[k]) * (1 + it) + 1e-10, (it, k, a[k], b[k])
    we = tr_e.model.state_dict()["cat_covariate0_embedder.weight"]
    wg = tr_g.model.state_dict()["cat_covariate0_embedder.weight"]
    assert float((we - wg).abs().max()) <= 2e-3 * 8 * 1e-3 + 1.6e-2 and not torch.equal(wg, torch.zeros_like(wg))
