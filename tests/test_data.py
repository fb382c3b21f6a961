"""Data path: splits (CPU, bit-exact vs the oracle spec) and the GPU sampler (bit-exact batches)."""
import numpy as np
import pytest
import torch

from nichecompass_b200.data import (adjacency_to_edge_index, initialize_dataloaders, prepare_data)
from oracle import sampler_ref as R
from tests import helpers as H


def test_edge_index_matches_oracle():
    adata = H.tiny_problem(n_cells=200, n_genes=20)
    adj = adata.obsp["spatial_connectivities"]
    assert torch.equal(adjacency_to_edge_index(adj), R.adjacency_to_edge_index(adj))


def test_splits_match_oracle_bit_exact():
    adata = H.tiny_problem(n_cells=300, n_genes=20, seed=2)
    n = adata.n_obs
    torch.manual_seed(7)
    d = prepare_data(adata, [], edge_val_ratio=0.1, node_val_ratio=0.1, device="cpu")
    torch.manual_seed(7)
    ei = R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    sp = R.edge_level_split(ei, n, 0.1)
    tm, vm = R.node_level_split_mask(n, 0.1)
    for name, key in (("train", "edge_train_data"), ("val", "edge_val_data")):
        assert torch.equal(d[key].edge_index, sp[name].edge_index)
        assert torch.equal(d[key].edge_label_index, sp[name].edge_label_index)
        assert torch.equal(d[key].edge_label, sp[name].edge_label)
    assert torch.equal(d["node_masked_data"].train_mask, tm)
    assert torch.equal(d["node_masked_data"].val_mask, vm)
    # validation message passing runs on the TRAIN edges
    assert torch.equal(d["edge_val_data"].edge_index, d["edge_train_data"].edge_index)


def test_zero_ratio_split_for_inference():
    adata = H.tiny_problem(n_cells=100, n_genes=10)
    d = prepare_data(adata, [], edge_val_ratio=0., node_val_ratio=0., device="cpu")
    assert d["edge_val_data"].edge_label_index.size(1) == 0
    assert int(d["node_masked_data"].train_mask.sum()) == 100


@pytest.mark.gpu
@pytest.mark.parametrize("k,hops", [(-1, 1), (3, 1), (-1, 2), (2, 2)])
def test_gpu_sampler_matches_oracle_bit_exact(k, hops):
    adata = H.tiny_problem(n_cells=400, n_genes=30, k=5, seed=4)
    n = adata.n_obs
    torch.manual_seed(11)
    d = prepare_data(adata, [], device="cuda")
    loaders = initialize_dataloaders(d["node_masked_data"], d["edge_train_data"], d["edge_val_data"],
                                     edge_batch_size=32, node_batch_size=16, n_direct_neighbors=k, n_hops=hops)
    x = H.dense_x(adata)
    nm = d["node_masked_data"]
    ref_node = R.NodeLoaderRef(x, nm.edge_index, n, nm.train_mask, 16, [k] * hops)
    et = d["edge_train_data"]
    ref_link = R.LinkLoaderRef(x, et.edge_index, n, et.edge_label_index, 32, [k] * hops)

    torch.manual_seed(3)
    got_n = [b for _, b in zip(range(4), loaders["node_train_loader"])]
    torch.manual_seed(3)
    exp_n = [b for _, b in zip(range(4), ref_node)]
    for g, e in zip(got_n, exp_n):
        assert torch.equal(g.n_id.cpu().long(), e.n_id)
        assert torch.equal(g.edge_index.cpu(), e.edge_index)
        assert g.batch_size == e.batch_size
        assert torch.equal(g.x.cpu(), e.x)

    torch.manual_seed(5)
    got_l = [b for _, b in zip(range(4), loaders["edge_train_loader"])]
    torch.manual_seed(5)
    exp_l = [b for _, b in zip(range(4), ref_link)]
    for g, e in zip(got_l, exp_l):
        assert torch.equal(g.n_id.cpu().long(), e.n_id)
        assert torch.equal(g.edge_index.cpu(), e.edge_index)
        assert torch.equal(g.edge_label_index.cpu().long(), e.edge_label_index)
        assert torch.equal(g.edge_label.cpu(), e.edge_label)
        assert g.n_seed == int(e.edge_label_index.max()) + 1


@pytest.mark.gpu
def test_sampler_leaves_workspace_clean_and_handles_last_short_batch():
    adata = H.tiny_problem(n_cells=130, n_genes=10, k=4, seed=1)
    torch.manual_seed(0)
    d = prepare_data(adata, [], device="cuda")
    loaders = initialize_dataloaders(d["node_masked_data"], d["edge_train_data"], None,
                                     edge_batch_size=50, node_batch_size=50)
    sizes = [b.batch_size for b in loaders["node_train_loader"]]
    assert sum(sizes) == int(d["node_masked_data"].train_mask.sum()) and sizes[-1] <= 50
    n_edges = sum(b.edge_label.numel() // 2 for b in loaders["edge_train_loader"])
    assert n_edges == d["edge_train_data"].edge_label_index.size(1)
    for g in (d["node_masked_data"].device_graph(), d["edge_train_data"].device_graph()):
        assert bool((g.slot_ws() == 2 ** 31 - 1).all())


def test_edge_label_adjacency_restricts_the_positive_edges():
    """data/dataprocessors.py:72-78, data/dataloaders.py:86: with ``adata.obsp['edge_label_spatial_connectivities']``
    only the edges that are also in that adjacency are labelled 1 and sampled as positives."""
    import scipy.sparse as sp
    adata = H.tiny_problem(n_cells=150, n_genes=30, k=4)
    adj = sp.csr_matrix(adata.obsp["spatial_connectivities"])
    keep = sp.triu(adj, k=1).tocoo()
    sel = np.arange(keep.nnz) % 3 == 0                                     # a third of the undirected edges
    lab = sp.csr_matrix((np.ones(sel.sum()), (keep.row[sel], keep.col[sel])), shape=adj.shape)
    lab = lab.maximum(lab.T)
    adata.obsp["edge_label_spatial_connectivities"] = lab
    torch.manual_seed(0)
    from nichecompass_b200.data.datasets import SpatialAnnTorchDataset
    from nichecompass_b200.data.dataprocessors import GraphData, edge_level_split
    ds = SpatialAnnTorchDataset(adata, [], counts_key=None, device="cpu")
    assert ds.edge_label_adj is not None
    train, val, _ = edge_level_split(GraphData(ds, ds.edge_index), ds.edge_label_adj, val_ratio=0.1)
    for g in (train, val):
        src, dst = g.edge_label_index[0].numpy().copy(), g.edge_label_index[1].numpy().copy()
        want = np.asarray(lab[src, dst]).reshape(-1) == 1.0
        assert np.array_equal(g.edge_label.numpy().astype(bool), want)
    assert 0 < int(train.edge_label.sum()) < train.edge_label.numel()
