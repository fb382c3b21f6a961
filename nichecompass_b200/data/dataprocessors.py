"""
Edge-level / node-level splits and ``prepare_data`` (reference: data/dataprocessors.py:17-221).

The splits consume the global torch CPU generator exactly like the PyG transforms the reference
calls (RandomLinkSplit then RandomNodeSplit: one ``torch.randperm`` each),
so a seeded run yields the same train / validation partition as the reference.
"""
from typing import List, Optional, Tuple

import scipy.sparse as sp
import numpy as np
import torch

from .datasets import SpatialAnnTorchDataset
from .sampler import DeviceGraph


class GraphData:
    """Host-side description of one split (the role PyG ``Data`` plays in the reference) with a
    lazily built device CSC for the sampler.  ``x`` is replaced by the shared HBM count store."""

    def __init__(self, dataset: SpatialAnnTorchDataset, edge_index: torch.Tensor):
        self.dataset = dataset
        self.store_owner = dataset
        self.edge_index = edge_index
        self.num_nodes = dataset.n_obs
        self.cat_covariates_cats = dataset.cat_covariates_cats
        self.edge_label_index: Optional[torch.Tensor] = None
        self.edge_label: Optional[torch.Tensor] = None
        self.train_mask: Optional[torch.Tensor] = None
        self.val_mask: Optional[torch.Tensor] = None
        self.test_mask: Optional[torch.Tensor] = None
        self._graph = None
        self._cats_dev = None

    @property
    def store(self):
        return self.dataset.store

    @property
    def x(self):
        return self.dataset.x

    @property
    def device(self):
        return self.dataset.device

    def device_graph(self) -> DeviceGraph:
        if self._graph is None:
            self._graph = DeviceGraph(self.edge_index, self.num_nodes, self.device)
        return self._graph

    def cats_device(self):
        if self.cat_covariates_cats is None:
            return None
        if self._cats_dev is None:
            self._cats_dev = self.cat_covariates_cats.to(self.device)
        return self._cats_dev


def edge_level_split(data: GraphData, edge_label_adj=None, val_ratio: float = 0.1, test_ratio: float = 0.,
                     is_undirected: bool = True, neg_sampling_ratio: float = 0.
                     ) -> Tuple[GraphData, GraphData, GraphData]:
    """Train / validation / test ``GraphData`` with: self loops stripped before the split and
    re-appended to the message-passing edges afterwards; one direction (row <= col) of every
    undirected edge permuted and cut into train | val | test supervision edges; message passing
    on the TRAIN edges (both directions) for train and val, train + val edges for test."""
    if not is_undirected or neg_sampling_ratio != 0:
        raise NotImplementedError("only the reference's configuration is implemented "
                                  "(undirected split, loader-side negatives)")
    n = data.num_nodes
    ei = data.edge_index
    ei = ei[:, ei[0] != ei[1]]
    cand = (ei[0] <= ei[1]).nonzero(as_tuple=False).view(-1)
    perm = cand[torch.randperm(cand.size(0))]
    num_val = int(val_ratio * perm.numel())
    num_test = int(test_ratio * perm.numel())
    num_train = perm.numel() - num_val - num_test
    train_e = perm[:num_train]
    val_e = perm[num_train:num_train + num_val]
    test_e = perm[num_train + num_val:]
    loop = torch.arange(n, dtype=torch.int64)
    loops = torch.stack((loop, loop))

    def build(label_e, mp_e):
        e = ei[:, mp_e]
        g = GraphData(data.dataset, torch.cat((e, e.flip([0]), loops), dim=1))
        g.edge_label_index = ei[:, label_e]
        if edge_label_adj is None:
            g.edge_label = torch.ones(label_e.numel())
        else:
            # data/dataprocessors.py:72-78: 1 where the edge is also an edge of ``edge_label_adj``, else 0; the
            # loaders then draw their positives from ``edge_label_index[:, edge_label.bool()]`` only
            # (data/dataloaders.py:86,96).  One vectorised sparse lookup instead of a Python loop over the edges.
            src, dst = g.edge_label_index[0].numpy().copy(), g.edge_label_index[1].numpy().copy()
            if src.size == 0:
                g.edge_label = torch.zeros(0)
            else:
                vals = np.asarray(sp.csr_matrix(edge_label_adj)[src, dst]).reshape(-1)
                g.edge_label = torch.from_numpy((vals == 1.0).astype(np.float32))
        return g

    return build(train_e, train_e), build(val_e, train_e), build(test_e, perm[:num_train + num_val])


def node_level_split_mask(data: GraphData, val_ratio: float = 0.1, test_ratio: float = 0.,
                          split_key: str = "x") -> GraphData:
    """Adds ``train_mask`` / ``val_mask`` / ``test_mask`` (PyG RandomNodeSplit 'train_rest')."""
    n = data.num_nodes
    num_val, num_test = round(n * val_ratio), round(n * test_ratio)
    perm = torch.randperm(n)
    data.val_mask = torch.zeros(n, dtype=torch.bool)
    data.test_mask = torch.zeros(n, dtype=torch.bool)
    data.train_mask = torch.zeros(n, dtype=torch.bool)
    data.val_mask[perm[:num_val]] = True
    data.test_mask[perm[num_val:num_val + num_test]] = True
    data.train_mask[perm[num_val + num_test:]] = True
    return data


def prepare_data(adata,
                 cat_covariates_label_encoders: List[dict],
                 adata_atac=None,
                 counts_key: Optional[str] = "counts",
                 adj_key: str = "spatial_connectivities",
                 cat_covariates_keys: Optional[List[str]] = None,
                 edge_val_ratio: float = 0.1,
                 edge_test_ratio: float = 0.,
                 node_val_ratio: float = 0.1,
                 node_test_ratio: float = 0.,
                 device=None,
                 store_location: str = "device") -> dict:
    """Dictionary with ``edge_train_data``, ``edge_val_data``, ``edge_test_data`` and
    ``node_masked_data`` exactly as the reference's ``prepare_data``; all splits share one HBM
    count store."""
    dataset = SpatialAnnTorchDataset(adata=adata, adata_atac=adata_atac, counts_key=counts_key, adj_key=adj_key,
                                     cat_covariates_keys=cat_covariates_keys,
                                     cat_covariates_label_encoders=cat_covariates_label_encoders, device=device,
                                     store_location=store_location)
    data = GraphData(dataset, dataset.edge_index)
    train, val, test = edge_level_split(data, dataset.edge_label_adj, val_ratio=edge_val_ratio,
                                        test_ratio=edge_test_ratio)
    out = {"edge_train_data": train, "edge_val_data": val, "edge_test_data": test}
    out["node_masked_data"] = node_level_split_mask(data, val_ratio=node_val_ratio, test_ratio=node_test_ratio)
    return out
