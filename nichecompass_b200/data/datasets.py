"""
Dataset extraction from an AnnData(-like) object (reference: data/datasets.py:16-117).

Same inputs, same ``edge_index`` construction (row-major sorted COO of the adjacency, self loops
removed and re-appended last) and categorical covariate encoding.  The one structural change:
counts are NOT densified on the host (``torch.tensor(x.toarray())``, data/datasets.py:59-71);
they are uploaded once as a CSR store in HBM (``self.store``).  ``self.x`` stays available as a
lazily densified view for small data / compatibility.
"""
from typing import List, Optional

import numpy as np
import scipy.sparse as sp
import torch

from .store import CSRFeatureStore
from .utils import encode_labels


def adjacency_to_edge_index(adj, self_loops: bool = True) -> torch.Tensor:
    """scipy sparse adjacency -> int64 [2, E]: entries in (row, col) order; with ``self_loops``
    existing diagonal entries are dropped and one (v, v) per node appended at the end."""
    coo = sp.coo_matrix(adj)
    n = coo.shape[0]
    row = coo.row.astype(np.int64)
    col = coo.col.astype(np.int64)
    order = np.argsort(row * n + col, kind="stable")
    row, col = row[order], col[order]
    if self_loops:
        keep = row != col
        loop = np.arange(n, dtype=np.int64)
        row = np.concatenate((row[keep], loop))
        col = np.concatenate((col[keep], loop))
    return torch.from_numpy(np.stack((row, col)))


class SpatialAnnTorchDataset:
    def __init__(self,
                 adata,
                 cat_covariates_label_encoders: List[dict],
                 adata_atac=None,
                 counts_key: Optional[str] = "counts",
                 adj_key: str = "spatial_connectivities",
                 edge_label_adj_key: str = "edge_label_spatial_connectivities",
                 self_loops: bool = True,
                 cat_covariates_keys: Optional[List[str]] = None,
                 device: Optional[torch.device] = None,
                 store_location: str = "device"):
        device = torch.device(device) if device is not None else torch.device(
            "cuda" if torch.cuda.is_available() else "cpu")
        x = adata.X if counts_key is None else adata.layers[counts_key]
        self._x_rna, self._x_atac = x, (adata_atac.X if adata_atac is not None else None)
        self.n_obs = x.shape[0]
        self.n_node_features = x.shape[1] + (self._x_atac.shape[1] if self._x_atac is not None else 0)
        self.device = device
        self.store_location = store_location
        self._store = None
        adj = adata.obsp[adj_key]
        adj = adj if sp.issparse(adj) else sp.csr_matrix(adj)
        # data/datasets.py:79-83: an optional second adjacency restricting the edges that serve as positives
        self.edge_label_adj = sp.csr_matrix(adata.obsp[edge_label_adj_key]) if edge_label_adj_key in adata.obsp else None
        # The reference's symmetry check compares nnz with nnz of the transpose (data/datasets.py:87-88),
        # which always holds: an asymmetric adjacency is accepted there, and therefore here too.
        if adj.nnz != adj.T.nnz:
            raise ImportError("The input adjacency matrix has to be symmetric.")
        self.edge_index = adjacency_to_edge_index(adj, self_loops=self_loops)
        if cat_covariates_keys is not None:
            cols = [torch.tensor(encode_labels(adata, enc, key), dtype=torch.long)
                    for key, enc in zip(cat_covariates_keys, cat_covariates_label_encoders)]
            self.cat_covariates_cats = torch.stack(cols, dim=1)
        else:
            self.cat_covariates_cats = None

    @property
    def store(self) -> CSRFeatureStore:
        if self._store is None:
            self._store = CSRFeatureStore.from_scipy(self._x_rna, self.device, self._x_atac, self.store_location)
        return self._store

    @property
    def x(self) -> torch.Tensor:
        """Dense host copy (compatibility only; never used by the training path)."""
        xs = [np.asarray(sp.csr_matrix(m).todense()) for m in (self._x_rna, self._x_atac) if m is not None]
        return torch.tensor(np.concatenate(xs, axis=1), dtype=torch.float32)

    def __len__(self):
        return self.n_obs
