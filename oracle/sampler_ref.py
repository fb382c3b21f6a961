"""
CPU specification of the NicheCompass data path (TEST INFRASTRUCTURE -- see oracle/__init__.py).

Restates, on plain torch CPU tensors:
  * data/datasets.py:74-97   adjacency -> row-major sorted COO edge_index, self loops removed
                             and re-appended LAST
  * data/dataprocessors.py:63-95   edge_level_split  (PyG RandomLinkSplit, is_undirected=True,
                             key="edge_label", neg_sampling_ratio=0, num_test=0)
  * data/dataprocessors.py:125-129 node_level_split_mask (PyG RandomNodeSplit "train_rest")
  * data/dataloaders.py:64-100     NeighborLoader / LinkNeighborLoader(neg_sampling_ratio=1,
                             directed=False)  -- SURVEY.md App. A.5
The PyG loaders themselves are third-party (pyg-lib C++); their semantics are restated, and the
one part that is *not* reproducible outside pyg-lib (uniform neighbour sub-sampling when
``num_neighbors > 0``) is replaced by a documented counter-based rule (``_subsample_keys``)
that the CUDA sampler implements bit-for-bit.

All random draws use the torch global CPU generator in the same order as the reference, so a
seeded run of this file and of the reference-with-shim yield identical batches.
"""
import math
from types import SimpleNamespace
from typing import Iterator, List, Optional

import numpy as np
import torch

MASK64 = (1 << 64) - 1


# --------------------------------------------------------------------------- dataset / splits
def adjacency_to_edge_index(adj_csr, self_loops: bool = True) -> torch.Tensor:
    """data/datasets.py:74-97.  ``adj_csr``: scipy sparse.  Result int64 [2, E]."""
    coo = adj_csr.tocoo()
    row = torch.from_numpy(np.asarray(coo.row, dtype=np.int64))
    col = torch.from_numpy(np.asarray(coo.col, dtype=np.int64))
    n = adj_csr.shape[0]
    order = torch.argsort(row * n + col, stable=True)       # coalesced COO order
    ei = torch.stack([row[order], col[order]])
    # (the reference's symmetry check compares nnz with nnz of the transpose: vacuous)
    if self_loops:
        ei = ei[:, ei[0] != ei[1]]
        loop = torch.arange(n, dtype=torch.int64)
        ei = torch.cat([ei, torch.stack([loop, loop])], dim=1)
    return ei


def edge_level_split(edge_index: torch.Tensor, num_nodes: int, val_ratio: float):
    """data/dataprocessors.py:63-95 with RandomLinkSplit semantics (App. A.5).
    Consumes ONE ``torch.randperm`` from the global generator.
    Returns dict(train=..., val=...) each with ``edge_index`` (message passing, self loops
    re-appended) and ``edge_label_index`` / ``edge_label``."""
    ei = edge_index[:, edge_index[0] != edge_index[1]]
    cand = (ei[0] <= ei[1]).nonzero(as_tuple=False).view(-1)
    perm = cand[torch.randperm(cand.size(0))]
    num_val = int(val_ratio * perm.numel())
    num_train = perm.numel() - num_val
    train_e, val_e = perm[:num_train], perm[num_train:num_train + num_val]
    loop = torch.arange(num_nodes, dtype=torch.int64)
    loops = torch.stack([loop, loop])

    def mp(idx):
        e = ei[:, idx]
        return torch.cat([e, e.flip([0]), loops], dim=1)

    out = {}
    for name, lab in (("train", train_e), ("val", val_e)):
        out[name] = SimpleNamespace(
            edge_index=mp(train_e),                       # val also passes messages on TRAIN edges
            edge_label_index=ei[:, lab],
            edge_label=torch.ones(lab.numel()),
            num_nodes=num_nodes)
    return out


def node_level_split_mask(num_nodes: int, val_ratio: float):
    """data/dataprocessors.py:125-129 (RandomNodeSplit, split='train_rest').
    Consumes ONE ``torch.randperm(num_nodes)``."""
    num_val = round(num_nodes * val_ratio)
    perm = torch.randperm(num_nodes)
    val_mask = torch.zeros(num_nodes, dtype=torch.bool)
    train_mask = torch.zeros(num_nodes, dtype=torch.bool)
    val_mask[perm[:num_val]] = True
    train_mask[perm[num_val:]] = True
    return train_mask, val_mask


# --------------------------------------------------------------------------- CSC + sampling
def to_csc(edge_index: torch.Tensor, num_nodes: int):
    """Incoming-neighbour lists: sorted by (target, source).  Returns colptr [n+1], row [E]
    (source ids) and perm (position in ``edge_index`` of each CSC entry == e_id)."""
    src, dst = edge_index[0], edge_index[1]
    perm = torch.argsort(dst * num_nodes + src, stable=True)
    row = src[perm]
    counts = torch.bincount(dst, minlength=num_nodes)
    colptr = torch.zeros(num_nodes + 1, dtype=torch.int64)
    colptr[1:] = torch.cumsum(counts, 0)
    return colptr, row, perm


def _mix64(x: int) -> int:
    """splitmix64 finaliser (the CUDA sampler uses the identical integer recipe)."""
    x &= MASK64
    x ^= x >> 30
    x = (x * 0xBF58476D1CE4E5B9) & MASK64
    x ^= x >> 27
    x = (x * 0x94D049BB133111EB) & MASK64
    x ^= x >> 31
    return x


def _subsample_keys(seed: int, batch_id: int, node: int, deg: int) -> List[int]:
    """Key of neighbour slot t of ``node``: mix64(mix64(seed ^ batch_id*C1) ^ node*C2 ^ t)."""
    base = _mix64((seed ^ (batch_id * 0x9E3779B97F4A7C15)) & MASK64)
    nb = (node * 0xD1B54A32D192ED03) & MASK64
    return [_mix64(base ^ nb ^ t) for t in range(deg)]


def neighbor_sample(colptr, row, perm, seeds: torch.Tensor, num_neighbors: List[int],
                    rng_seed: int = 0, batch_id: int = 0):
    """pyg-lib ``neighbor_sample`` + induced ``subgraph`` (directed=False), homogeneous,
    non-disjoint (App. A.5).  Local node order: seeds in given order, then newly discovered
    in-neighbours in discovery order.  ``num_neighbors[h] < 0`` keeps all; otherwise the
    ``k`` slots with the smallest ``_subsample_keys`` are kept (in CSC order).
    Returns n_id, local edge_index [2, E_sub] (grouped by local target), e_id."""
    colptr_l, row_l = colptr.tolist(), row.tolist()
    n_id: List[int] = []
    local = {}
    for s in seeds.tolist():
        if s not in local:              # seeds are unique in both loaders
            local[s] = len(n_id)
            n_id.append(s)
    begin, end = 0, len(n_id)
    for k in num_neighbors:
        for li in range(begin, end):
            v = n_id[li]
            lo, hi = colptr_l[v], colptr_l[v + 1]
            deg = hi - lo
            slots = range(deg)
            if 0 <= k < deg:
                keys = _subsample_keys(rng_seed, batch_id, v, deg)
                slots = sorted(sorted(range(deg), key=lambda t: (keys[t], t))[:k])
            for t in slots:
                u = row_l[lo + t]
                if u not in local:
                    local[u] = len(n_id)
                    n_id.append(u)
        begin, end = end, len(n_id)
    # induced subgraph: every edge of the full graph with both endpoints sampled
    src_l, dst_l, e_id = [], [], []
    perm_l = perm.tolist()
    for li, v in enumerate(n_id):
        for p in range(colptr_l[v], colptr_l[v + 1]):
            u = row_l[p]
            lu = local.get(u)
            if lu is not None:
                src_l.append(lu)
                dst_l.append(li)
                e_id.append(perm_l[p])
    return (torch.tensor(n_id, dtype=torch.int64),
            torch.tensor([src_l, dst_l], dtype=torch.int64).view(2, -1),
            torch.tensor(e_id, dtype=torch.int64))


# --------------------------------------------------------------------------- loaders
def _epoch_permutation(n: int, shuffle: bool) -> torch.Tensor:
    """torch DataLoader(shuffle=True) -> RandomSampler: draw a 64-bit seed from the global
    generator, then randperm with a private generator."""
    if not shuffle:
        return torch.arange(n)
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n, generator=g)


class NodeLoaderRef:
    """PyG NeighborLoader(data, num_neighbors=[n]*hops, batch_size, directed=False, shuffle,
    input_nodes=mask) -- data/dataloaders.py:64-79."""

    def __init__(self, x, edge_index, num_nodes, input_mask, batch_size, num_neighbors=(-1,),
                 shuffle=True, cat_covariates_cats=None, rng_seed=0):
        self.x, self.cats = x, cat_covariates_cats
        self.colptr, self.row, self.perm = to_csc(edge_index, num_nodes)
        self.input_nodes = input_mask.nonzero(as_tuple=False).view(-1)
        self.batch_size, self.num_neighbors = batch_size, list(num_neighbors)
        self.shuffle, self.rng_seed = shuffle, rng_seed
        self._batch_counter = 0

    def __len__(self):
        return math.ceil(self.input_nodes.numel() / self.batch_size)

    def __iter__(self) -> Iterator[SimpleNamespace]:
        order = _epoch_permutation(self.input_nodes.numel(), self.shuffle)
        for b in range(len(self)):
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            seeds = self.input_nodes[idx]
            n_id, ei, e_id = neighbor_sample(self.colptr, self.row, self.perm, seeds,
                                             self.num_neighbors, self.rng_seed,
                                             self._batch_counter)
            self._batch_counter += 1
            batch = SimpleNamespace(x=self.x[n_id], edge_index=ei, n_id=n_id, e_id=e_id,
                                    batch_size=seeds.numel(), input_id=idx)
            if self.cats is not None:
                batch.cat_covariates_cats = self.cats[n_id]
            yield batch


class LinkLoaderRef:
    """PyG LinkNeighborLoader(data, num_neighbors, batch_size, edge_label=None,
    edge_label_index=pos, directed=False, shuffle, neg_sampling_ratio=1.0)
    -- data/dataloaders.py:82-100."""

    def __init__(self, x, edge_index, num_nodes, edge_label_index, batch_size,
                 num_neighbors=(-1,), shuffle=True, neg_sampling_ratio=1.0,
                 cat_covariates_cats=None, rng_seed=0):
        self.x, self.cats, self.num_nodes = x, cat_covariates_cats, num_nodes
        self.colptr, self.row, self.perm = to_csc(edge_index, num_nodes)
        self.pos = edge_label_index
        self.batch_size, self.num_neighbors = batch_size, list(num_neighbors)
        self.shuffle, self.ratio, self.rng_seed = shuffle, neg_sampling_ratio, rng_seed
        self._batch_counter = 0

    def __len__(self):
        return math.ceil(self.pos.size(1) / self.batch_size)

    def __iter__(self) -> Iterator[SimpleNamespace]:
        order = _epoch_permutation(self.pos.size(1), self.shuffle)
        for b in range(len(self)):
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            src, dst = self.pos[0, idx], self.pos[1, idx]
            num_pos = src.numel()
            num_neg = math.ceil(num_pos * self.ratio)
            src_neg = torch.randint(self.num_nodes, (num_neg,))
            dst_neg = torch.randint(self.num_nodes, (num_neg,))
            src = torch.cat([src, src_neg])
            dst = torch.cat([dst, dst_neg])
            edge_label = torch.cat([torch.ones(num_pos), torch.zeros(num_neg)])
            seed, inv = torch.cat([src, dst]).unique(return_inverse=True)   # sorted ascending
            n_id, ei, e_id = neighbor_sample(self.colptr, self.row, self.perm, seed,
                                             self.num_neighbors, self.rng_seed,
                                             self._batch_counter)
            self._batch_counter += 1
            batch = SimpleNamespace(x=self.x[n_id], edge_index=ei, n_id=n_id, e_id=e_id,
                                    edge_label_index=inv.view(2, -1), edge_label=edge_label,
                                    input_id=idx)
            if self.cats is not None:
                batch.cat_covariates_cats = self.cats[n_id]
            yield batch


def auto_node_batch_size(edge_batch_size: int, n_edges_train: int, n_nodes_train: int) -> int:
    """train/trainer.py:194-196."""
    return int(edge_batch_size / math.floor(n_edges_train / n_nodes_train))


def cycle_iterable(iterable):
    """train/utils.py:219-225."""
    it = iter(iterable)
    while True:
        try:
            yield next(it)
        except StopIteration:
            it = iter(iterable)
