"""
Edge-level and node-level mini-batch loaders (reference: data/dataloaders.py:11-102).

``initialize_dataloaders`` keeps the reference signature and dictionary keys.  The loaders follow
the PyG semantics the reference relies on -- shuffling through a private
generator seeded from the global torch CPU generator, two ``torch.randint`` draws of negatives
per link batch, sorted-unique seeds, seeds-first local ordering, induced subgraph -- but sampling
runs on the GPU (``nc_sample_subgraph``) and a batch carries ids into the HBM count store instead
of dense rows.
"""
import math
from typing import Iterator, List, Optional

import torch

from .dataprocessors import GraphData
from .sampler import SampledBatch, sample_subgraph, unique_sorted


def _epoch_permutation(n: int, shuffle: bool) -> torch.Tensor:
    """torch DataLoader(shuffle=True) -> RandomSampler: 64-bit seed from the global generator, then
    ``randperm`` with a private generator."""
    if not shuffle:
        return torch.arange(n)
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n, generator=g)


class NodeLoader:
    """NeighborLoader(data, num_neighbors=[n] * hops, batch_size, directed=False, shuffle,
    input_nodes=mask)."""

    def __init__(self, data: GraphData, input_mask: torch.Tensor, batch_size: int, num_neighbors: List[int],
                 shuffle: bool = True, rng_seed: int = 0):
        self.data = data
        self.input_nodes = input_mask.nonzero(as_tuple=False).view(-1)
        self.batch_size = int(batch_size)
        self.num_neighbors = list(num_neighbors)
        self.shuffle = shuffle
        self.rng_seed = rng_seed
        self._batch_counter = 0
        self._input_dev = None
        self.rank, self.world_size = 0, 1

    def set_shard(self, rank: int, world_size: int):
        """Data parallelism: this process yields batches rank, rank + W, ... of the common
        permutation (wrapping around so that every rank yields the same number of batches)."""
        self.rank, self.world_size = int(rank), int(world_size)

    @property
    def n_batches_total(self) -> int:
        return math.ceil(self.input_nodes.numel() / self.batch_size)

    def __len__(self) -> int:
        return math.ceil(self.n_batches_total / self.world_size)

    def __iter__(self) -> Iterator[SampledBatch]:
        g = self.data.device_graph()
        dev = g.device
        if self._input_dev is None:
            self._input_dev = self.input_nodes.to(dev, torch.int32)
        order = _epoch_permutation(self.input_nodes.numel(), self.shuffle).to(dev)
        cats = self.data.cats_device()
        for i in range(len(self)):
            b = (i * self.world_size + self.rank) % self.n_batches_total
            batch_id = self._batch_counter + self.rank
            self._batch_counter += self.world_size
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            seeds = self._input_dev[idx].contiguous()
            n_seed = seeds.numel()
            n_id, graph = sample_subgraph(g, seeds, n_seed, self.num_neighbors, self.rng_seed, batch_id)
            batch = SampledBatch(self.data.store, n_id, graph, n_seed, cats)
            batch.batch_size = n_seed
            batch.input_id = idx
            yield batch


class LinkLoader:
    """LinkNeighborLoader(data, num_neighbors, batch_size, edge_label=None, edge_label_index=pos,
    directed=False, shuffle, neg_sampling_ratio)."""

    def __init__(self, data: GraphData, edge_label_index: torch.Tensor, batch_size: int,
                 num_neighbors: List[int], shuffle: bool = True, neg_sampling_ratio: float = 1.0,
                 rng_seed: int = 0):
        self.data = data
        self.pos = edge_label_index
        self.batch_size = int(batch_size)
        self.num_neighbors = list(num_neighbors)
        self.shuffle = shuffle
        self.ratio = neg_sampling_ratio
        self.rng_seed = rng_seed
        self._batch_counter = 0
        self._pos_dev = None
        self.rank, self.world_size = 0, 1

    def set_shard(self, rank: int, world_size: int):
        """See ``NodeLoader.set_shard``.  Every rank draws the negatives of ALL batches so that the
        host generators of the ranks stay in lock-step."""
        self.rank, self.world_size = int(rank), int(world_size)

    @property
    def n_batches_total(self) -> int:
        return math.ceil(self.pos.size(1) / self.batch_size)

    def __len__(self) -> int:
        return math.ceil(self.n_batches_total / self.world_size)

    def __iter__(self) -> Iterator[SampledBatch]:
        g = self.data.device_graph()
        dev = g.device
        if self._pos_dev is None:
            self._pos_dev = self.pos.to(dev, torch.int32)
        order = _epoch_permutation(self.pos.size(1), self.shuffle).to(dev)
        cats = self.data.cats_device()
        n_nodes = self.data.num_nodes
        n_total = self.n_batches_total
        n_round = len(self) * self.world_size
        for q in range(n_round):
            b = q % n_total
            n_pos_b = min(self.batch_size, self.pos.size(1) - b * self.batch_size)
            num_neg = math.ceil(n_pos_b * self.ratio)
            # same two draws, same order, as PyG's binary negative sampling
            src_neg = torch.randint(n_nodes, (num_neg,))
            dst_neg = torch.randint(n_nodes, (num_neg,))
            batch_id = self._batch_counter
            self._batch_counter += 1
            if q % self.world_size != self.rank:
                continue
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            pos = self._pos_dev[:, idx]
            num_pos = pos.size(1)
            neg = torch.stack((src_neg, dst_neg)).to(torch.int32)
            if dev.type == "cuda":
                neg = neg.pin_memory().to(dev, non_blocking=True)
            ids = torch.cat((pos[0], neg[0], pos[1], neg[1])).contiguous()
            seeds, inv, n_seed = unique_sorted(ids)
            n_id, graph = sample_subgraph(g, seeds.contiguous(), n_seed, self.num_neighbors, self.rng_seed,
                                          batch_id)
            batch = SampledBatch(self.data.store, n_id, graph, n_seed, cats)
            batch.edge_label_index = inv.view(2, -1)
            batch.edge_label = torch.cat((torch.ones(num_pos, device=dev), torch.zeros(num_neg, device=dev)))
            batch.input_id = idx
            yield batch


def initialize_dataloaders(node_masked_data: GraphData,
                           edge_train_data: Optional[GraphData] = None,
                           edge_val_data: Optional[GraphData] = None,
                           edge_batch_size: Optional[int] = 64,
                           node_batch_size: int = 64,
                           n_direct_neighbors: int = -1,
                           n_hops: int = 1,
                           shuffle: bool = True,
                           edges_directed: bool = False,
                           neg_edge_sampling_ratio: float = 1.) -> dict:
    """Same keys as the reference: ``node_train_loader`` (+ ``node_val_loader`` if the validation
    mask is non-empty), ``edge_train_loader`` / ``edge_val_loader`` when edge data is given and,
    for validation, non-empty."""
    if edges_directed:
        raise NotImplementedError("directed sampling is not used by the reference trainer")
    nn_ = [n_direct_neighbors] * n_hops
    out = {"node_train_loader": NodeLoader(node_masked_data, node_masked_data.train_mask, node_batch_size, nn_,
                                           shuffle)}
    if node_masked_data.val_mask.sum() != 0:
        out["node_val_loader"] = NodeLoader(node_masked_data, node_masked_data.val_mask, node_batch_size, nn_,
                                            shuffle)
    if edge_train_data is not None:
        eli = edge_train_data.edge_label_index[:, edge_train_data.edge_label.bool()]
        out["edge_train_loader"] = LinkLoader(edge_train_data, eli, edge_batch_size, nn_, shuffle,
                                              neg_edge_sampling_ratio)
    if edge_val_data is not None and edge_val_data.edge_label.sum() != 0:
        eli = edge_val_data.edge_label_index[:, edge_val_data.edge_label.bool()]
        out["edge_val_loader"] = LinkLoader(edge_val_data, eli, edge_batch_size, nn_, shuffle,
                                            neg_edge_sampling_ratio)
    return out
