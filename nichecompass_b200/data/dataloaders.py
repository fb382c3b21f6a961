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
from collections import deque

from .sampler import (SampledBatch, sample_subgraph_finish, sample_subgraph_launch, unique_sorted_finish,
                      unique_sorted_launch)


def _epoch_permutation(n: int, shuffle: bool) -> torch.Tensor:
    """torch DataLoader(shuffle=True) -> RandomSampler: 64-bit seed from the global generator, then
    ``randperm`` with a private generator."""
    if not shuffle:
        return torch.arange(n)
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n, generator=g)


def shard_batch_ids(n_batches_total: int, rank: int, world_size: int) -> List[int]:
    """Batches of the common epoch permutation that rank ``rank`` of ``world_size`` processes: rank, rank + W,
    ... wrapping around so that every rank yields ceil(total / W) batches (the all-reduce needs equal counts)."""
    per_rank = math.ceil(n_batches_total / world_size)
    return [(i * world_size + rank) % n_batches_total for i in range(per_rank)]


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
        self.pipeline = False

    def set_pipeline(self, on: bool):
        """Software pipelining of the sampler: the kernels of batch k + 1 are launched before the sizes
        of batch k are read back, so the read-back never waits for the device (the sampling kernels
        share the GPU with the training step and would otherwise finish late).  The host generators
        are consumed in the same order, one batch earlier."""
        self.pipeline = bool(on)

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
        inflight, depth = deque(), (1 if self.pipeline else 0)
        for b in shard_batch_ids(self.n_batches_total, self.rank, self.world_size):
            batch_id = self._batch_counter + self.rank
            self._batch_counter += self.world_size
            idx = order[b * self.batch_size:(b + 1) * self.batch_size]
            seeds = self._input_dev[idx].contiguous()
            n_seed = seeds.numel()
            inflight.append((sample_subgraph_launch(g, seeds, n_seed, self.num_neighbors, self.rng_seed, batch_id),
                             n_seed, idx))
            if len(inflight) > depth:
                yield self._finish(inflight.popleft(), cats)
        while inflight:
            yield self._finish(inflight.popleft(), cats)

    def _finish(self, rec, cats) -> SampledBatch:
        handle, n_seed, idx = rec
        n_id, graph = sample_subgraph_finish(handle)
        batch = SampledBatch(self.data.store, n_id, graph, n_seed, cats)
        batch.batch_size = n_seed
        batch.input_id = idx
        return batch


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
        self.pipeline = False

    def set_pipeline(self, on: bool):
        """See ``NodeLoader.set_pipeline``.  A link batch has two dependent read-backs (number of unique
        seeds, then the sub-graph sizes), hence three stages: unique(k + 2) | sample(k + 1) | finish(k)."""
        self.pipeline = bool(on)

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
        qa, qb, depth = deque(), deque(), (1 if self.pipeline else 0)
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
            # stage A: unique seeds of batch q
            qa.append((unique_sorted_launch(ids), batch_id, idx, num_pos, num_neg))
            if len(qa) > depth:
                qb.append(self._stage_b(qa.popleft(), g))
            if len(qb) > depth:
                yield self._stage_c(qb.popleft(), cats, dev)
        while qa:
            qb.append(self._stage_b(qa.popleft(), g))
            if len(qb) > depth:
                yield self._stage_c(qb.popleft(), cats, dev)
        while qb:
            yield self._stage_c(qb.popleft(), cats, dev)

    def _stage_b(self, rec, g):
        """Number of unique seeds is known -> launch the neighbourhood sampler."""
        uh, batch_id, idx, num_pos, num_neg = rec
        seeds, inv, n_seed = unique_sorted_finish(uh)
        sh = sample_subgraph_launch(g, seeds.contiguous(), n_seed, self.num_neighbors, self.rng_seed, batch_id)
        return sh, inv, n_seed, idx, num_pos, num_neg

    def _stage_c(self, rec, cats, dev) -> SampledBatch:
        sh, inv, n_seed, idx, num_pos, num_neg = rec
        n_id, graph = sample_subgraph_finish(sh)
        batch = SampledBatch(self.data.store, n_id, graph, n_seed, cats)
        batch.edge_label_index = inv.view(2, -1)
        batch.edge_label = torch.cat((torch.ones(num_pos, device=dev), torch.zeros(num_neg, device=dev)))
        batch.input_id = idx
        return batch


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
