"""
Device-resident graph + GPU neighbourhood sampler front-end (C ABI: nc_sample_subgraph,
nc_unique_sorted_i32).  Replaces the PyG NeighborLoader / LinkNeighborLoader machinery the
reference builds in data/dataloaders.py:64-100: batches are index tensors into the HBM count
store plus a local dst-CSR, never a materialised dense ``Data`` object.
"""
from typing import List, Optional

import numpy as np
import torch

from .. import ops
from .._lib import call
from ..functional import SubgraphCSR
from .store import CSRFeatureStore

INT32_MAX = 2 ** 31 - 1


class DeviceGraph:
    """CSC (incoming neighbours, sorted by (target, source)) of a message-passing graph on the
    device, built once from a host ``edge_index`` [2, E] int64 (row 0 = source, 1 = target)."""

    def __init__(self, edge_index: torch.Tensor, num_nodes: int, device):
        src = edge_index[0].numpy()
        dst = edge_index[1].numpy()
        order = np.argsort(dst.astype(np.int64) * num_nodes + src.astype(np.int64), kind="stable")
        counts = np.bincount(dst, minlength=num_nodes)
        rowptr = np.zeros(num_nodes + 1, dtype=np.int64)
        np.cumsum(counts, out=rowptr[1:])
        assert rowptr[-1] < INT32_MAX, "graph too large for int32 edge offsets"
        self.num_nodes = int(num_nodes)
        self.num_edges = int(rowptr[-1])
        self.max_deg = int(counts.max()) if num_nodes > 0 else 0
        self.rowptr = torch.from_numpy(rowptr.astype(np.int32)).to(device)
        self.col = torch.from_numpy(src[order].astype(np.int32)).to(device)
        self.e_perm_host = torch.from_numpy(order)            # CSC position -> edge_index position
        self.device = torch.device(device)
        self._slot = None

    def slot_ws(self) -> torch.Tensor:
        if self._slot is None:
            self._slot = torch.empty(max(self.num_nodes, 1), dtype=torch.int32, device=self.device)
            call("nc_fill_i32", self._slot, self._slot.numel(), INT32_MAX)
        return self._slot


class SampledBatch:
    """One sampled mini-batch: ids + local CSR; dense views are materialised only on request
    (compatibility with code written against PyG ``Data`` batches)."""

    def __init__(self, store: CSRFeatureStore, n_id: torch.Tensor, graph: SubgraphCSR, n_seed: int,
                 cats_all: Optional[torch.Tensor] = None):
        self.store, self.n_id, self.graph, self.n_seed = store, n_id, graph, int(n_seed)
        self.cat_covariates_cats = cats_all[n_id.long()] if cats_all is not None else None
        self.batch_size: Optional[int] = None
        self.edge_label_index: Optional[torch.Tensor] = None
        self.edge_label: Optional[torch.Tensor] = None
        self.input_id: Optional[torch.Tensor] = None

    @property
    def num_nodes(self) -> int:
        return self.graph.n_nodes

    @property
    def x(self) -> torch.Tensor:
        return ops.csr_gather_rows(self.store, self.n_id, self.graph.n_nodes, False, torch.float32)

    @property
    def edge_index(self) -> torch.Tensor:
        rp = self.graph.rowptr.long()
        n = self.graph.n_nodes
        dst = torch.repeat_interleave(torch.arange(n, device=rp.device), rp[1:n + 1] - rp[:n])
        return torch.stack((self.graph.col[:dst.numel()].long(), dst))

    def to(self, device):
        return self

    def tensors(self):
        """Every device tensor of the batch (stream bookkeeping of the prefetcher)."""
        out = [self.n_id, self.graph.rowptr, self.graph.col]
        for t in (self.cat_covariates_cats, self.edge_label_index, self.edge_label, self.input_id):
            if t is not None and t.is_cuda:
                out.append(t)
        return out


class _PinnedRing:
    """A few pinned int32 slots for asynchronous read-backs (cudaHostAlloc once, not per batch)."""

    def __init__(self, width: int, slots: int = 16):
        self.buf = torch.empty(slots, width, dtype=torch.int32).pin_memory() if torch.cuda.is_available() \
            else torch.empty(slots, width, dtype=torch.int32)
        self.i = 0

    def next(self) -> torch.Tensor:
        self.i = (self.i + 1) % self.buf.size(0)
        return self.buf[self.i]


_RING = {}


def _readback_async(t: torch.Tensor):
    """Starts the D2H copy of a small int32 device vector on the current stream; returns (pinned view,
    event).  ``event.synchronize()`` then costs nothing when the kernels were launched a step earlier."""
    ring = _RING.get(t.numel())
    if ring is None:
        ring = _RING[t.numel()] = _PinnedRing(t.numel())
    host = ring.next()
    host.copy_(t, non_blocking=True)
    ev = torch.cuda.Event()
    ev.record()
    return host, ev


def unique_sorted_launch(ids: torch.Tensor):
    """First half of ``unique_sorted``: the kernel and an asynchronous read-back of the count."""
    n = ids.numel()
    if n > 16384:
        u, inv = torch.unique(ids, return_inverse=True)          # plumbing for very large link batches
        return ("done", u.to(torch.int32), inv.to(torch.int32), int(u.numel()))
    uniq = torch.empty(n, dtype=torch.int32, device=ids.device)
    inv = torch.empty(n, dtype=torch.int32, device=ids.device)
    cnt = torch.empty(1, dtype=torch.int32, device=ids.device)
    call("nc_unique_sorted_i32", ids, n, uniq, inv, cnt)
    host, ev = _readback_async(cnt)
    return ("pending", uniq, inv, host, ev, ids)


def unique_sorted_finish(h):
    if h[0] == "done":
        return h[1], h[2], h[3]
    _, uniq, inv, host, ev, _ids = h
    ev.synchronize()
    nu = int(host[0])
    return uniq[:nu], inv, nu


def unique_sorted(ids: torch.Tensor):
    """(unique sorted ids, inverse, n_unique) of an int32 device vector -- one kernel, one sync."""
    return unique_sorted_finish(unique_sorted_launch(ids))


def sample_subgraph(g: DeviceGraph, seeds: torch.Tensor, n_seed: int, num_neighbors: List[int],
                    rng_seed: int = 0, batch_id: int = 0):
    """Returns (n_id int32 [n_sub], SubgraphCSR with n_dst = n_seed)."""
    return sample_subgraph_finish(sample_subgraph_launch(g, seeds, n_seed, num_neighbors, rng_seed, batch_id))


def sample_subgraph_launch(g: DeviceGraph, seeds: torch.Tensor, n_seed: int, num_neighbors: List[int],
                           rng_seed: int = 0, batch_id: int = 0):
    """Launches the sampling kernels and the asynchronous read-back of the sub-graph sizes."""
    hops = len(num_neighbors)
    k = int(num_neighbors[0])
    if any(int(v) != k for v in num_neighbors):
        raise NotImplementedError("per-hop neighbour counts must be equal (the reference uses [n] * n_hops)")
    d = max(g.max_deg, 1)
    d_eff = d if k < 0 else min(d, k)
    frontier, cap_nodes, cap_cand = n_seed, n_seed, 1
    for _ in range(hops):
        cand = min(frontier * d_eff, g.num_nodes * d_eff)
        cap_cand = max(cap_cand, cand)
        frontier = min(cand, g.num_nodes)
        cap_nodes = min(cap_nodes + frontier, g.num_nodes)
    cap_nodes = max(cap_nodes, n_seed)
    cap_edges = max(min(cap_nodes * d, g.num_edges), 1)
    dev = g.device
    n_id = torch.empty(cap_nodes, dtype=torch.int32, device=dev)
    rowptr = torch.empty(cap_nodes + 1, dtype=torch.int32, device=dev)
    col = torch.empty(cap_edges, dtype=torch.int32, device=dev)
    ws = torch.empty(2 * cap_nodes + 3 * cap_cand + 2, dtype=torch.int32, device=dev)
    counts = torch.empty(8, dtype=torch.int32, device=dev)
    call("nc_sample_subgraph", g.rowptr, g.col, seeds, n_seed, hops, k, int(rng_seed), int(batch_id),
         g.slot_ws(), n_id, cap_nodes, rowptr, col, cap_edges, ws, cap_cand, counts)
    host, ev = _readback_async(counts)
    return (n_id, rowptr, col, n_seed, host, ev, seeds, ws, counts)


def sample_subgraph_finish(h):
    n_id, rowptr, col, n_seed, host, ev = h[:6]
    ev.synchronize()                                       # the one host wait of a batch
    c = host.tolist()
    n_sub, e_sub, overflow, max_row = int(c[0]), int(c[1]), int(c[2]), int(c[3])
    if overflow:
        raise RuntimeError("nc_sample_subgraph: capacity overflow (graph degree bound violated)")
    graph = SubgraphCSR(rowptr[:n_sub + 1], col[:max(e_sub, 1)][:e_sub] if e_sub > 0 else col[:0], n_sub,
                        n_seed, max_row, e_dst=int(c[4]))
    return n_id[:n_sub], graph
