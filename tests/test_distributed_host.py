"""
World-size-2 host logic on CPU (gloo): the spatial-tile partition, the rank-local graphs and the halo
exchange of the layer-wise inference path (models/tiled_inference.py), the mean all-reduce of the
flat gradient buffer.  No CUDA calls.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from nichecompass_b200.data.synthetic import knn_adjacency
from nichecompass_b200.models import tiled_inference as TI


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _graph(n=600, k=5, seed=0):
    rng = np.random.default_rng(seed)
    coords = rng.random((n, 2))
    adj = knn_adjacency(coords, k).tocsr()
    adj.setdiag(1.0)                                    # the dataset appends self loops
    adj.sort_indices()
    csc = adj.tocsc()                                   # incoming neighbours per target
    csc.sort_indices()
    return coords, adj, torch.from_numpy(csc.indptr.astype(np.int64)), torch.from_numpy(csc.indices.astype(np.int64))


def test_morton_order_is_a_permutation_with_compact_tiles():
    coords, adj, rowptr, col = _graph()
    perm = TI.morton_order(coords)
    assert sorted(perm.tolist()) == list(range(len(coords)))
    # contiguous quarters of the curve are compact: far fewer cut edges than random quarters
    bounds = TI.tile_bounds(len(coords), 4)
    owner = np.empty(len(coords), dtype=np.int64)
    for r in range(4):
        owner[perm[bounds[r]:bounds[r + 1]]] = r
    coo = adj.tocoo()
    cut = int((owner[coo.row] != owner[coo.col]).sum())
    rnd = np.random.default_rng(1).integers(0, 4, len(coords))
    cut_rnd = int((rnd[coo.row] != rnd[coo.col]).sum())
    assert cut < 0.35 * cut_rnd, (cut, cut_rnd)
    assert TI.tile_bounds(10, 3) == [0, 3, 6, 10]


def test_partition_order_methods():
    coords, adj, _, _ = _graph(200)
    for m in ("spatial", "rcm", "none", "auto"):
        p = TI.partition_order(200, coords, adj, m)
        assert sorted(p.tolist()) == list(range(200)), m
    assert np.array_equal(TI.partition_order(5, None, None, "auto"), np.arange(5))
    with pytest.raises(ValueError):
        TI.partition_order(5, None, None, "spatial")


def test_local_graph_single_rank_is_the_permuted_graph():
    coords, adj, rowptr, col = _graph(150)
    perm = torch.from_numpy(TI.morton_order(coords))
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(perm.numel())
    lg = TI.build_local_graph(rowptr, col, perm, inv, 0, perm.numel())
    assert lg.n_halo == 0 and lg.n_own == 150
    for i_new in range(150):
        nbr_new = lg.col[lg.rowptr[i_new]:lg.rowptr[i_new + 1]].long()
        i_old = int(perm[i_new])
        expect = col[rowptr[i_old]:rowptr[i_old + 1]]
        assert torch.equal(perm[nbr_new], expect)               # entry order of the row is kept
    assert lg.max_deg == int((rowptr[1:] - rowptr[:-1]).max())


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        coords, adj, rowptr, col = _graph()
        n = len(coords)
        perm = torch.from_numpy(TI.partition_order(n, coords, adj, "auto"))
        inv = torch.empty_like(perm)
        inv[perm] = torch.arange(n)
        bounds = TI.tile_bounds(n, world)
        lo, hi = bounds[rank], bounds[rank + 1]
        lg = TI.build_local_graph(rowptr, col, perm, inv, lo, hi)
        feats = torch.arange(n, dtype=torch.float32)[:, None] * torch.tensor([1.0, 0.5, -2.0])    # row i = f(i)
        own_rows = feats[perm[lo:hi]]
        halo = TI.exchange_halo_rows(own_rows, lg, bounds)
        rows = torch.cat((own_rows, halo))
        ok = True
        for i_loc in range(lg.n_own):
            nbr = lg.col[lg.rowptr[i_loc]:lg.rowptr[i_loc + 1]].long()
            i_old = int(perm[lo + i_loc])
            expect_old = col[rowptr[i_old]:rowptr[i_old + 1]]
            ok &= bool(torch.equal(rows[nbr], feats[expect_old]))
        # the mean all-reduce of FlatAdam on a CPU buffer
        from nichecompass_b200.train.optim import FlatAdam
        fa = FlatAdam.__new__(FlatAdam)
        fa.flat_g = torch.full((7,), float(rank + 1))
        fa.all_reduce_grads()
        ok &= bool(torch.allclose(fa.flat_g, torch.full((7,), (1 + world) / 2)))
        ret[rank] = (ok, lg.n_own, lg.n_halo)
    finally:
        dist.destroy_process_group()


def test_halo_exchange_and_grad_allreduce_world2():
    world = 2
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), ret), nprocs=world, join=True)
    assert len(ret) == world
    n_halo = 0
    for r in range(world):
        ok, n_own, nh = ret[r]
        assert ok, f"rank {r}: halo rows differ from the owners' rows"
        assert n_own == 300
        n_halo += nh
    assert 0 < n_halo < 200          # boundary only: the tiles are spatially compact


def test_loader_shards_cover_every_batch_with_equal_counts():
    """Data-parallel training: rank r takes batches r, r + W, ... of the common permutation; the
    union over the ranks covers every batch, every rank yields the same number (the gradient all-reduce is
    collective), and the link loader's round arithmetic (negatives drawn for ALL batches on every rank so the
    host generators stay in lock-step; dataloaders.py LinkLoader.__iter__) selects the same batches."""
    from nichecompass_b200.data.dataloaders import shard_batch_ids
    import math
    for total in (1, 2, 7, 16, 125, 16218):
        for world in (1, 2, 3, 4, 8):
            shards = [shard_batch_ids(total, r, world) for r in range(world)]
            per = math.ceil(total / world)
            assert all(len(s) == per for s in shards)
            assert set(b for s in shards for b in s) == set(range(total))
            if total % world == 0:
                assert sorted(b for s in shards for b in s) == list(range(total))       # no batch twice
            for r in range(world):                                                        # link-loader rounds
                mine = [q % total for q in range(per * world) if q % world == r]
                assert mine == shards[r]
