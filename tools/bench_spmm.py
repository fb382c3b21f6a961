"""One-hop aggregator SpMM at scale (BASELINE.json metric 2: SpMM HBM GB/s): the gcn-norm aggregation of
n_rows seed cells over their sampled neighbourhood, (a) from a dense fp32 / bf16 x tile (the reference's
layout, nn/aggregators.py:205-213), (b) straight from the CSR count store.  CUDA events, L2 flushed."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                                                    # noqa: E402
from nichecompass_b200 import ops                               # noqa: E402
from nichecompass_b200.data import make_synthetic_adata_large   # noqa: E402
from nichecompass_b200.data.store import CSRFeatureStore        # noqa: E402
from nichecompass_b200.data.sampler import DeviceGraph, sample_subgraph  # noqa: E402
from nichecompass_b200.data.datasets import adjacency_to_edge_index      # noqa: E402

dev = torch.device("cuda")
_flush = []


def _flush_buf():
    if not _flush:
        _flush.append(torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev))
    return _flush[0]



def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        _flush_buf().zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2]


def main(n_cells=200_000, n_genes=5000, n_rows=16384):
    adata = make_synthetic_adata_large(n_cells, n_genes, 8, 1500, seed=0, lib_mean=300.)
    store = CSRFeatureStore.from_scipy(adata.X.tocsr(), dev)
    ei = adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    g = DeviceGraph(ei, n_cells, dev)
    return measure(store, g, n_rows)


def measure(store, g, n_rows=16384, verbose=True):
    """-> {"dense_f32" | "dense_bf16" | "csr_store": {"ms", "algorithmic_bytes", "gbs"}}"""
    n_cells, n_genes = g.num_nodes, store.n_cols
    res = {}
    say = print if verbose else (lambda *a, **k: None)
    seeds = torch.randperm(n_cells, device=dev)[:n_rows].to(torch.int32).sort().values
    n_id, graph = sample_subgraph(g, seeds, n_rows, [-1])
    n_sub = graph.n_nodes
    w, wself = graph.gcn_weights(use_out_degree=True, n_rows=n_rows)
    deg = (graph.rowptr[1:n_rows + 1] - graph.rowptr[:n_rows]).float()
    entries = float(deg.sum()) + n_rows                         # gathered rows incl. the self term
    nnz_row = store.nnz / store.n_rows
    say(f"n_rows {n_rows}, n_sub {n_sub}, gathered rows {entries:.0f} ({entries / n_rows:.2f}/row), store nnz/row {nnz_row:.1f}")
    for dt, s in ((torch.float32, 4), (torch.bfloat16, 2)):
        x = ops.csr_gather_rows(store, n_id, n_sub, False, dt)
        out = torch.empty(n_rows, n_genes, dtype=torch.float32, device=dev)
        ms = timeit(lambda: ops.spmm_csr(x, graph.rowptr, graph.col, w, wself, True, n_rows, out=out))
        alg = entries * n_genes * s + n_rows * n_genes * 4 + entries * 8
        say(f"dense x {dt}: {ms:.3f} ms, algorithmic {alg / 1e6:.0f} MB -> {alg / ms / 1e6:.0f} GB/s")
        res["dense_f32" if s == 4 else "dense_bf16"] = {"ms": ms, "algorithmic_bytes": alg, "gbs": alg / ms / 1e6}
    ms = timeit(lambda: ops.spmm_csr_store(store, n_id, graph.rowptr, graph.col, w, wself, True, n_rows))
    rp = store.rowptr
    ids = n_id.long()
    nnz_n = (rp[ids + 1] - rp[ids])
    live = torch.arange(graph.col.numel(), device=dev) < graph.rowptr[n_rows]
    gathered_nnz = float((nnz_n[graph.col.long()] * live).sum() + nnz_n[:n_rows].sum())
    alg = gathered_nnz * 8 + n_rows * n_genes * 4 + entries * 8
    say(f"CSR store: {ms:.3f} ms, algorithmic {alg / 1e6:.0f} MB ({gathered_nnz:.0f} nnz) -> {alg / ms / 1e6:.0f} GB/s")
    res["csr_store"] = {"ms": ms, "algorithmic_bytes": alg, "gbs": alg / ms / 1e6}
    res["shape"] = {"n_rows": n_rows, "n_sub": n_sub, "gathered_rows": entries, "n_genes": n_genes}
    return res


if __name__ == "__main__":
    main()
