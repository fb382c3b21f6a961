"""
One-hop node-label aggregators (reference: nn/aggregators.py).  The sum / GCN-norm variants are
parameter free: one weighted CSR SpMM (``nc_spmm_csr`` on dense rows, ``nc_spmm_csr_store``
straight from the HBM count store).  The attention variant keeps the reference parameters
(``linear_l_l``, ``linear_r_l``, ``attn``) and runs the fused GATv2-style edge kernels with
the raw features as messages.
"""
from typing import Literal, Optional

import torch
import torch.nn as nn

from .. import ops
from ..functional import AttentionAggregatorFn, SubgraphCSR, new_cache_key
from .conv import _Lin, glorot_


def _as_graph(edge_index_or_graph, n_nodes: int) -> SubgraphCSR:
    if isinstance(edge_index_or_graph, SubgraphCSR):
        return edge_index_or_graph
    return SubgraphCSR.from_edge_index(edge_index_or_graph, n_nodes)


def _entry_weights_in_edge_order(graph: SubgraphCSR, w: torch.Tensor, wself: torch.Tensor) -> torch.Tensor:
    """Aggregation weights as a flat vector (entries grouped by target, the appended self term
    last per target) -- only built when ``return_agg_weights`` is requested."""
    return torch.cat((w[:graph.col.numel()], wself))


class OneHopGCNNormNodeLabelAggregator(nn.Module):
    """x_nbr[c] = sum_r d_r^-1/2 d_c^-1/2 x[r] over the induced subgraph with exactly one self
    loop per node and subgraph-local degrees (nn/aggregators.py:205-213)."""

    def __init__(self, modality: Literal["rna", "atac"]):
        super().__init__()
        print(f"ONE HOP GCN NORM {modality.upper()} NODE LABEL AGGREGATOR")

    def forward(self, x: torch.Tensor, edge_index, return_agg_weights: bool = False, n_rows: Optional[int] = None):
        graph = _as_graph(edge_index, x.size(0))
        n_rows = graph.n_nodes if n_rows is None else n_rows
        w, wself = graph.gcn_weights(use_out_degree=True, n_rows=n_rows)
        out = ops.spmm_csr(x.detach(), graph.rowptr, graph.col, w, wself, True, n_rows)
        return out, (_entry_weights_in_edge_order(graph, w, wself) if return_agg_weights else None)

    def forward_store(self, store, n_id, graph: SubgraphCSR, n_rows: int, out=None) -> torch.Tensor:
        w, wself = graph.gcn_weights(use_out_degree=True, n_rows=n_rows)
        return ops.spmm_csr_store(store, n_id, graph.rowptr, graph.col, w, wself, True, n_rows, out=out)


class OneHopSumNodeLabelAggregator(nn.Module):
    """x_nbr[c] = sum over edges (r -> c) of x[r], self loops exactly as present
    (nn/aggregators.py:260-264)."""

    def __init__(self, modality: Literal["rna", "atac"]):
        super().__init__()
        print(f"ONE HOP SUM {modality.upper()} NODE LABEL AGGREGATOR")

    def forward(self, x: torch.Tensor, edge_index, return_agg_weights: bool = False, n_rows: Optional[int] = None):
        graph = _as_graph(edge_index, x.size(0))
        n_rows = graph.n_nodes if n_rows is None else n_rows
        return ops.spmm_csr(x.detach(), graph.rowptr, graph.col, None, None, False, n_rows), None

    def forward_store(self, store, n_id, graph: SubgraphCSR, n_rows: int, out=None) -> torch.Tensor:
        return ops.spmm_csr_store(store, n_id, graph.rowptr, graph.col, None, None, False, n_rows, out=out)


class OneHopAttentionNodeLabelAggregator(nn.Module):
    """Additive-attention aggregation of the neighbours' raw feature vectors
    (nn/aggregators.py:17-156): g_l = linear_l_l(x), g_r = linear_r_l(x),
    alpha = softmax_i(attn . leaky_relu(g_r[i] + g_l[j])), out[i] = mean_h sum_j alpha x[j]."""

    def __init__(self, modality: Literal["rna", "atac"], n_input: int, n_heads: int = 4,
                 leaky_relu_negative_slope: float = 0.2, dropout_rate: float = 0.):
        super().__init__()
        if dropout_rate != 0.:
            raise NotImplementedError("attention dropout is not implemented in the CUDA path")
        self.n_input, self.n_heads = n_input, n_heads
        self.leaky_relu_negative_slope = leaky_relu_negative_slope
        self.linear_l_l = _Lin(n_input, n_input * n_heads, bias=False)
        self.linear_r_l = _Lin(n_input, n_input * n_heads, bias=False)
        self.attn = nn.Parameter(glorot_(torch.empty(1, n_heads, n_input)))
        self.compute_dtype = torch.float32
        self._key = new_cache_key(self)
        print(f"ONE HOP ATTENTION {modality.upper()} NODE LABEL AGGREGATOR -> "
              f"n_input: {n_input}, n_heads: {n_heads}")

    def forward(self, x: torch.Tensor, edge_index, return_agg_weights: bool = False, n_rows: Optional[int] = None):
        graph = _as_graph(edge_index, x.size(0))
        if n_rows is not None and n_rows != graph.n_dst:
            graph = SubgraphCSR(graph.rowptr, graph.col, graph.n_nodes, n_rows, graph.max_deg)
        out, alpha = AttentionAggregatorFn.apply(x.detach(), graph, self.n_heads, self.leaky_relu_negative_slope,
                                                 self.compute_dtype, self._key, bool(return_agg_weights),
                                                 self.linear_l_l.weight, self.linear_r_l.weight, self.attn)
        return out, (alpha if return_agg_weights else None)

    def forward_store(self, store, n_id, graph: SubgraphCSR, n_rows: int) -> torch.Tensor:
        """Count rows of all sampled cells gathered densely from the store (the attention scores need
        linear maps of every neighbour's full row), then the same kernels as ``forward``."""
        x = ops.csr_gather_rows(store, n_id, graph.n_nodes, False, torch.float32)
        return self.forward(x[:, :self.n_input], graph, n_rows=n_rows)[0]
