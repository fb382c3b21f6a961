"""
CPU restatement of the torch_geometric / torch_sparse arithmetic used on the NicheCompass hot
path (TEST INFRASTRUCTURE -- see oracle/__init__.py).  torch_geometric is an un-vendored,
unpinned dependency of the reference (pyproject.toml:31; torch-sparse 0.6.17 / torch-scatter
2.1.1 only in environment.yaml:8-12); its published semantics are restated here from
SURVEY.md App. A.  Flow is source->target: ``j = edge_index[0]`` (neighbour),
``i = edge_index[1]`` (centre).

Everything is plain dense torch on whatever dtype the inputs carry (fp64 for the tight
oracle-vs-reference check, fp32 for the timed CPU baseline).
"""
from typing import Optional, Tuple

import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- utils
def remove_self_loops(edge_index: torch.Tensor,
                      edge_attr: Optional[torch.Tensor] = None):
    """torch_geometric.utils.remove_self_loops: drop columns with row == col."""
    keep = edge_index[0] != edge_index[1]
    ei = edge_index[:, keep]
    return ei, (None if edge_attr is None else edge_attr[keep])


def add_self_loops(edge_index: torch.Tensor,
                   edge_attr: Optional[torch.Tensor] = None,
                   fill_value=None,
                   num_nodes: Optional[int] = None):
    """torch_geometric.utils.add_self_loops: append (v, v) for v = 0..n-1 at the END."""
    n = int(edge_index.max()) + 1 if num_nodes is None else int(num_nodes)
    loop = torch.arange(n, dtype=edge_index.dtype, device=edge_index.device)
    ei = torch.cat([edge_index, torch.stack([loop, loop])], dim=1)
    if edge_attr is not None:
        fv = 1.0 if fill_value is None else fill_value
        loop_attr = torch.full((n,) + tuple(edge_attr.shape[1:]), fv,
                               dtype=edge_attr.dtype, device=edge_attr.device)
        edge_attr = torch.cat([edge_attr, loop_attr], dim=0)
    return ei, edge_attr


def segment_softmax(src: torch.Tensor, index: torch.Tensor, num_nodes: int) -> torch.Tensor:
    """torch_geometric.utils.softmax(src, index): softmax over entries sharing ``index``
    along dim 0 (max-subtracted; denominator + 1e-16)."""
    shape = (num_nodes,) + tuple(src.shape[1:])
    idx = index.view(-1, *([1] * (src.dim() - 1))).expand_as(src)
    mx = torch.full(shape, float("-inf"), dtype=src.dtype, device=src.device)
    mx = mx.scatter_reduce(0, idx, src, reduce="amax", include_self=True)
    out = (src - mx.index_select(0, index)).exp()
    den = torch.zeros(shape, dtype=src.dtype, device=src.device).index_add_(0, index, out)
    return out / (den.index_select(0, index) + 1e-16)


def glorot_(t: torch.Tensor) -> torch.Tensor:
    """torch_geometric.nn.inits.glorot: U(-a, a), a = sqrt(6 / (size(-2) + size(-1)))."""
    a = (6.0 / (t.size(-2) + t.size(-1))) ** 0.5
    with torch.no_grad():
        t.uniform_(-a, a)
    return t


# --------------------------------------------------------------------------- GATv2Conv
def gatv2_conv(x: torch.Tensor, edge_index: torch.Tensor,
               lin_l_w: torch.Tensor, lin_l_b: Optional[torch.Tensor],
               lin_r_w: torch.Tensor, lin_r_b: Optional[torch.Tensor],
               att: torch.Tensor, bias: Optional[torch.Tensor],
               heads: int, negative_slope: float = 0.2,
               return_alpha: bool = False, round_fn=None):
    """GATv2Conv(in, C, heads=H, concat=False, add_self_loops=True, share_weights=False)
    (SURVEY.md App. A.1).  Returns [n, C] (head mean + bias).
    ``round_fn`` (oracle/bf16_emulation.py, tests of the bf16 mode only): applied to x_l and x_r, the two tensors
    the CUDA tensor-core mode stores in bf16; ``None`` = the reference arithmetic."""
    n = x.size(0)
    H = heads
    C = lin_l_w.size(0) // H
    x_l = F.linear(x, lin_l_w, lin_l_b)
    x_r = F.linear(x, lin_r_w, lin_r_b)
    if round_fn is not None:
        x_l, x_r = round_fn(x_l), round_fn(x_r)
    x_l, x_r = x_l.view(n, H, C), x_r.view(n, H, C)
    ei, _ = remove_self_loops(edge_index)
    ei, _ = add_self_loops(ei, num_nodes=n)
    j, i = ei[0], ei[1]
    e = F.leaky_relu(x_l.index_select(0, j) + x_r.index_select(0, i), negative_slope)
    a = (e * att.view(1, H, C)).sum(-1)                      # [E, H]
    alpha = segment_softmax(a, i, n)
    msg = x_l.index_select(0, j) * alpha.unsqueeze(-1)       # [E, H, C]
    out = torch.zeros(n, H, C, dtype=x.dtype, device=x.device).index_add_(0, i, msg)
    out = out.mean(dim=1)
    if bias is not None:
        out = out + bias
    if return_alpha:
        return out, (ei, alpha)
    return out


# --------------------------------------------------------------------------- GCNConv
def gcn_norm_edge_index(edge_index: torch.Tensor, num_nodes: int, dtype) -> Tuple[torch.Tensor, torch.Tensor]:
    """gcn_norm(edge_index, None, n, improved=False, add_self_loops=True) for a plain
    edge_index (App. A.2): add_remaining_self_loops (every node ends with exactly one self
    loop of weight 1), deg over targets, w = deg^-1/2[src] * deg^-1/2[dst] (inf -> 0)."""
    keep = edge_index[0] != edge_index[1]
    loop = torch.arange(num_nodes, dtype=edge_index.dtype, device=edge_index.device)
    ei = torch.cat([edge_index[:, keep], torch.stack([loop, loop])], dim=1)
    w = torch.ones(ei.size(1), dtype=dtype, device=edge_index.device)
    deg = torch.zeros(num_nodes, dtype=dtype, device=edge_index.device).index_add_(0, ei[1], w)
    dis = deg.pow(-0.5)
    dis = torch.where(torch.isinf(dis), torch.zeros_like(dis), dis)
    return ei, dis[ei[0]] * w * dis[ei[1]]


def gcn_conv(x: torch.Tensor, edge_index: torch.Tensor, lin_w: torch.Tensor,
             bias: Optional[torch.Tensor]) -> torch.Tensor:
    """GCNConv(in, C) defaults (App. A.2): out[i] = sum_j w_ji (x W^T)[j] + bias."""
    n = x.size(0)
    ei, w = gcn_norm_edge_index(edge_index, n, x.dtype)
    xw = F.linear(x, lin_w)
    out = torch.zeros(n, xw.size(1), dtype=x.dtype, device=x.device)
    out.index_add_(0, ei[1], xw.index_select(0, ei[0]) * w.unsqueeze(-1))
    if bias is not None:
        out = out + bias
    return out


# --------------------------------------------------------------------------- aggregators
def sparse_adj_t_matmul(edge_index: torch.Tensor, value: Optional[torch.Tensor],
                        x: torch.Tensor) -> torch.Tensor:
    """``SparseTensor.from_edge_index(edge_index, value).t().matmul(x)``:
    out[c] = sum over entries (r, c) of value * x[r]  (multi-edges each count)."""
    n = x.size(0)
    msg = x.index_select(0, edge_index[0])
    if value is not None:
        msg = msg * value.unsqueeze(-1)
    return torch.zeros(n, x.size(1), dtype=x.dtype, device=x.device).index_add_(0, edge_index[1], msg)


def gcn_norm_sparse(edge_index: torch.Tensor, num_nodes: int, dtype):
    """gcn_norm(SparseTensor adj) branch (App. A.3): values := 1, fill_diag(adj, 1)
    (existing diagonal entries replaced by exactly one of value 1), deg = ROW sum,
    adj_norm[r, c] = deg[r]^-1/2 * deg[c]^-1/2.  Returns the entries in SparseTensor
    (row-major sorted) order as (edge_index_sorted, value)."""
    keep = edge_index[0] != edge_index[1]
    loop = torch.arange(num_nodes, dtype=edge_index.dtype, device=edge_index.device)
    ei = torch.cat([edge_index[:, keep], torch.stack([loop, loop])], dim=1)
    order = torch.argsort(ei[0] * num_nodes + ei[1], stable=True)
    ei = ei[:, order]
    deg = torch.zeros(num_nodes, dtype=dtype, device=edge_index.device).index_add_(
        0, ei[0], torch.ones(ei.size(1), dtype=dtype, device=edge_index.device))
    dis = deg.pow(-0.5)
    dis = torch.where(torch.isinf(dis), torch.zeros_like(dis), dis)
    return ei, dis[ei[0]] * dis[ei[1]]


def one_hop_gcn_norm_aggregate(x: torch.Tensor, edge_index: torch.Tensor,
                               return_agg_weights: bool = False):
    """nn/aggregators.py:205-213."""
    ei, val = gcn_norm_sparse(edge_index, x.size(0), x.dtype)
    out = sparse_adj_t_matmul(ei, val, x)
    return out, (val if return_agg_weights else None)


def one_hop_sum_aggregate(x: torch.Tensor, edge_index: torch.Tensor):
    """nn/aggregators.py:260-264 (self loops exactly as present in edge_index)."""
    return sparse_adj_t_matmul(edge_index, None, x), None


def one_hop_attention_aggregate(x: torch.Tensor, edge_index: torch.Tensor,
                                w_l: torch.Tensor, w_r: torch.Tensor, attn: torch.Tensor,
                                heads: int, negative_slope: float = 0.2,
                                return_agg_weights: bool = False):
    """nn/aggregators.py:113-156 (App. A.4): additive attention on raw features, no self-loop
    manipulation, mean over heads."""
    n, Fd = x.shape
    H = heads
    g_l = F.linear(x, w_l).view(n, H, Fd)
    g_r = F.linear(x, w_r).view(n, H, Fd)
    j, i = edge_index[0], edge_index[1]
    g = F.leaky_relu(g_r.index_select(0, i) + g_l.index_select(0, j), negative_slope)
    a = (g * attn.view(1, H, Fd)).sum(-1)
    alpha = segment_softmax(a, i, n)
    msg = x.index_select(0, j).unsqueeze(1) * alpha.unsqueeze(-1)   # [E, H, F]
    out = torch.zeros(n, H, Fd, dtype=x.dtype, device=x.device).index_add_(0, i, msg)
    return out.mean(dim=1), (alpha if return_agg_weights else None)
