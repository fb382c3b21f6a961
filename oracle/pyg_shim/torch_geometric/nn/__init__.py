import torch

from oracle import pyg_semantics as S
from . import conv, dense, inits  # noqa: F401
from .dense.linear import Linear
from .inits import glorot


class GATv2Conv(torch.nn.Module):
    """Parameter layout of torch_geometric.nn.GATv2Conv (share_weights=False, no edge_dim):
    att [1,H,C], lin_l / lin_r Linear(in, H*C, bias=True, glorot), bias [C] (concat=False)."""

    def __init__(self, in_channels, out_channels, heads=1, concat=True, negative_slope=0.2,
                 dropout=0.0, add_self_loops=True, bias=True, **kwargs):
        super().__init__()
        assert not concat and add_self_loops and dropout == 0.0
        self.heads, self.out_channels, self.negative_slope = heads, out_channels, negative_slope
        self.lin_l = Linear(in_channels, heads * out_channels, bias=True, weight_initializer="glorot")
        self.lin_r = Linear(in_channels, heads * out_channels, bias=True, weight_initializer="glorot")
        self.att = torch.nn.Parameter(torch.empty(1, heads, out_channels))
        self.bias = torch.nn.Parameter(torch.zeros(out_channels)) if bias else None
        glorot(self.att)

    def forward(self, x, edge_index):
        return S.gatv2_conv(x, edge_index, self.lin_l.weight, self.lin_l.bias, self.lin_r.weight,
                            self.lin_r.bias, self.att, self.bias, self.heads, self.negative_slope)


class GCNConv(torch.nn.Module):
    """torch_geometric.nn.GCNConv defaults: lin = Linear(in, out, bias=False, glorot), bias [out]."""

    def __init__(self, in_channels, out_channels, **kwargs):
        super().__init__()
        self.lin = Linear(in_channels, out_channels, bias=False, weight_initializer="glorot")
        self.bias = torch.nn.Parameter(torch.zeros(out_channels))

    def forward(self, x, edge_index):
        return S.gcn_conv(x, edge_index, self.lin.weight, self.bias)
