"""
Parameter containers with the state_dict layout of the PyG convolutions the reference
instantiates (nn/encoders.py:105-140): ``GATv2Conv`` -> att [1,H,C], lin_l/lin_r
{weight [H*C, in], bias [H*C]}, bias [C];  ``GCNConv`` -> lin.weight [C, in], bias [C].
Initialisation follows PyG (glorot weights / attention vector, zero biases).  The arithmetic
lives in the fused kernels driven by ``Encoder``; these modules own the parameters only.
"""
import math

import torch
import torch.nn as nn


def glorot_(t: torch.Tensor) -> torch.Tensor:
    bound = math.sqrt(6.0 / (t.size(-2) + t.size(-1)))
    with torch.no_grad():
        return t.uniform_(-bound, bound)


class _Lin(nn.Module):
    def __init__(self, n_in: int, n_out: int, bias: bool):
        super().__init__()
        self.weight = nn.Parameter(glorot_(torch.empty(n_out, n_in)))
        if bias:
            self.bias = nn.Parameter(torch.zeros(n_out))
        else:
            self.register_parameter("bias", None)


class GATv2Conv(nn.Module):
    def __init__(self, in_channels: int, out_channels: int, heads: int = 1, concat: bool = False):
        super().__init__()
        if concat:
            raise NotImplementedError("the reference only uses concat=False")
        self.in_channels, self.out_channels, self.heads = in_channels, out_channels, heads
        self.lin_l = _Lin(in_channels, heads * out_channels, bias=True)
        self.lin_r = _Lin(in_channels, heads * out_channels, bias=True)
        self.att = nn.Parameter(glorot_(torch.empty(1, heads, out_channels)))
        self.bias = nn.Parameter(torch.zeros(out_channels))

    def param_tuple(self):
        return (self.lin_l.weight, self.lin_l.bias, self.lin_r.weight, self.lin_r.bias, self.att, self.bias)


class GCNConv(nn.Module):
    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.lin = _Lin(in_channels, out_channels, bias=False)
        self.bias = nn.Parameter(torch.zeros(out_channels))
