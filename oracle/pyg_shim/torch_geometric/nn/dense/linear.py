import math

import torch
import torch.nn.functional as F

from ..inits import glorot


class Linear(torch.nn.Module):
    """torch_geometric.nn.dense.linear.Linear (weight [out, in]; glorot or kaiming-uniform)."""

    def __init__(self, in_channels, out_channels, bias=True, weight_initializer=None,
                 bias_initializer=None):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.weight_initializer = weight_initializer
        self.weight = torch.nn.Parameter(torch.empty(out_channels, in_channels))
        self.bias = torch.nn.Parameter(torch.empty(out_channels)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        if self.weight_initializer == "glorot":
            glorot(self.weight)
        else:
            torch.nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            self.bias.data.zero_()

    def forward(self, x):
        return F.linear(x, self.weight, self.bias)
