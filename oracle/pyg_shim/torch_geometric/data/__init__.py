import copy

import torch


class Data:
    """Attribute bag with the handful of torch_geometric.data.Data methods the reference uses."""

    def __init__(self, x=None, edge_index=None, edge_attr=None, **kwargs):
        if x is not None:
            self.x = x
        if edge_index is not None:
            self.edge_index = edge_index
        if edge_attr is not None:
            self.edge_attr = edge_attr
        for k, v in kwargs.items():
            setattr(self, k, v)

    @property
    def num_nodes(self):
        return self.x.size(0)

    def keys(self):
        return list(self.__dict__.keys())

    def clone(self):
        out = Data()
        for k, v in self.__dict__.items():
            setattr(out, k, v.clone() if torch.is_tensor(v) else copy.deepcopy(v))
        return out

    def to(self, device):
        for k, v in self.__dict__.items():
            if torch.is_tensor(v):
                setattr(self, k, v.to(device))
        return self

    def __contains__(self, key):
        return key in self.__dict__
