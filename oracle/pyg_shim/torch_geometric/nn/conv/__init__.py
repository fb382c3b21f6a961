import inspect

import torch


class MessagePassing(torch.nn.Module):
    """Minimal torch_geometric MessagePassing: flow source_to_target, aggr='add'.
    ``propagate(edge_index, name=(src_tensor, dst_tensor) | tensor)`` builds ``name_j`` /
    ``name_i`` by gathering along ``node_dim`` and sums ``message(...)`` at the targets."""

    def __init__(self, aggr="add", node_dim=-2, **kwargs):
        super().__init__()
        self.aggr, self.node_dim = aggr, node_dim

    def propagate(self, edge_index, size=None, **kwargs):
        j, i = edge_index[0], edge_index[1]
        params = inspect.signature(self.message).parameters
        args, n_dst = {}, None
        for name, val in kwargs.items():
            src, dst = val if isinstance(val, tuple) else (val, val)
            n_dst = dst.size(self.node_dim) if n_dst is None else n_dst
            if f"{name}_j" in params:
                args[f"{name}_j"] = src.index_select(self.node_dim, j)
            if f"{name}_i" in params:
                args[f"{name}_i"] = dst.index_select(self.node_dim, i)
        if "index" in params:
            args["index"] = i
        msg = self.message(**args)
        shape = list(msg.shape)
        shape[self.node_dim] = n_dst
        out = torch.zeros(shape, dtype=msg.dtype, device=msg.device)
        return out.index_add_(self.node_dim, i, msg)
