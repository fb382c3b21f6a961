from oracle.pyg_semantics import gcn_norm_edge_index, gcn_norm_sparse


def gcn_norm(adj_or_edge_index, edge_weight=None, num_nodes=None, improved=False,
             add_self_loops=True, flow="source_to_target", dtype=None):
    from torch_sparse import SparseTensor
    if isinstance(adj_or_edge_index, SparseTensor):
        adj = adj_or_edge_index
        ei, val = gcn_norm_sparse(adj.edge_index, adj.n, adj.dtype)
        return SparseTensor(ei, val, adj.n, adj.dtype)
    return gcn_norm_edge_index(adj_or_edge_index, num_nodes, dtype)
