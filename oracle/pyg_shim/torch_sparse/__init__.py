"""Stand-in for torch_sparse.SparseTensor: entries kept in row-major sorted COO order."""
import torch

from oracle.pyg_semantics import sparse_adj_t_matmul


# dtype of implicit values (torch_sparse uses float32; the fp64 harness raises it so the
# oracle-vs-reference comparison is not limited by fp32 rounding of the norm weights)
DEFAULT_DTYPE = torch.float32


class SparseTensor:
    def __init__(self, edge_index, value, n, dtype=None, transposed=False):
        dtype = DEFAULT_DTYPE if dtype is None else dtype
        self.edge_index, self.value, self.n, self.dtype = edge_index, value, n, dtype
        self.transposed = transposed

    @classmethod
    def from_edge_index(cls, edge_index, edge_attr=None, sparse_sizes=None, **kw):
        n = sparse_sizes[0]
        order = torch.argsort(edge_index[0] * n + edge_index[1], stable=True)
        val = None if edge_attr is None else edge_attr[order]
        return cls(edge_index[:, order], val, n)

    @classmethod
    def from_torch_sparse_coo_tensor(cls, mat, has_value=True):
        mat = mat.coalesce()
        return cls(mat.indices(), mat.values() if has_value else None, mat.size(0), mat.dtype)

    def to_torch_sparse_coo_tensor(self):
        val = self.value if self.value is not None else torch.ones(self.edge_index.size(1))
        return torch.sparse_coo_tensor(self.edge_index, val, (self.n, self.n)).coalesce()

    def nnz(self):
        return self.edge_index.size(1)

    def t(self):
        return SparseTensor(self.edge_index, self.value, self.n, self.dtype, not self.transposed)

    def coo(self):
        return self.edge_index[0], self.edge_index[1], self.value

    def matmul(self, x):
        ei = self.edge_index if self.transposed else self.edge_index.flip([0])
        # transposed: out[c] = sum_(r,c) v x[r];  plain: out[r] = sum_(r,c) v x[c]
        return sparse_adj_t_matmul(ei, None if self.value is None else self.value.to(x.dtype), x)
