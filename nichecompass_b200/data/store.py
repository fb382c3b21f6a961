"""
HBM-resident CSR count store.

The reference densifies the whole count matrix on the host (data/datasets.py:59-71
``torch.tensor(x.toarray())``: 20 GB at 1M x 5k, 440 GB at 2M x 55k) and slices dense rows per
batch on the CPU before a pageable H2D copy.  Here the counts stay sparse on the device
(int64 row pointers, int32 column ids, fp32 values: 8 B per non-zero) and batches are gathered
by ``nc_csr_gather_rows`` / aggregated by ``nc_spmm_csr_store`` straight from this store.
"""
from typing import Optional

import numpy as np
import scipy.sparse as sp
import torch


class CSRFeatureStore:
    def __init__(self, rowptr: torch.Tensor, col: torch.Tensor, val: torch.Tensor, n_cols: int):
        assert rowptr.dtype == torch.int64 and col.dtype == torch.int32 and val.dtype == torch.float32
        self.rowptr, self.col, self.val, self.n_cols = rowptr, col, val, int(n_cols)

    @property
    def location(self) -> str:
        """"device": non-zeros in HBM.  "host": non-zeros in pinned host memory, read by the
        gather / SpMM kernels straight over the host link (zero-copy) -- for count matrices larger
        than HBM and for the end-to-end leg of bench.py (every step pulls its inputs from the host)."""
        return "device" if self.val.is_cuda else "host"

    @property
    def n_rows(self) -> int:
        return self.rowptr.numel() - 1

    @property
    def nnz(self) -> int:
        return self.col.numel()

    @property
    def device(self):
        return self.rowptr.device

    @classmethod
    def from_scipy(cls, x, device, x2=None, location: str = "device") -> "CSRFeatureStore":
        """``x`` (and optionally ``x2`` appended column-wise, the RNA ++ ATAC layout of
        data/datasets.py:65-71): scipy sparse or ndarray."""
        if location not in ("device", "host"):
            raise ValueError(f"unknown store location {location}")
        x = sp.csr_matrix(x)
        if x2 is not None:
            x = sp.hstack([x, sp.csr_matrix(x2)], format="csr")
        x.sort_indices()
        pin = torch.cuda.is_available()

        def up(a, dt, keep_on_host=False):
            t = torch.from_numpy(np.ascontiguousarray(a.astype(dt, copy=False)))
            if pin:
                t = t.pin_memory()
            return t if keep_on_host else t.to(device, non_blocking=True)
        host = location == "host"
        return cls(up(x.indptr, np.int64), up(x.indices, np.int32, host), up(x.data, np.float32, host),
                   x.shape[1])

    def to_dense(self, rows: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Debug / compatibility helper (torch index plumbing, not a hot path)."""
        n = self.n_rows
        counts = (self.rowptr[1:] - self.rowptr[:-1])
        r = torch.repeat_interleave(torch.arange(n, device=self.device), counts)
        d = torch.zeros(n, self.n_cols, dtype=torch.float32, device=self.device)
        d[r, self.col.to(self.device).long()] = self.val.to(self.device)
        return d if rows is None else d[rows]
