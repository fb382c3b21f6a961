"""
HBM-resident CSR count store.

The reference densifies the whole count matrix on the host (data/datasets.py:59-71
``torch.tensor(x.toarray())``: 20 GB at 1M x 5k, 440 GB at 2M x 55k) and slices dense rows per
batch on the CPU before a pageable H2D copy.  Here the counts stay sparse on the device
(int64 row pointers, int32 column ids, fp32 values: 8 B per non-zero) and batches are gathered
by ``nc_csr_gather_rows`` / aggregated by ``nc_spmm_csr_store`` straight from this store.
"""
import ctypes
from typing import Optional

import numpy as np
import scipy.sparse as sp
import torch


class PinnedBlock:
    """One ``cudaHostAlloc`` block (portable, device-mapped) owned by a host-resident store: allocated through
    ``nc_host_alloc``, released with ``nc_host_free`` (``cudaFreeHost`` waits for the device) when the last
    tensor view is gone.  Unlike ``Tensor.pin_memory()`` the block never passes through torch's caching host
    allocator, which knows nothing about kernels launched through the C ABI and could hand a block that a
    zero-copy kernel is still reading to the next ``pin_memory()`` call."""

    def __init__(self, nbytes: int):
        from .._lib import NicheCompassCudaError, lib
        self._lib = lib()
        p = ctypes.c_void_p()
        rc = self._lib.nc_host_alloc(ctypes.c_int64(int(nbytes)), ctypes.byref(p))
        if rc != 0:
            raise NicheCompassCudaError(self._lib.nc_last_error().decode())
        self.ptr, self.nbytes = p.value, int(nbytes)

    def __del__(self):
        if getattr(self, "ptr", None):
            self._lib.nc_host_free(ctypes.c_void_p(self.ptr))
            self.ptr = None

    @classmethod
    def from_numpy(cls, a: np.ndarray) -> torch.Tensor:
        """Copy ``a`` into a fresh block; the returned tensor keeps the block alive."""
        a = np.ascontiguousarray(a)
        if a.nbytes == 0:
            return torch.from_numpy(a.copy()).pin_memory()
        blk = cls(a.nbytes)
        buf = (ctypes.c_char * a.nbytes).from_address(blk.ptr)
        view = np.frombuffer(buf, dtype=a.dtype).reshape(a.shape)
        view[...] = a
        t = torch.from_numpy(view)
        t._nc_block = blk                      # tensor -> block; numpy view -> ctypes buffer (no copy anywhere)
        _BLOCKS[t.data_ptr()] = blk
        return t


# data_ptr -> block for every live host-resident array (views made by torch slicing drop python attributes;
# the store below additionally holds its blocks itself)
_BLOCKS = {}


class CSRFeatureStore:
    def __init__(self, rowptr: torch.Tensor, col: torch.Tensor, val: torch.Tensor, n_cols: int):
        assert rowptr.dtype == torch.int64 and col.dtype == torch.int32 and val.dtype == torch.float32
        self.rowptr, self.col, self.val, self.n_cols = rowptr, col, val, int(n_cols)
        self._blocks = [getattr(t, "_nc_block", None) for t in (col, val)]

    def __del__(self):
        for t in (getattr(self, "col", None), getattr(self, "val", None)):
            if t is not None and not t.is_cuda:
                _BLOCKS.pop(t.data_ptr(), None)

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
            a = np.ascontiguousarray(a.astype(dt, copy=False))
            if keep_on_host:
                if not pin:
                    raise RuntimeError("a host-resident count store needs a CUDA device (pinned, device-mapped memory)")
                return PinnedBlock.from_numpy(a)
            t = torch.from_numpy(a)
            # pageable -> device: a synchronous copy through the driver's own staging buffers (no pinned block of
            # torch's caching host allocator is involved, nothing to recycle)
            return t.to(device)
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
