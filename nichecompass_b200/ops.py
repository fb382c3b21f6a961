"""
Thin tensor-level wrappers over the C ABI (raw, non-differentiable).  The autograd Functions
the ``nn`` modules use live in ``nichecompass_b200/functional.py`` and are built from these.
PyTorch is used for allocation, streams and index plumbing only.
"""
from typing import Optional, Tuple

import torch

from . import _lib
from ._lib import call, dtype_code


def _i32(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """Index argument of the C ABI: int32 and contiguous.  The sampler emits exactly that (no work here);
    an int64 / strided index tensor of foreign (PyG-style) code is converted instead of being read as garbage
    through its raw pointer."""
    if t is None or (t.dtype == torch.int32 and t.is_contiguous()):
        return t
    if t.dtype not in (torch.int32, torch.int64, torch.int16, torch.uint8):
        raise TypeError(f"index tensor of dtype {t.dtype}")
    return t.to(torch.int32).contiguous()


def pad8(n: int) -> int:
    return (n + 7) // 8 * 8


# --------------------------------------------------------------------------- graph plumbing
def coo_to_dst_csr(edge_index: torch.Tensor, n_nodes: int) -> Tuple[torch.Tensor, torch.Tensor]:
    """edge_index [2, E] (row 0 = source, row 1 = target) -> dst-CSR (rowptr int32 [n+1],
    col int32 [E]); entries of a row keep their edge_index order (stable).  Format conversion
    of a foreign (PyG-style) batch: the GPU sampler emits dst-CSR directly."""
    src, dst = edge_index[0], edge_index[1]
    if dst.numel() > 1 and bool((dst[1:] < dst[:-1]).any()):
        order = torch.sort(dst, stable=True).indices
        src, dst = src[order], dst[order]
    counts = torch.bincount(dst, minlength=n_nodes)
    rowptr = torch.zeros(n_nodes + 1, dtype=torch.int32, device=edge_index.device)
    rowptr[1:] = torch.cumsum(counts, 0)
    return rowptr, src.to(torch.int32).contiguous()


def gcn_norm_weights(rowptr, col, n_nodes: int, n_rows: int, use_out_degree: bool):
    rowptr, col = _i32(rowptr), _i32(col)
    dev = rowptr.device
    w = torch.empty(max(col.numel(), 1), dtype=torch.float32, device=dev)
    wself = torch.empty(max(n_rows, 1), dtype=torch.float32, device=dev)
    deg = torch.empty(max(n_nodes, 1), dtype=torch.int32, device=dev)
    call("nc_gcn_norm_weights", rowptr, col, n_nodes, n_rows, int(use_out_degree), deg, w, wself)
    return w, wself


def reverse_csr(rowptr, col, n_nodes: int, n_rows: int, add_self: bool, n_entries: Optional[int] = None):
    """``n_entries``: upper bound of the number of reverse entries (host int; avoids a sync)."""
    rowptr, col = _i32(rowptr), _i32(col)
    dev = rowptr.device
    if n_entries is None:
        n_entries = col.numel() + (n_rows if add_self else 0)
    rptr = torch.empty(n_nodes + 1, dtype=torch.int32, device=dev)
    # zero-filled: entries past rptr[n_nodes] (the buffers are sized by an upper bound) stay valid ids
    rdst = torch.zeros(max(n_entries, 1), dtype=torch.int32, device=dev)
    rslot = torch.zeros(max(n_entries, 1), dtype=torch.int32, device=dev)
    cursor = torch.empty(max(n_nodes, 1), dtype=torch.int32, device=dev)
    call("nc_reverse_csr", rowptr, col, n_nodes, n_rows, int(add_self), rptr, rdst, rslot, cursor)
    return rptr, rdst, rslot


def gatv2_bwd_ws_bytes(n_slots: int, n_dst: int, n_conv: int, heads: int) -> int:
    """Scratch of ``nc_gatv2_bwd``'s ring kernels (sign bits per (edge, element) + per-target constants)."""
    return int(_lib.lib().nc_gatv2_bwd_ws_bytes(int(n_slots), int(n_dst), int(n_conv), int(heads)))


# --------------------------------------------------------------------------- SpMM / gathers
def spmm_csr(x: torch.Tensor, rowptr, col, w, wself, skip_self: bool, n_rows: int,
             out_dtype: torch.dtype = torch.float32, out: Optional[torch.Tensor] = None,
             bias: Optional[torch.Tensor] = None, n_cols: Optional[int] = None) -> torch.Tensor:
    assert x.dim() == 2 and x.stride(1) == 1
    rowptr, col = _i32(rowptr), _i32(col)
    n_cols = x.size(1) if n_cols is None else n_cols
    if out is None:
        out = torch.empty(n_rows, n_cols, dtype=out_dtype, device=x.device)
    call("nc_spmm_csr", x, dtype_code(x.dtype), x.stride(0), n_cols, rowptr, col, w, wself, bias,
         int(skip_self), n_rows, out, dtype_code(out.dtype), out.stride(0))
    return out


def spmm_csr_store(store, n_id, rowptr, col, w, wself, skip_self: bool, n_rows: int,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    n_id, rowptr, col = _i32(n_id), _i32(rowptr), _i32(col)
    if out is None:
        out = torch.empty(n_rows, store.n_cols, dtype=torch.float32, device=rowptr.device)
    call("nc_spmm_csr_store", store.rowptr, store.col, store.val, store.n_cols, n_id, rowptr, col,
         w, wself, int(skip_self), n_rows, out, out.stride(0))
    return out


BG_MAX_CTAS = 0         # grid bound of the side-stream gather (0: the library's default, one CTA per two SMs)


def csr_gather_rows(store, n_id, n_rows: int, log1p: bool, dtype: torch.dtype, ld: Optional[int] = None,
                    background: bool = False, out: Optional[torch.Tensor] = None):
    """``background``: the low-footprint kernel for a side stream (``nc_csr_gather_rows_bg``)."""
    ld = store.n_cols if ld is None else ld
    n_id = _i32(n_id)
    if out is None:
        out = torch.empty(n_rows, ld, dtype=dtype, device=store.device)
    if background:
        call("nc_csr_gather_rows_bg", store.rowptr, store.col, store.val, store.n_cols, n_id, n_rows,
             int(log1p), out, dtype_code(dtype), ld, BG_MAX_CTAS)
        return out
    call("nc_csr_gather_rows", store.rowptr, store.col, store.val, store.n_cols, n_id, n_rows,
         int(log1p), out, dtype_code(dtype), ld)
    return out


def dense_gather_rows(x: torch.Tensor, n_id, n_rows: int, log1p: bool, dtype: torch.dtype,
                      ld: Optional[int] = None):
    assert x.dtype == torch.float32 and x.stride(1) == 1
    ld = x.size(1) if ld is None else ld
    n_id = _i32(n_id)
    out = torch.empty(n_rows, ld, dtype=dtype, device=x.device)
    call("nc_dense_gather_rows", x, x.stride(0), x.size(1), n_id, n_rows, int(log1p), out,
         dtype_code(dtype), ld)
    return out


# --------------------------------------------------------------------------- GEMM
def gemm(A: torch.Tensor, B: torch.Tensor, trans_a: bool = False, trans_b: bool = False,
         bias: Optional[torch.Tensor] = None, relu: bool = False, mask: Optional[torch.Tensor] = None,
         out: Optional[torch.Tensor] = None, out_dtype: Optional[torch.dtype] = None,
         accumulate: bool = False, M: Optional[int] = None, N: Optional[int] = None,
         K: Optional[int] = None) -> torch.Tensor:
    """C[M,N] (+)= act(opA(A) @ opB(B) + bias) (* mask).  A: [M,K] or (trans_a) [K,M];
    B: [N,K] (weight layout) or (trans_b) [K,N].  fp32 operands -> SIMT kernel, bf16 operands
    -> tcgen05 kernel.  M/N/K may be given to ignore padding columns/rows of the buffers."""
    assert A.dim() == 2 and B.dim() == 2 and A.stride(1) == 1 and B.stride(1) == 1
    assert A.dtype == B.dtype
    if M is None:
        M = A.size(1) if trans_a else A.size(0)
    if K is None:
        K = A.size(0) if trans_a else A.size(1)
    if N is None:
        N = B.size(1) if trans_b else B.size(0)
    if out is None:
        od = out_dtype or (torch.float32 if A.dtype == torch.float32 else torch.bfloat16)
        ldc = pad8(N)
        buf = torch.empty(M, ldc, dtype=od, device=A.device)
        out = buf[:, :N]
    assert out.stride(1) == 1
    ldmask = mask.stride(0) if mask is not None else 0
    if A.dtype == torch.float32:
        assert out.dtype == torch.float32
        call("nc_gemm_f32", A, A.stride(0), int(trans_a), B, B.stride(0), int(trans_b), out,
             out.stride(0), M, N, K, bias, int(relu), mask, ldmask, int(accumulate))
    elif A.dtype == torch.bfloat16:
        call("nc_gemm_bf16", A, A.stride(0), int(trans_a), B, B.stride(0), int(trans_b), out,
             dtype_code(out.dtype), out.stride(0), M, N, K, bias, int(relu), mask, ldmask,
             int(accumulate))
    else:
        raise TypeError(f"gemm: unsupported dtype {A.dtype}")
    return out
