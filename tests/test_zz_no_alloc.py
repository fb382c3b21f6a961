"""
``nc_gcn_norm_weights_noalloc``: the GCN-normalisation weights with the degree scratch in a static device array of
the library (no allocation inside the call; include/nichecompass_b200.h).  No counterpart in the reference.
Checked through the C ABI against ``nc_gcn_norm_weights`` (caller-provided
scratch): integer degrees and the same three kernels, so the outputs must be bit-identical (the tolerance the other
tests use for this kernel, 1e-6 against fp64, is asserted against plain torch as well).  Two calls in a row on
DIFFERENT graphs: the second one must not see the first one's degrees in the shared scratch.

Collected last on purpose (file name).
"""
import pytest
import torch

from nichecompass_b200._lib import call

pytestmark = pytest.mark.gpu
DEV = "cuda"
MAX_NODES = 1 << 20


def _random_csr(n_nodes, max_deg, seed):
    g = torch.Generator().manual_seed(seed)
    deg = torch.randint(0, max_deg + 1, (n_nodes,), generator=g)
    rowptr = torch.zeros(n_nodes + 1, dtype=torch.int64)
    rowptr[1:] = torch.cumsum(deg, 0)
    col = torch.randint(0, n_nodes, (int(rowptr[-1]),), generator=g)
    # some explicit self entries: their weight is 0 and they do not count towards a degree
    rows = torch.repeat_interleave(torch.arange(n_nodes), deg)
    flip = torch.rand(col.numel(), generator=g) < 0.05
    col[flip] = rows[flip]
    return rowptr.to(torch.int32).to(DEV), col.to(torch.int32).to(DEV), rows.to(DEV)


def _fp64_reference(rowptr, col, rows, n_nodes, n_rows, use_out_degree):
    c, r = col.long(), rows.long()
    nonself = c != r
    counted = c[nonself] if use_out_degree else r[nonself]
    deg = torch.bincount(counted, minlength=n_nodes).double() + 1.0
    keep = r < n_rows
    w = torch.where(nonself, deg[c].rsqrt() * deg[r].rsqrt(), torch.zeros((), dtype=torch.float64, device=col.device))
    return w[keep], 1.0 / deg[:n_rows]


def _both(rowptr, col, n_nodes, n_rows, use_out_degree):
    n_e = int(rowptr[n_rows])
    out = []
    for name in ("nc_gcn_norm_weights", "nc_gcn_norm_weights_noalloc"):
        w = torch.full((max(n_e, 1),), float("nan"), device=DEV)
        wself = torch.full((max(n_rows, 1),), float("nan"), device=DEV)
        if name == "nc_gcn_norm_weights":
            ws = torch.full((n_nodes,), 12345, dtype=torch.int32, device=DEV)        # dirty scratch on purpose
            call(name, rowptr, col, n_nodes, n_rows, int(use_out_degree), ws, w, wself)
        else:
            call(name, rowptr, col, n_nodes, n_rows, int(use_out_degree), w, wself)
        out.append((w[:n_e], wself[:n_rows]))
    torch.cuda.synchronize()
    return out


@pytest.mark.parametrize("use_out_degree", [True, False])
def test_noalloc_matches_existing_entry_point_twice_in_a_row(use_out_degree):
    # first call: the larger graph (leaves non-zero degrees in all of the scratch the second call uses)
    for n_nodes, n_rows, max_deg, seed in ((5000, 5000, 9, 0), (1200, 700, 6, 1)):
        rowptr, col, rows = _random_csr(n_nodes, max_deg, seed)
        (w0, s0), (w1, s1) = _both(rowptr, col, n_nodes, n_rows, use_out_degree)
        assert torch.equal(w0, w1) and torch.equal(s0, s1), (n_nodes, use_out_degree)
        w_ref, s_ref = _fp64_reference(rowptr, col, rows, n_nodes, n_rows, use_out_degree)
        assert float((w1.double() - w_ref).abs().max()) < 1e-6
        assert float((s1.double() - s_ref).abs().max()) < 1e-6
    # the same graph twice: identical answers (the scratch is cleared by every call)
    rowptr, col, rows = _random_csr(3000, 8, 2)
    a = _both(rowptr, col, 3000, 3000, use_out_degree)[1]
    b = _both(rowptr, col, 3000, 3000, use_out_degree)[1]
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])


def test_noalloc_allocates_nothing_and_rejects_oversize():
    rowptr, col, _ = _random_csr(2000, 7, 3)
    w = torch.empty(col.numel(), device=DEV)
    wself = torch.empty(2000, device=DEV)
    call("nc_gcn_norm_weights_noalloc", rowptr, col, 2000, 2000, 1, w, wself)       # resolves the symbol address
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    n_alloc0 = torch.cuda.memory_stats().get("allocation.all.allocated", 0)
    for _ in range(3):
        call("nc_gcn_norm_weights_noalloc", rowptr, col, 2000, 2000, 1, w, wself)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0                                    # no cudaMalloc behind our back
    assert torch.cuda.memory_stats().get("allocation.all.allocated", 0) == n_alloc0  # nor through torch
    with pytest.raises(Exception, match="exceeds the static scratch"):
        call("nc_gcn_norm_weights_noalloc", rowptr, col, MAX_NODES + 1, 0, 1, w, wself)
    # the limit itself is accepted (n_rows = 0: only the degree kernels run over the given rowptr range)
    big_rowptr = torch.zeros(MAX_NODES + 1, dtype=torch.int32, device=DEV)
    call("nc_gcn_norm_weights_noalloc", big_rowptr, col, MAX_NODES, 0, 1, w, wself)
    torch.cuda.synchronize()
    # and the next small call is still right
    (w0, s0), (w1, s1) = _both(rowptr, col, 2000, 2000, True)
    assert torch.equal(w0, w1) and torch.equal(s0, s1)
