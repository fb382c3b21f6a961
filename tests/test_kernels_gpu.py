"""GPU parity of the individual kernels (through the C ABI) against the CPU oracle / fp64 torch."""
import numpy as np
import pytest
import torch

from nichecompass_b200 import ops
from nichecompass_b200._lib import call
from nichecompass_b200.data.store import CSRFeatureStore
from oracle import pyg_semantics as S
from oracle import sampler_ref as R
from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _subgraph(n_cells=300, n_genes=50, seed=0, node_bs=24):
    adata = H.tiny_problem(n_cells=n_cells, n_genes=n_genes, k=5, seed=seed)
    torch.manual_seed(seed)
    x = H.dense_x(adata)
    ei = R.adjacency_to_edge_index(adata.obsp["spatial_connectivities"])
    mask = torch.ones(n_cells, dtype=torch.bool)
    b = next(iter(R.NodeLoaderRef(x, ei, n_cells, mask, node_bs)))
    return adata, b


@pytest.mark.parametrize("n_genes", [50, 1037])
@pytest.mark.parametrize("mode", ["norm", "sum"])
def test_aggregator_spmm_matches_oracle(mode, n_genes):
    adata, b = _subgraph(n_genes=n_genes)
    n_sub = b.x.size(0)
    ref = (S.one_hop_gcn_norm_aggregate if mode == "norm" else S.one_hop_sum_aggregate)(
        b.x.double(), b.edge_index)[0]
    ei = b.edge_index.to(DEV)
    rowptr, col = ops.coo_to_dst_csr(ei, n_sub)
    x = b.x.to(DEV)
    if mode == "norm":
        w, wself = ops.gcn_norm_weights(rowptr, col, n_sub, n_sub, use_out_degree=True)
        out = ops.spmm_csr(x, rowptr, col, w, wself, True, n_sub)
    else:
        out = ops.spmm_csr(x, rowptr, col, None, None, False, n_sub)
    assert H.rel_err(out, ref) < 1e-6
    # only the seed rows (what VGPGAE consumes)
    nb = b.batch_size
    if mode == "norm":
        out_b = ops.spmm_csr(x, rowptr, col, w, wself, True, nb)
        assert torch.equal(out_b, out[:nb])
    # the same aggregation straight from the CSR count store
    store = CSRFeatureStore.from_scipy(adata.X, DEV)
    n_id = b.n_id.to(DEV, torch.int32)
    if mode == "norm":
        out_s = ops.spmm_csr_store(store, n_id, rowptr, col, w, wself, True, nb)
    else:
        out_s = ops.spmm_csr_store(store, n_id, rowptr, col, None, None, False, nb)
    assert H.rel_err(out_s, ref[:nb]) < 1e-6


def test_spmm_bf16_and_unaligned():
    _, b = _subgraph(n_genes=61)
    n_sub = b.x.size(0)
    rowptr, col = ops.coo_to_dst_csr(b.edge_index.to(DEV), n_sub)
    xb = b.x.to(DEV).to(torch.bfloat16)
    ref = S.one_hop_sum_aggregate(xb.float().cpu().double(), b.edge_index)[0]
    out = ops.spmm_csr(xb, rowptr, col, None, None, False, n_sub)
    assert H.rel_err(out, ref) < 1e-6


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gather_rows(dtype):
    adata, b = _subgraph(n_genes=77)
    store = CSRFeatureStore.from_scipy(adata.X, DEV)
    n_id = b.n_id.to(DEV, torch.int32)
    ld = ops.pad8(77)
    out = ops.csr_gather_rows(store, n_id, n_id.numel(), True, dtype, ld)
    ref = torch.log(1 + b.x)
    tol = 1e-6 if dtype == torch.float32 else 4e-3
    assert H.rel_err(out[:, :77].float(), ref) < tol
    assert float(out[:, 77:].float().abs().max()) == 0.0
    xd = torch.tensor(np.asarray(adata.X.todense()), dtype=torch.float32, device=DEV)
    out2 = ops.dense_gather_rows(xd, n_id, n_id.numel(), True, dtype, ld)
    assert torch.equal(out2, out)
    out3 = ops.csr_gather_rows(store, None, 10, False, torch.float32)
    assert torch.equal(out3.cpu(), H.dense_x(adata)[:10])
    # the side-stream variant (no shared memory, bounded grid) and a store in pinned host memory: same bits.
    # No retry: a mismatch fails on the first pass and names the variant, the differing (row, column) pairs
    # and the global row ids behind them.
    def check(name, got, want, ids):
        bad = (got.float() != want.float()).nonzero()
        if bad.numel():
            rows = bad[:, 0].unique()[:8]
            raise AssertionError(
                f"test_gather_rows[{dtype}]: {name}: {bad.size(0)} differing elements; first (row, col): "
                f"{bad[:8].tolist()}; n_id of those rows: {ids[rows].tolist()}; got {got[bad[:4, 0], bad[:4, 1]].tolist()} "
                f"want {want[bad[:4, 0], bad[:4, 1]].tolist()}")

    check("side-stream variant, device store",
          ops.csr_gather_rows(store, n_id, n_id.numel(), True, dtype, ld, background=True), out, n_id)
    host = CSRFeatureStore.from_scipy(adata.X, DEV, location="host")
    assert host.location == "host" and host.val.is_pinned()
    check("side-stream variant, host store",
          ops.csr_gather_rows(host, n_id, n_id.numel(), True, dtype, ld, background=True), out, n_id)
    check("standard kernel, host store", ops.csr_gather_rows(host, n_id, n_id.numel(), True, dtype, ld), out, n_id)
    big = torch.arange(adata.n_obs, device=DEV, dtype=torch.int32).repeat(9)       # more rows than CTAs
    check("rows > CTAs", ops.csr_gather_rows(store, big, big.numel(), True, dtype, ld, background=True),
          ops.csr_gather_rows(store, big, big.numel(), True, dtype, ld), big)


def test_host_store_blocks_are_never_recycled_with_stale_contents():
    """A host-resident store that is dropped and rebuilt with DIFFERENT counts of the same shape (what bench.py
    and the trainer tests do between cases): every gather must see the new contents.  The blocks come from
    cudaHostAlloc / cudaFreeHost (data/store.py PinnedBlock), not from torch's caching host allocator, so a
    kernel launched through the C ABI can never overlap the next tenant's writes."""
    import scipy.sparse as sp
    rng = np.random.default_rng(0)
    n, F = 600, 77
    ld = ops.pad8(F)
    n_id = torch.from_numpy(rng.permutation(n).astype(np.int32)).to(DEV)
    for rep in range(12):
        dense = rng.poisson(0.3, size=(n, F)).astype(np.float32) * (rep + 1)
        host = CSRFeatureStore.from_scipy(sp.csr_matrix(dense), DEV, location="host")
        want = torch.from_numpy(dense).to(DEV)[n_id.long()]           # raw counts: a pure data-movement check
        for background in (True, False):
            got = ops.csr_gather_rows(host, n_id, n, False, torch.float32, ld, background=background)
            assert torch.equal(got[:, :F], want), (rep, background, int((got[:, :F] != want).sum()))
            assert float(got[:, F:].abs().max()) == 0.0
        del host                                   # no synchronize: nc_host_free itself waits for the device


def test_reverse_csr():
    _, b = _subgraph()
    n_sub, nb = b.x.size(0), b.batch_size
    rowptr, col = ops.coo_to_dst_csr(b.edge_index.to(DEV), n_sub)
    rptr, rdst, rslot = ops.reverse_csr(rowptr, col, n_sub, nb, True)
    rp, cl = rowptr.cpu().tolist(), col.cpu().tolist()
    expect = {j: [] for j in range(n_sub)}
    for i in range(nb):
        for e in range(rp[i], rp[i + 1]):
            if cl[e] != i:
                expect[cl[e]].append((e + i, i))
        expect[i].append((rp[i + 1] + i, i))
    rptr, rdst, rslot = rptr.cpu().tolist(), rdst.cpu().tolist(), rslot.cpu().tolist()
    for j in range(n_sub):
        got = [(rslot[p], rdst[p]) for p in range(rptr[j], rptr[j + 1])]
        assert got == sorted(expect[j]), j
    assert rptr[n_sub] == sum(len(v) for v in expect.values())


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
def test_gemm_f32(ta, tb):
    g = torch.Generator().manual_seed(0)
    M, N, K = 150, 93, 70
    A = torch.randn((K, M) if ta else (M, K), generator=g)
    B = torch.randn((K, N) if tb else (N, K), generator=g)
    bias = torch.randn(N, generator=g)
    mask = (torch.rand(M, N, generator=g) > 0.5).float()
    ref = torch.relu((A.t() if ta else A).double() @ (B if tb else B.t()).double() + bias.double()) * mask.double()
    out = ops.gemm(A.to(DEV), B.to(DEV), ta, tb, bias=bias.to(DEV), relu=True, mask=mask.to(DEV))
    assert H.rel_err(out, ref) < 1e-6
    out2 = ops.gemm(A.to(DEV), B.to(DEV), ta, tb, out=out.clone(), accumulate=True)
    ref2 = ref + (A.t() if ta else A).double() @ (B if tb else B.t()).double()
    assert H.rel_err(out2, ref2) < 1e-6


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("shape", [(128, 128, 64), (300, 200, 136), (1000, 1608, 520), (64, 5000, 1600),
                                   (64, 1608, 5000), (5000, 1500, 72)])
def test_gemm_bf16_tcgen05(ta, tb, shape):
    g = torch.Generator().manual_seed(1)
    M, N, K = shape
    def mk(r, c):
        buf = torch.randn(r, ops.pad8(c), generator=g).to(torch.bfloat16)
        return buf.to(DEV)[:, :c]
    A = mk(K, M) if ta else mk(M, K)
    B = mk(K, N) if tb else mk(N, K)
    bias = torch.randn(N, generator=g).to(DEV)
    Ad = (A.t() if ta else A).double().cpu()
    Bd = (B if tb else B.t()).double().cpu()
    ref = Ad @ Bd
    out = ops.gemm(A, B, ta, tb, out_dtype=torch.float32)
    assert H.rel_err(out, ref) < 1e-5, "plain"
    out_b = ops.gemm(A, B, ta, tb, bias=bias, relu=True)
    ref_b = torch.relu(ref + bias.double().cpu())
    assert out_b.dtype == torch.bfloat16
    assert H.rel_err(out_b.float(), ref_b) < 6e-3, "bias+relu bf16"
    mask = (torch.rand(M, N, generator=g) > 0.5).float().to(DEV)
    acc = out.clone()
    ops.gemm(A, B, ta, tb, mask=mask, out=acc, accumulate=True)
    assert H.rel_err(acc, ref + ref * mask.double().cpu()) < 1e-5, "mask+accumulate"


@pytest.mark.parametrize("ta,tb", [(False, False), (True, False), (False, True), (True, True)])
@pytest.mark.parametrize("shape", [(3900, 2500, 520), (8704, 1600, 1000)])
def test_gemm_bf16_cta_pair_tiles(ta, tb, shape):
    """Shapes large enough for the CTA-pair configuration (256 x 256 tiles, tcgen05.mma.cta_group::2): odd
    number of 128-row tiles (the pair's second CTA is entirely out of bounds), ragged N / K, every epilogue."""
    g = torch.Generator().manual_seed(3)
    M, N, K = shape
    def mk(r, c):
        return torch.randn(r, ops.pad8(c), generator=g).to(torch.bfloat16).to(DEV)[:, :c]
    A = mk(K, M) if ta else mk(M, K)
    B = mk(K, N) if tb else mk(N, K)
    bias = torch.randn(N, generator=g).to(DEV)
    ref = (A.t() if ta else A).float() @ (B if tb else B.t()).float()      # torch fp32 reference of the same op
    out = ops.gemm(A, B, ta, tb, out_dtype=torch.float32)
    assert H.rel_err(out, ref) < 1e-4, "plain"
    out_b = ops.gemm(A, B, ta, tb, bias=bias, relu=True)
    assert out_b.dtype == torch.bfloat16
    assert H.rel_err(out_b.float(), torch.relu(ref + bias)) < 6e-3, "bias+relu bf16"
    mask = (torch.rand(M, N, generator=g) > 0.5).float().to(DEV)
    acc = out.clone()
    ops.gemm(A, B, ta, tb, mask=mask, out=acc, accumulate=True)
    assert H.rel_err(acc, ref + ref * mask) < 1e-4, "mask+accumulate"


def test_gemm_bf16_large_persistent():
    g = torch.Generator().manual_seed(2)
    M, N, K = 4100, 3080, 1600
    A = torch.randn(M, K, generator=g).to(torch.bfloat16).to(DEV)
    B = torch.randn(N, K, generator=g).to(torch.bfloat16).to(DEV)
    out = ops.gemm(A, B, out_dtype=torch.float32)
    ref = A.float() @ B.float().t()          # torch fp32 reference of the same op
    assert H.rel_err(out, ref) < 1e-4


@pytest.mark.parametrize("nb", [64, 300])
def test_masked_sddmm_matches_dense(nb):
    """Sparse-mask weight gradient == (dlogits^T @ z) * mask of MaskedLinear's backward."""
    from nichecompass_b200.functional import mask_info
    g = torch.Generator().manual_seed(4)
    F, P = 500, 96
    mask = (torch.rand(F, P, generator=g) < 0.03).float().to(DEV)
    mask[7] = 0                                              # an empty feature row
    dl = torch.randn(nb, F, generator=g).to(DEV)
    z = torch.randn(nb, P, generator=g).to(DEV)
    info = mask_info(mask)
    assert info[0] == "sparse"
    dW = torch.full((F, P), 0.5, device=DEV)                 # accumulates
    dlT, zT = dl.t().contiguous(), z.t().contiguous()
    call("nc_masked_sddmm", dlT, dlT.stride(0), zT, zT.stride(0), nb, info[1], info[2], info[3], F, dW, dW.stride(0))
    ref = 0.5 + (dl.double().t() @ z.double()) * mask.double()
    assert H.rel_err(dW, ref) < 1e-5
    assert mask_info(torch.ones(4, 4, device=DEV))[0] == "none"
    assert mask_info((torch.rand(50, 50, generator=g) < 0.5).float().to(DEV))[0] == "dense"


def test_cat_covariates_contrastive_loss_matches_oracle():
    """modules/losses.py:15-120: value and logit gradient, two covariates (labels relabelled in place)."""
    from nichecompass_b200.modules.losses import compute_cat_covariates_contrastive_loss
    from oracle.vgpgae_ref import contrastive_loss
    g = torch.Generator().manual_seed(9)
    n = 200
    logits = (torch.rand(n, generator=g) * 2 - 1)
    labels = torch.cat((torch.ones(n // 2), torch.zeros(n // 2)))
    same = [torch.rand(n, generator=g) < 0.6, torch.rand(n, generator=g) < 0.4]
    lo = logits.double().clone().requires_grad_(True)
    ref = contrastive_loss(lo, labels.double().clone(), [s.clone() for s in same], 0.125, 0.25)
    ref.backward()
    lc = logits.to(DEV).requires_grad_(True)
    lab = labels.to(DEV).clone()
    out = compute_cat_covariates_contrastive_loss(lc, lab, [s.to(DEV) for s in same], 0.125, 0.25)
    out.backward()
    assert H.rel_err(out, ref) < 1e-5
    assert H.rel_err(lc.grad, lo.grad) < 1e-5


def test_store_kernels_on_rows_wider_than_shared_memory():
    """RNA + ATAC layout of config 4 (5 k genes + 50 k peaks = 55 k columns > the 51 200 floats a CTA can stage):
    the gather takes the zero-then-scatter kernel, the label aggregation runs in column chunks."""
    import scipy.sparse as sp
    rng = np.random.default_rng(3)
    n, F = 400, 55_003
    x = sp.random(n, F, density=0.004, format="csr", random_state=7, dtype=np.float32)
    x.data = np.ceil(x.data * 5).astype(np.float32)
    store = CSRFeatureStore.from_scipy(x, DEV)
    n_id = torch.from_numpy(rng.permutation(n)[:96].astype(np.int32)).to(DEV)
    dense = torch.from_numpy(np.asarray(x.todense())).to(DEV)
    ld = ops.pad8(F)
    got = ops.csr_gather_rows(store, n_id, 96, False, torch.float32, ld)
    assert torch.equal(got[:, :F], dense[n_id.long()]) and float(got[:, F:].abs().max()) == 0.0
    gotb = ops.csr_gather_rows(store, n_id, 96, True, torch.bfloat16, ld)
    assert H.rel_err(gotb[:, :F].float(), torch.log1p(dense[n_id.long()])) < 4e-3
    # aggregation over a random local graph of the 96 gathered cells
    g = torch.Generator().manual_seed(0)
    deg = torch.randint(1, 7, (32,), generator=g)
    rowptr = torch.zeros(33, dtype=torch.int32)
    rowptr[1:] = torch.cumsum(deg, 0)
    col = torch.randint(0, 96, (int(rowptr[-1]),), generator=g).to(torch.int32)
    w = torch.rand(col.numel(), generator=g)
    wself = torch.rand(32, generator=g)
    out = ops.spmm_csr_store(store, n_id, rowptr.to(DEV), col.to(DEV), w.to(DEV), wself.to(DEV), False, 32)
    xs = dense[n_id.long()].double().cpu()
    ref = torch.zeros(32, F, dtype=torch.float64)
    for i in range(32):
        for e in range(int(rowptr[i]), int(rowptr[i + 1])):
            ref[i] += float(w[e]) * xs[int(col[e])]
        ref[i] += float(wself[i]) * xs[i]
    assert H.rel_err(out, ref) < 1e-6
