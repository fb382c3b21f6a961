"""
Independent known-answer tests for the third-party arithmetic the reference calls but does not vendor
(torch_geometric GATv2Conv / GCNConv / gcn_norm(SparseTensor) / softmax) and for the
closed-form losses (negative binomial vs ``scipy.stats.nbinom.logpmf``, BCE-with-logits with pos_weight,
KL).  The expected numbers are written here -- as literals / closed-form expressions evaluated with
``math`` on python floats, or by per-target scalar loops that share no code with ``oracle/pyg_semantics.py``
-- so that the oracle (CPU, ``-m "not gpu"``) and the CUDA kernels (``-m gpu``) each face a witness that is
neither of them.  (PyG itself cannot be installed in this image: these KATs restate its DOCUMENTED
behaviour; they pin the oracle against a second, independent reading of that documentation.)
"""
import math

import numpy as np
import pytest
import torch

from oracle import pyg_semantics as S
from oracle import vgpgae_ref as V

SLOPE = 0.2


def lrelu(v):
    return v if v > 0 else SLOPE * v


# ------------------------------------------------------------------------------------------------ graphs
# 6 nodes; (source j -> target i).  Contains: an existing self loop (2 -> 2) that GATv2 / GCN must not count
# twice, a node without incoming edges (5), a multi-neighbour target (0) and an asymmetric pair (4 -> 3 only).
EDGES = [(1, 0), (2, 0), (3, 0), (0, 1), (2, 2), (1, 2), (4, 3), (0, 3), (3, 4), (0, 4)]
N = 6


def edge_index(dtype=torch.int64):
    return torch.tensor(EDGES, dtype=dtype).t().contiguous()


def incoming(i, with_self: str):
    """Source list of target i.  'replace': self loops removed, then exactly one added (GATv2Conv / GCNConv);
    'as-is': the edges as given (sum / attention aggregators)."""
    src = [j for (j, t) in EDGES if t == i]
    if with_self == "replace":
        src = [j for j in src if j != i] + [i]
    return src


# ------------------------------------------------------------------------------------------------ A.1
def gatv2_scalar(xl, xr, att, bias, H, C, want_alpha=False):
    """Per-target scalar loops over python floats (App. A.1)."""
    out = [[0.0] * C for _ in range(N)]
    alphas = {}
    for i in range(N):
        src = incoming(i, "replace")
        for h in range(H):
            sc = []
            for j in src:
                sc.append(sum(att[h][c] * lrelu(xl[j][h][c] + xr[i][h][c]) for c in range(C)))
            m = max(sc)
            ex = [math.exp(s - m) for s in sc]
            den = sum(ex) + 1e-16
            al = [e / den for e in ex]
            alphas[(i, h)] = al
            for c in range(C):
                out[i][c] += sum(a * xl[j][h][c] for a, j in zip(al, src)) / H
        for c in range(C):
            out[i][c] += bias[c]
    return (out, alphas) if want_alpha else out


def _gat_inputs(H=2, C=3, Din=4, seed=3):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(N, Din, generator=g, dtype=torch.float64)
    Wl = torch.randn(H * C, Din, generator=g, dtype=torch.float64) * 0.7
    Wr = torch.randn(H * C, Din, generator=g, dtype=torch.float64) * 0.7
    bl = torch.randn(H * C, generator=g, dtype=torch.float64) * 0.1
    br = torch.randn(H * C, generator=g, dtype=torch.float64) * 0.1
    att = torch.randn(1, H, C, generator=g, dtype=torch.float64)
    bias = torch.randn(C, generator=g, dtype=torch.float64)
    return x, Wl, bl, Wr, br, att, bias


def _lin(x, W, b, H, C):
    """x W^T + b as nested python lists [n][H][C] (scalar loops)."""
    xs, Ws, bs = x.tolist(), W.tolist(), b.tolist()
    return [[[sum(xs[n][d] * Ws[h * C + c][d] for d in range(len(xs[n]))) + bs[h * C + c] for c in range(C)]
             for h in range(H)] for n in range(len(xs))]


def test_gatv2_closed_form_three_nodes():
    """H = C = 1, numbers by hand.  Target 0 hears 1, 2 and itself; x_l = (-3, 2, 3), x_r[0] = 0.5, att = 2:
    scores 2 * lrelu(-2.5) = -1, 2 * 2.5 = 5, 2 * 3.5 = 7  ->  alpha = softmax(5, 7, -1) in edge order (1, 2, self);
    out[0] = alpha . (2, 3, -3) + bias."""
    ei = torch.tensor([[1, 2, 0], [0, 0, 0]])                    # 1 -> 0, 2 -> 0 and an existing self loop 0 -> 0
    xl = torch.tensor([[-3.0], [2.0], [3.0]], dtype=torch.float64)
    # lin_l = identity (1x1 weight 1), lin_r maps x -> fixed x_r through its bias with a zero weight
    out = S.gatv2_conv(xl, ei, torch.ones(1, 1, dtype=torch.float64), None,
                       torch.zeros(1, 1, dtype=torch.float64), torch.tensor([0.5], dtype=torch.float64),
                       torch.tensor([[[2.0]]], dtype=torch.float64), torch.tensor([0.25], dtype=torch.float64), heads=1)
    e = [math.exp(5 - 7), 1.0, math.exp(-1 - 7)]
    den = sum(e)
    want0 = (e[0] * 2 + e[1] * 3 + e[2] * -3) / den + 0.25
    assert abs(float(out[0, 0]) - want0) < 1e-14
    assert abs(want0 - 3.1290599657) < 1e-9                      # the literal, for the record
    # targets 1 and 2 have only their added self loop: alpha = 1, out = x_l[i] + bias
    assert abs(float(out[1, 0]) - 2.25) < 1e-14 and abs(float(out[2, 0]) - 3.25) < 1e-14


def test_gatv2_oracle_matches_scalar_loops():
    H, C = 2, 3
    x, Wl, bl, Wr, br, att, bias = _gat_inputs(H, C)
    out, (ei, alpha) = S.gatv2_conv(x, edge_index(), Wl, bl, Wr, br, att, bias, heads=H, return_alpha=True)
    want, al = gatv2_scalar(_lin(x, Wl, bl, H, C), _lin(x, Wr, br, H, C), att[0].tolist(), bias.tolist(), H, C, True)
    assert torch.allclose(out, torch.tensor(want, dtype=torch.float64), atol=1e-13, rtol=0)
    # attention coefficients edge by edge: PyG's order is (non-self edges as given) ++ (self loops 0..n-1)
    for e in range(ei.size(1)):
        j, i = int(ei[0, e]), int(ei[1, e])
        src = incoming(i, "replace")
        for h in range(H):
            assert abs(float(alpha[e, h]) - al[(i, h)][src.index(j)]) < 1e-14
    assert ei.size(1) == len([1 for (j, i) in EDGES if j != i]) + N


# ------------------------------------------------------------------------------------------------ A.2
def test_gcn_conv_closed_form():
    """deg[i] = #non-self incoming + 1: (4, 2, 2, 3, 3, 1).  out[i] = sum_j x'[j] / sqrt(deg j deg i) + bias."""
    deg = [len(incoming(i, "replace")) for i in range(N)]
    assert deg == [4, 2, 2, 3, 3, 1]
    g = torch.Generator().manual_seed(5)
    x = torch.randn(N, 3, generator=g, dtype=torch.float64)
    W = torch.randn(2, 3, generator=g, dtype=torch.float64)
    b = torch.tensor([0.5, -1.0], dtype=torch.float64)
    out = S.gcn_conv(x, edge_index(), W, b)
    xw = [[sum(x[n, d].item() * W[c, d].item() for d in range(3)) for c in range(2)] for n in range(N)]
    for i in range(N):
        for c in range(2):
            want = sum(xw[j][c] / math.sqrt(deg[j] * deg[i]) for j in incoming(i, "replace")) + b[c].item()
            assert abs(out[i, c].item() - want) < 1e-13
    # node 5 (isolated): only its self loop with weight 1 / deg = 1
    assert abs(out[5, 0].item() - (xw[5][0] + 0.5)) < 1e-13


# ------------------------------------------------------------------------------------------------ A.3
def test_one_hop_norm_aggregator_path_graph_literals():
    """Path 0 - 1 - 2 (both directions, no self loops given), x = (1, 2, 4): row sums after fill_diag = (2, 3, 2).
    x_nbr[0] = 1/2 + 2/sqrt6, x_nbr[1] = 1/sqrt6 + 2/3 + 4/sqrt6, x_nbr[2] = 2/sqrt6 + 4/2."""
    ei = torch.tensor([[0, 1, 1, 2], [1, 0, 2, 1]])
    x = torch.tensor([[1.0], [2.0], [4.0]], dtype=torch.float64)
    out, _ = S.one_hop_gcn_norm_aggregate(x, ei)
    r6 = math.sqrt(6.0)
    want = [0.5 + 2 / r6, 1 / r6 + 2 / 3 + 4 / r6, 2 / r6 + 2.0]
    assert torch.allclose(out[:, 0], torch.tensor(want, dtype=torch.float64), atol=1e-15, rtol=0)
    assert abs(want[0] - 1.3164965809) < 1e-9 and abs(want[1] - 2.7079081189) < 1e-9


def test_one_hop_norm_aggregator_uses_row_sums_of_an_asymmetric_graph():
    """deg = ROW sum (= out-degree + 1 after fill_diag) -- differs from the GCNConv in-degree on the asymmetric
    pair 4 -> 3.  adj_norm[r, c] = deg[r]^-1/2 deg[c]^-1/2, x_nbr[c] = sum_r adj_norm[r, c] x[r]."""
    out_deg = [len([1 for (j, i) in EDGES if j == v and i != v]) + 1 for v in range(N)]
    assert out_deg == [4, 3, 2, 3, 2, 1]
    g = torch.Generator().manual_seed(7)
    x = torch.rand(N, 2, generator=g, dtype=torch.float64)
    out, _ = S.one_hop_gcn_norm_aggregate(x, edge_index())
    for c_ in range(N):
        for f in range(2):
            want = sum(x[r, f].item() / math.sqrt(out_deg[r] * out_deg[c_]) for r in incoming(c_, "replace"))
            assert abs(out[c_, f].item() - want) < 1e-14
    s, _ = S.one_hop_sum_aggregate(x, edge_index())
    for c_ in range(N):
        assert abs(s[c_, 0].item() - sum(x[r, 0].item() for r in incoming(c_, "as-is"))) < 1e-14


# ------------------------------------------------------------------------------------------------ A.4
def attention_aggregate_scalar(x, Wl, Wr, attn, H):
    Fd = len(x[0])
    gl = [[[sum(x[n][d] * Wl[h * Fd + f][d] for d in range(Fd)) for f in range(Fd)] for h in range(H)] for n in range(N)]
    gr = [[[sum(x[n][d] * Wr[h * Fd + f][d] for d in range(Fd)) for f in range(Fd)] for h in range(H)] for n in range(N)]
    out = [[0.0] * Fd for _ in range(N)]
    for i in range(N):
        src = incoming(i, "as-is")
        if not src:
            continue
        for h in range(H):
            sc = [sum(attn[h][f] * lrelu(gr[i][h][f] + gl[j][h][f]) for f in range(Fd)) for j in src]
            m = max(sc)
            ex = [math.exp(s - m) for s in sc]
            den = sum(ex) + 1e-16
            for f in range(Fd):
                out[i][f] += sum(e / den * x[j][f] for e, j in zip(ex, src)) / H
    return out


def test_attention_aggregator_oracle_matches_scalar_loops():
    H, Fd = 2, 3
    g = torch.Generator().manual_seed(11)
    x = torch.rand(N, Fd, generator=g, dtype=torch.float64) * 3
    Wl = torch.randn(H * Fd, Fd, generator=g, dtype=torch.float64)
    Wr = torch.randn(H * Fd, Fd, generator=g, dtype=torch.float64)
    attn = torch.randn(1, H, Fd, generator=g, dtype=torch.float64)
    out, _ = S.one_hop_attention_aggregate(x, edge_index(), Wl, Wr, attn, heads=H)
    want = attention_aggregate_scalar(x.tolist(), Wl.tolist(), Wr.tolist(), attn[0].tolist(), H)
    assert torch.allclose(out, torch.tensor(want, dtype=torch.float64), atol=1e-13, rtol=0)
    assert float(out[5].abs().max()) == 0.0          # no incoming edge, no self-loop manipulation: zero row


# ------------------------------------------------------------------------------------------------ losses
def test_nb_loss_matches_scipy_nbinom_logpmf():
    """NB(mean mu, inverse dispersion theta) == scipy nbinom(n = theta, p = theta / (theta + mu));
    modules/losses.py:169-212 adds eps = 1e-8 inside three logarithms (relative effect < 1e-7 here)."""
    from scipy.stats import nbinom
    rng = np.random.default_rng(0)
    x = rng.poisson(3.0, size=(7, 9)).astype(np.float64)
    mu = rng.uniform(0.2, 8.0, size=(7, 9))
    theta = rng.uniform(0.3, 5.0, size=9)
    want = float(np.mean(-nbinom.logpmf(x, theta[None, :], theta[None, :] / (theta[None, :] + mu)).sum(-1)))
    got = float(V.nb_recon_loss(torch.from_numpy(x), torch.from_numpy(mu), torch.from_numpy(theta)))
    assert abs(got - want) < 1e-6 * abs(want)


def test_bce_pos_weight_closed_form():
    """logits (2, -1, 0.5, -3), labels (1, 1, 0, 0): pos_weight = 2/2 = 1 -> mean of softplus terms;
    with labels (1, 0, 0, 0): pos_weight = 3, loss = [3 softplus(-2) + softplus(-1) + softplus(.5) + softplus(-3)] / 4."""
    sp = lambda v: math.log1p(math.exp(v))                                        # noqa: E731
    lg = torch.tensor([2.0, -1.0, 0.5, -3.0], dtype=torch.float64)
    got = float(V.edge_recon_loss(lg, torch.tensor([1.0, 0.0, 0.0, 0.0], dtype=torch.float64)))
    want = (3 * sp(-2.0) + sp(-1.0) + sp(0.5) + sp(-3.0)) / 4
    assert abs(got - want) < 1e-14 and abs(want - 0.4291775141) < 1e-9
    # edge_incl drops the third pair BEFORE pos_weight is counted: pos_weight = 2
    incl = torch.tensor([True, True, False, True])
    got = float(V.edge_recon_loss(lg, torch.tensor([1.0, 0.0, 0.0, 0.0], dtype=torch.float64), incl))
    assert abs(got - (2 * sp(-2.0) + sp(-1.0) + sp(-3.0)) / 3) < 1e-14


def test_kl_closed_form():
    """-1/2 mean_rows sum_p (1 + 2 ls - mu^2 - exp(ls)^2); mu = (1, 0), ls = (0, ln 2): -1/2 (1 + 0 - 1 - 1 + 1 + 2 ln2 - 0 - 4)."""
    mu = torch.tensor([[1.0, 0.0]], dtype=torch.float64)
    ls = torch.tensor([[0.0, math.log(2.0)]], dtype=torch.float64)
    assert abs(float(V.kl_reg_loss(mu, ls)) - (-0.5 * (-1.0 + 2 * math.log(2.0) - 3.0))) < 1e-15


def test_cosine_decoder_literals():
    a = torch.tensor([[3.0, 4.0], [1.0, 0.0]], dtype=torch.float64)
    b = torch.tensor([[4.0, 3.0], [0.0, 2.0]], dtype=torch.float64)
    got = V.cosine_similarity_rows(a, b)
    assert abs(float(got[0]) - 24.0 / 25.0) < 1e-15 and abs(float(got[1])) < 1e-15


# ================================================================================================ GPU
@pytest.mark.gpu
def test_gatv2_kernels_match_scalar_loops():
    """nc_gatv2_fwd (generic fp32 kernels) on the 6-node graph against the scalar loops above."""
    from nichecompass_b200 import ops
    from nichecompass_b200._lib import HostPtr, call, dtype_code
    H, C = 2, 4                                                   # C % 4 == 0: also a legal ring shape in bf16
    x, Wl, bl, Wr, br, att, bias = _gat_inputs(H, C)
    xl_l, xr_l = _lin(x, Wl, bl, H, C), _lin(x, Wr, br, H, C)
    want, al = gatv2_scalar(xl_l, xr_l, att[0].tolist(), bias.tolist(), H, C, True)
    dev = "cuda"
    rowptr, col = ops.coo_to_dst_csr(edge_index().to(dev), N)
    ld = ops.pad8(H * C)
    xl = torch.zeros(N, ld, device=dev)
    xr = torch.zeros(N, ld, device=dev)
    xl[:, :H * C] = torch.tensor(xl_l).reshape(N, H * C).to(dev)
    xr[:, :H * C] = torch.tensor(xr_l).reshape(N, H * C).to(dev)
    attf = torch.zeros(ld, device=dev)
    attf[:H * C] = att.reshape(-1).float().to(dev)
    seg_host = torch.tensor([[0, C, 0, 0]], dtype=torch.int32)
    out = torch.empty(N, C, device=dev)
    n_slots = col.numel() + N
    alpha = torch.zeros(n_slots * H, device=dev)
    deg = int((rowptr[1:] - rowptr[:-1]).max())
    call("nc_gatv2_fwd", xl, xr, dtype_code(torch.float32), ld, attf, bias.float().to(dev), rowptr, col, N,
         seg_host.to(dev), HostPtr(seg_host), 1, H, deg, SLOPE, out, out.stride(0), alpha, None)
    assert torch.allclose(out.cpu().double(), torch.tensor(want, dtype=torch.float64), atol=2e-6, rtol=1e-6)
    # slots of target i: its CSR entries (removed self entries carry 0) then the added self edge
    rp, cl = rowptr.cpu().tolist(), col.cpu().tolist()
    a = alpha.view(n_slots, H).cpu()
    for i in range(N):
        src = incoming(i, "replace")
        for h in range(H):
            for t, e in enumerate(range(rp[i], rp[i + 1])):
                j = cl[e]
                exp_a = 0.0 if j == i else al[(i, h)][src.index(j)]
                assert abs(float(a[e + i, h]) - exp_a) < 2e-6
            assert abs(float(a[rp[i + 1] + i, h]) - al[(i, h)][-1]) < 2e-6


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["norm", "sum"])
def test_aggregator_kernels_match_closed_form(mode):
    from nichecompass_b200 import ops
    dev = "cuda"
    g = torch.Generator().manual_seed(7)
    x = torch.rand(N, 8, generator=g)
    rowptr, col = ops.coo_to_dst_csr(edge_index().to(dev), N)
    if mode == "norm":
        w, wself = ops.gcn_norm_weights(rowptr, col, N, N, True)
        out = ops.spmm_csr(x.to(dev), rowptr, col, w, wself, True, N)
        out_deg = [len([1 for (j, i) in EDGES if j == v and i != v]) + 1 for v in range(N)]
        want = [[sum(x[r, f].item() / math.sqrt(out_deg[r] * out_deg[c_]) for r in incoming(c_, "replace"))
                 for f in range(8)] for c_ in range(N)]
    else:
        out = ops.spmm_csr(x.to(dev), rowptr, col, None, None, False, N)
        want = [[sum(x[r, f].item() for r in incoming(c_, "as-is")) for f in range(8)] for c_ in range(N)]
    assert torch.allclose(out.cpu().double(), torch.tensor(want, dtype=torch.float64), atol=1e-6, rtol=1e-6)


@pytest.mark.gpu
def test_nb_kernel_matches_scipy_nbinom():
    """nc_nb_loss_fused (softmax over all F -> library size over the selected columns -> NB NLL) against
    scipy.stats.nbinom.logpmf evaluated on the same means."""
    from scipy.stats import nbinom
    from nichecompass_b200._lib import call
    rng = np.random.default_rng(1)
    nb, F = 5, 24
    logits = rng.normal(0, 1.5, size=(nb, F)).astype(np.float32)
    x = rng.poisson(2.5, size=(nb, F)).astype(np.float32)
    theta_param = rng.normal(0, 0.7, size=F).astype(np.float32)
    sel = (rng.uniform(size=F) < 0.7).astype(np.uint8)
    sm = np.exp(logits.astype(np.float64) - logits.max(1, keepdims=True))
    sm /= sm.sum(1, keepdims=True)
    lib = (x.astype(np.float64) * sel).sum(1, keepdims=True)
    mu = (lib * sm)[:, sel.astype(bool)]
    th = np.exp(theta_param.astype(np.float64))[sel.astype(bool)]
    xs = x.astype(np.float64)[:, sel.astype(bool)]
    want = float(np.mean(-nbinom.logpmf(xs, th[None, :], th[None, :] / (th[None, :] + mu)).sum(-1)))
    dev = "cuda"
    t = lambda a: torch.from_numpy(a).to(dev)                                      # noqa: E731
    acc = torch.zeros(1, dtype=torch.float64, device=dev)
    dlog = torch.empty(nb, F, device=dev)
    dth = torch.zeros(F, device=dev)
    means = torch.empty(nb, F, device=dev)
    call("nc_nb_loss_fused", t(logits), F, t(x), F, t(sel), t(theta_param), nb, F, 1.0 / nb, means, F, acc, dlog, F,
         dth, None, 0)
    assert abs(float(acc) - want) < 2e-5 * abs(want)


@pytest.mark.gpu
def test_bce_and_cosine_kernels_closed_form():
    from nichecompass_b200._lib import call
    dev = "cuda"
    sp = lambda v: math.log1p(math.exp(v))                                        # noqa: E731
    lg = torch.tensor([2.0, -1.0, 0.5, -3.0], device=dev)
    lab = torch.tensor([1.0, 0.0, 0.0, 0.0], device=dev)
    loss = torch.empty(1, device=dev)
    dl = torch.empty(4, device=dev)
    call("nc_bce_fwd_bwd", lg, lab, None, 4, 1.0, loss, dl, 1)
    assert abs(float(loss) - (3 * sp(-2.0) + sp(-1.0) + sp(0.5) + sp(-3.0)) / 4) < 1e-6
    # d/dlogit: positives pos_weight * (sigmoid - 1) / n, negatives sigmoid / n
    sig = lambda v: 1 / (1 + math.exp(-v))                                        # noqa: E731
    want = [3 * (sig(2.0) - 1) / 4, sig(-1.0) / 4, sig(0.5) / 4, sig(-3.0) / 4]
    assert torch.allclose(dl.cpu(), torch.tensor(want), atol=1e-7)
    incl = torch.tensor([1, 1, 0, 1], dtype=torch.uint8, device=dev)
    call("nc_bce_fwd_bwd", lg, lab, incl, 4, 1.0, loss, dl, 1)
    assert abs(float(loss) - (2 * sp(-2.0) + sp(-1.0) + sp(-3.0)) / 3) < 1e-6
    z = torch.tensor([[3.0, 4.0], [1.0, 0.0], [4.0, 3.0], [0.0, 2.0]], device=dev)
    logits = torch.empty(2, device=dev)
    norms = torch.empty(4, device=dev)
    call("nc_cosine_fwd", z, z.stride(0), 2, 2, logits, norms)
    assert abs(float(logits[0]) - 0.96) < 1e-6 and abs(float(logits[1])) < 1e-7
