"""
GATv2 edge kernels at hot-path sizes (bf16 rows: the streaming kernels; fp32 rows: the generic ones)
against a plain PyTorch fp32 reference of the same op evaluated on the same (bf16-rounded) inputs:
forward output + attention coefficients, and the backward w.r.t. x_l, x_r, att, bias through autograd.
Tolerance: the kernels compute in fp32 from identical inputs -> 2e-4 relative on the forward (online
softmax reassociation), bf16 rounding (1e-2) on the bf16-stored gradients.
"""
import pytest
import torch

from nichecompass_b200 import ops
from nichecompass_b200._lib import HostPtr, call, dtype_code
from tests import helpers as H

pytestmark = pytest.mark.gpu
DEV = "cuda"


def random_graph(n_src, n_dst, max_deg, g):
    """dst-CSR with 1..max_deg incoming entries per target, some of them self entries (which the
    kernels must drop and replace by exactly one self edge)."""
    rows, rowptr = [], [0]
    for i in range(n_dst):
        d = int(torch.randint(1, max_deg + 1, (1,), generator=g))
        nb = torch.randint(0, n_src, (d,), generator=g)
        if i % 3 == 0:
            nb[0] = i
        rows.append(nb)
        rowptr.append(rowptr[-1] + d)
    return torch.tensor(rowptr, dtype=torch.int32), torch.cat(rows).to(torch.int32)


def torch_reference(xl, xr, att, bias, rowptr, col, n_dst, Cs, offs, out_offs, Hh, slope=0.2):
    """PyG GATv2Conv message passing for all convs, fp32 torch ops."""
    outs = []
    dst = torch.repeat_interleave(torch.arange(n_dst, device=xl.device), (rowptr[1:] - rowptr[:-1]).long())
    src = col.long()
    keep = src != dst
    src = torch.cat([src[keep], torch.arange(n_dst, device=xl.device)])
    dst = torch.cat([dst[keep], torch.arange(n_dst, device=xl.device)])
    width = sum(Cs)
    out = torch.zeros(n_dst, width, device=xl.device)
    for C, off, oo in zip(Cs, offs, out_offs):
        a = xl[:, off:off + Hh * C].view(-1, Hh, C)
        b = xr[:, off:off + Hh * C].view(-1, Hh, C)
        w = att[off:off + Hh * C].view(1, Hh, C)
        e = (torch.nn.functional.leaky_relu(a[src] + b[dst], slope) * w).sum(-1)          # [E, H]
        m = torch.full((n_dst, Hh), -float("inf"), device=xl.device).scatter_reduce(0, dst[:, None].expand(-1, Hh), e, "amax")
        ex = torch.exp(e - m[dst])
        den = torch.zeros(n_dst, Hh, device=xl.device).index_add_(0, dst, ex) + 1e-16
        al = ex / den[dst]
        msg = a[src] * al[:, :, None]
        agg = torch.zeros(n_dst, Hh, C, device=xl.device).index_add_(0, dst, msg)
        out[:, oo:oo + C] = agg.mean(1) + bias[oo:oo + C]
    return out


@pytest.mark.parametrize("dtype,Cs,ring,n_dst,max_deg", [
    (torch.bfloat16, [1500, 100, 1500, 100], True, 120, 14),      # C2 widths: ring kernels <12, 1>
    (torch.bfloat16, [1500, 100, 1500, 100], True, 333, 40),      # more targets than SMs, rows longer than a warp chunk
    (torch.bfloat16, [1500, 100, 1500, 100], False, 120, 14),     # same shape on the generic / streaming kernels
    (torch.bfloat16, [300, 100, 300, 100], True, 200, 9),         # C1 widths: <4, 1>
    (torch.bfloat16, [1500, 1500], True, 64, 14),                 # no add-on programs: <12, 0>
    (torch.bfloat16, [256, 256, 256, 256], True, 64, 5),          # equal segments: <4, 4>
    (torch.bfloat16, [24, 8], True, 120, 14),
    (torch.bfloat16, [24, 8], False, 120, 14),
    (torch.float32, [60, 12], False, 120, 14)])
def test_gatv2_fwd_bwd_vs_torch(dtype, Cs, ring, n_dst, max_deg):
    g = torch.Generator().manual_seed(0)
    Hh, n_src = 4, max(300, n_dst)
    NC = len(Cs)
    widths = [Hh * C for C in Cs]
    offs = [sum(widths[:c]) for c in range(NC)]
    out_offs = [sum(Cs[:c]) for c in range(NC)]
    NL, ldl, out_w = sum(widths), ops.pad8(sum(widths)), sum(Cs)
    rowptr, col = random_graph(n_src, n_dst, max_deg, g)
    rowptr, col = rowptr.to(DEV), col.to(DEV)
    xl = (torch.randn(n_src, ldl, generator=g) * 0.5).to(dtype).to(DEV)
    xr = (torch.randn(n_dst, ldl, generator=g) * 0.5).to(dtype).to(DEV)
    att = (torch.randn(ldl, generator=g) * 0.3).to(DEV)
    bias = torch.randn(out_w, generator=g).to(DEV)
    seg_host = torch.tensor([[offs[c], Cs[c], out_offs[c], 0] for c in range(NC)], dtype=torch.int32)
    seg = seg_host.to(DEV)
    deg = int((rowptr[1:] - rowptr[:-1]).max())
    n_slots = col.numel() + n_dst
    out = torch.empty(n_dst, out_w, device=DEV)
    alpha = torch.full((n_slots * NC * Hh,), 7.0, device=DEV)
    head_ws = torch.full((n_dst, ldl), float("nan"), device=DEV) if ring else None
    call("nc_gatv2_fwd", xl, xr, dtype_code(dtype), ldl, att, bias, rowptr, col, n_dst, seg, HostPtr(seg_host), NC, Hh,
         deg, 0.2, out, out.stride(0), alpha, head_ws)
    if ring:        # the ring kernels ran (not the generic fallback): they filled the per-head outputs
        assert bool(torch.isfinite(head_ws[:, :NL]).all())
    xl32 = xl.float().requires_grad_(True)
    xr32 = xr.float().requires_grad_(True)
    att32 = att.clone().requires_grad_(True)
    bias32 = bias.clone().requires_grad_(True)
    ref = torch_reference(xl32, xr32, att32, bias32, rowptr, col, n_dst, Cs, offs, out_offs, Hh)
    assert H.rel_err(out, ref) < 2e-4
    # attention coefficients of every target sum to one per (conv, head); removed self entries carry 0
    al = alpha.view(n_slots, NC * Hh)
    slot_dst = torch.repeat_interleave(torch.arange(n_dst, device=DEV), (rowptr[1:] - rowptr[:-1]).long() + 1)
    sums = torch.zeros(n_dst, NC * Hh, device=DEV).index_add_(0, slot_dst, al)
    assert float((sums - 1).abs().max()) < 1e-4
    # backward
    dout = torch.randn(n_dst, out_w, generator=g).to(DEV)
    ref.backward(dout)
    rptr, rdst, rslot = ops.reverse_csr(rowptr, col, n_src, n_dst, True, n_slots)
    dxl = torch.empty_like(xl)
    dxr = torch.empty_like(xr)
    datt = torch.zeros(ldl, device=DEV)
    dbias = torch.zeros(out_w, device=DEV)
    ds = torch.empty_like(alpha)
    ws_bytes = ops.gatv2_bwd_ws_bytes(n_slots, n_dst, NC, Hh) if ring else 0
    bwd_ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV) if ring else None
    call("nc_gatv2_bwd", xl, xr, dtype_code(dtype), ldl, att, rowptr, col, n_dst, n_src, rptr, rdst, rslot, seg,
         HostPtr(seg_host), NC, Hh, deg, 0.2, dout, dout.stride(0), alpha, ds, dxl, dxr, datt, dbias, head_ws, bwd_ws,
         ws_bytes, n_slots)
    tol = 1e-2 if dtype == torch.bfloat16 else 2e-4
    print(f"gatv2 {dtype} Cs={Cs} ring={ring}: out {H.rel_err(out, ref):.2e} dxl {H.rel_err(dxl[:, :NL].float(), xl32.grad[:, :NL]):.2e} "
          f"dxr {H.rel_err(dxr[:, :NL].float(), xr32.grad[:, :NL]):.2e} datt {H.rel_err(datt[:NL], att32.grad[:NL]):.2e}")
    assert H.rel_err(dxl[:, :NL].float(), xl32.grad[:, :NL]) < tol
    assert H.rel_err(dxr[:, :NL].float(), xr32.grad[:, :NL]) < tol
    assert H.rel_err(datt[:NL], att32.grad[:NL]) < 2e-4
    assert H.rel_err(dbias, bias32.grad) < 1e-5
