"""
GATv2 edge kernels alone at the hot-path shape of a workload (default C2: 4 convs of 1500 / 100 / 1500 / 100
channels x 4 heads, ~1.09 k targets, ~8 k sub-graph rows, kNN-like degrees), bf16 rows.

    python tools/bench_gatv2.py [--iters 20] [--widths 1500,100,1500,100] [--n-dst 1090] [--n-src 8000] [--once]

Timing: CUDA events on the launching stream, L2 flushed (a 256 MB fill) before every timed launch.
Algorithmic bytes: forward = (pairs + n_dst) rows of NL bf16 read + n_dst * out_w fp32
written + slots * NC * H alpha fp32 written; backward = the same rows read once + n_dst + n_src rows of NL bf16
written (d x_r, d x_l) + alpha / d score traffic.  ``--once``: one forward + one backward, for ncu.
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nichecompass_b200 import ops  # noqa: E402
from nichecompass_b200._lib import HostPtr, call, dtype_code  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--widths", default="1500,100,1500,100")
    ap.add_argument("--n-dst", type=int, default=1090)
    ap.add_argument("--n-src", type=int, default=8000)
    ap.add_argument("--deg", type=int, default=10)
    ap.add_argument("--generic", action="store_true", help="first-generation kernels (no ring workspace)")
    ap.add_argument("--once", action="store_true")
    a = ap.parse_args()
    dev = "cuda"
    g = torch.Generator().manual_seed(0)
    Cs = [int(c) for c in a.widths.split(",")]
    Hh, NC = 4, len(Cs)
    widths = [Hh * C for C in Cs]
    offs = [sum(widths[:c]) for c in range(NC)]
    out_offs = [sum(Cs[:c]) for c in range(NC)]
    NL, out_w = sum(widths), sum(Cs)
    ldl = ops.pad8(NL)
    n_dst, n_src = a.n_dst, a.n_src
    degs = torch.randint(max(a.deg - 3, 1), a.deg + 4, (n_dst,), generator=g)
    rowptr = torch.zeros(n_dst + 1, dtype=torch.int32)
    rowptr[1:] = torch.cumsum(degs, 0)
    E = int(rowptr[-1])
    col = torch.randint(0, n_src, (E,), generator=g).to(torch.int32)
    rowptr, col = rowptr.to(dev), col.to(dev)
    xl = (torch.randn(n_src, ldl, generator=g) * 0.5).to(torch.bfloat16).to(dev)
    xr = (torch.randn(n_dst, ldl, generator=g) * 0.5).to(torch.bfloat16).to(dev)
    att = (torch.randn(ldl, generator=g) * 0.3).to(dev)
    bias = torch.randn(out_w, generator=g).to(dev)
    seg_host = torch.tensor([[offs[c], Cs[c], out_offs[c], 0] for c in range(NC)], dtype=torch.int32)
    seg = seg_host.to(dev)
    max_deg = int(degs.max())
    n_slots = E + n_dst
    out = torch.empty(n_dst, out_w, device=dev)
    alpha = torch.empty(n_slots * NC * Hh, device=dev)
    head_ws = None if a.generic else torch.empty(n_dst, ldl, device=dev)
    dout = torch.randn(n_dst, out_w, generator=g).to(dev)
    rptr, rdst, rslot = ops.reverse_csr(rowptr, col, n_src, n_dst, True, n_slots)
    dxl, dxr = torch.empty_like(xl), torch.empty_like(xr)
    datt = torch.zeros(ldl, device=dev)
    dbias = torch.zeros(out_w, device=dev)
    ds = torch.empty_like(alpha)
    ws_bytes = 0 if a.generic else ops.gatv2_bwd_ws_bytes(n_slots, n_dst, NC, Hh)
    bwd_ws = None if a.generic else torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def fwd():
        call("nc_gatv2_fwd", xl, xr, dtype_code(xl.dtype), ldl, att, bias, rowptr, col, n_dst, seg, HostPtr(seg_host),
             NC, Hh, max_deg, 0.2, out, out.stride(0), alpha, head_ws)

    def bwd():
        call("nc_gatv2_bwd", xl, xr, dtype_code(xl.dtype), ldl, att, rowptr, col, n_dst, n_src, rptr, rdst, rslot, seg,
             HostPtr(seg_host), NC, Hh, max_deg, 0.2, dout, dout.stride(0), alpha, ds, dxl, dxr, datt, dbias,
             head_ws, bwd_ws, ws_bytes, n_slots)

    if a.once:
        fwd()
        bwd()
        torch.cuda.synchronize()
        return
    for _ in range(3):
        fwd()
        bwd()
    torch.cuda.synchronize()

    def timed(fn):
        ts = []
        for _ in range(a.iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]

    t_f, t_b = timed(fwd), timed(bwd)
    row = NL * 2
    pairs = n_slots                                    # entries + one self edge per target (upper bound: self entries)
    b_f = (pairs + n_dst) * row + n_dst * out_w * 4 + n_slots * NC * Hh * 4
    b_b = (pairs + n_dst) * row + (n_dst + n_src) * row + n_dst * out_w * 4 + 3 * n_slots * NC * Hh * 4
    peak = 6468.0
    try:
        peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                           "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:  # noqa: BLE001
        pass
    for name, t, b in (("fwd", t_f, b_f), ("bwd", t_b, b_b)):
        print(f"gatv2 {name}: {t * 1e3:8.1f} us  {b / 1e6:7.1f} MB algorithmic  {b / t / 1e6:7.0f} GB/s  "
              f"frac {b / t / 1e6 / peak:.3f}  (n_dst {n_dst}, n_src {n_src}, slots {n_slots}, NL {NL}, "
              f"{'generic' if a.generic else 'ring'})")


if __name__ == "__main__":
    main()
