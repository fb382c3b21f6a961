"""
Differentiable building blocks of the VGPGAE hot path: ``torch.autograd.Function``s whose
forward and backward are calls into the C ABI (``nichecompass_b200/_lib.py``).  No tensor math is
done by PyTorch here beyond allocation, views and scalar bookkeeping.

Two precisions: ``torch.float32`` (SIMT GEMM, parity mode, 1e-5) and ``torch.bfloat16``
(activations + weights in bf16, tcgen05 GEMMs, fp32 accumulation / losses, 1e-3).
"""
from typing import List, Optional, Sequence

import torch

from . import ops
from ._lib import HostPtr, call, dtype_code
from .ops import pad8


# --------------------------------------------------------------------------- graph container
class SubgraphCSR:
    """Induced subgraph in dst-CSR form (int32, device).  ``n_dst``: the leading rows whose
    outputs are consumed (the seeds); ``max_deg``: upper bound of their row lengths."""

    def __init__(self, rowptr, col, n_nodes: int, n_dst: Optional[int] = None, max_deg: Optional[int] = None,
                 e_dst: Optional[int] = None):
        """``e_dst``: rowptr[n_dst] when the caller knows it on the host (the sampler reads it back with the
        sizes); bounds the attention slots by the LIVE entries instead of ``col.numel()``."""
        self.rowptr, self.col, self.n_nodes = rowptr, col, int(n_nodes)
        self.n_dst = self.n_nodes if n_dst is None else int(n_dst)
        self.e_dst = col.numel() if e_dst is None else min(int(e_dst), col.numel())
        if max_deg is None:
            max_deg = int((rowptr[1:] - rowptr[:-1]).max().item()) if self.n_nodes > 0 else 0
        self.max_deg = int(max_deg)
        self._rev = None
        self._norm = {}

    @classmethod
    def from_edge_index(cls, edge_index: torch.Tensor, n_nodes: int, n_dst: Optional[int] = None):
        rowptr, col = ops.coo_to_dst_csr(edge_index, n_nodes)
        return cls(rowptr, col, n_nodes, n_dst)

    @property
    def n_slots(self) -> int:
        """Attention slots of the first n_dst rows (entries + self edges; an upper bound when ``e_dst`` is not
        known)."""
        return self.e_dst + self.n_dst

    def reverse(self):
        if self._rev is None:
            self._rev = ops.reverse_csr(self.rowptr, self.col, self.n_nodes, self.n_dst, True, self.n_slots)
        return self._rev

    def gcn_weights(self, use_out_degree: bool, n_rows: Optional[int] = None):
        n_rows = self.n_dst if n_rows is None else n_rows
        key = (use_out_degree, n_rows)
        if key not in self._norm:
            self._norm[key] = ops.gcn_norm_weights(self.rowptr, self.col, self.n_nodes, n_rows, use_out_degree)
        return self._norm[key]


# --------------------------------------------------------------------------- packed weights
_PACK_CACHE = {}
_PARAM_EPOCH = 0
_KEY_COUNTER = [0]


def _purge_cache_key(k: int):
    for key in [q for q in _PACK_CACHE if q == k or (isinstance(q, tuple) and q and q[0] == k)]:
        _PACK_CACHE.pop(key, None)


def new_cache_key(owner=None) -> int:
    """Process-unique token for a module's packed-weight cache entries (``id()`` values are
    recycled after garbage collection and would let a new module hit a dead module's entry).
    ``owner``: the module; its packed copies are dropped from the cache when it is collected."""
    _KEY_COUNTER[0] += 1
    k = _KEY_COUNTER[0]
    if owner is not None:
        import weakref
        weakref.finalize(owner, _purge_cache_key, k)
    return k



def bump_param_epoch():
    """Called by code that updates parameters through raw pointers (the fused Adam kernel),
    which autograd's version counters cannot see: invalidates every packed weight copy."""
    global _PARAM_EPOCH
    _PARAM_EPOCH += 1


def _versions(ts: Sequence[Optional[torch.Tensor]]):
    return (_PARAM_EPOCH,) + tuple((t.data_ptr(), t._version) if t is not None else None for t in ts)


def arena_view(params: Sequence[Optional[torch.Tensor]], kind: str) -> Optional[torch.Tensor]:
    """1-D view over ``params`` inside a ``FlatAdam`` arena when they lie back to back there
    (``FlatAdam(groups=...)``): kind 'p' = fp32 parameters, 'g' = gradients (zeroed before every
    backward, kernels ACCUMULATE into it), 'b' = bf16 shadow kept current by the Adam kernel.
    ``None`` when the parameters are not arena residents (plain modules, frozen weights)."""
    if not params or params[0] is None:
        return None
    first = getattr(params[0], "_nc_arena", None)
    if first is None:
        return None
    opt, off = first
    if off % 8 != 0:
        return None
    o = off
    base_p, base_g = opt.flat_p.data_ptr(), opt.flat_g.data_ptr()
    for t in params:
        a = getattr(t, "_nc_arena", None) if t is not None else None
        if a is None or a[0] is not opt or a[1] != o or not t.requires_grad or t.data_ptr() != base_p + 4 * o:
            return None
        if kind == "g" and (t.grad is None or t.grad.data_ptr() != base_g + 4 * o):
            return None
        o += t.numel()
    if kind == "p":
        return opt.flat_p[off:o]
    if kind == "g":
        return opt.flat_g[off:o]
    if opt.flat_bf16 is None:
        return None
    if not opt.shadow_is_current():
        opt.refresh_shadow()
    return opt.flat_bf16[off:o]


def grad_view(params: Sequence[Optional[torch.Tensor]], needs: Sequence[bool]) -> Optional[torch.Tensor]:
    """Gradient arena view for a group of parameters that all need a gradient, else ``None``."""
    if not all(needs):
        return None
    return arena_view(params, "g")


def pack_rows(key, params: Sequence[torch.Tensor], dtype: torch.dtype, n_cols: int) -> torch.Tensor:
    """Stack 2-D parameters [r_i, n_cols] row-wise into one [sum r_i, pad8(n_cols)] buffer of
    ``dtype`` (bf16 shadow / fp32 packed copy); refreshed only when a parameter changed."""
    if n_cols % 8 == 0 and all(p.dim() == 2 and p.size(1) == n_cols for p in params):
        v = arena_view(params, "b" if dtype == torch.bfloat16 else "p")
        if v is not None:
            return v.view(-1, n_cols)
    ver = (_versions(params), dtype)
    hit = _PACK_CACHE.get(key)
    if hit is not None and hit[0] == ver:
        return hit[1]
    rows = sum(p.size(0) for p in params)
    ld = pad8(n_cols)
    buf = hit[1] if (hit is not None and hit[1].shape == (rows, ld) and hit[1].dtype == dtype) else \
        torch.zeros(rows, ld, dtype=dtype, device=params[0].device)
    r = 0
    with torch.no_grad():
        for p in params:
            p2 = p.detach().reshape(p.size(0), -1)
            if dtype == torch.bfloat16:
                call("nc_cast_bf16", p2, p2.stride(0), None, 0, p2.size(0), p2.size(1), buf[r:], ld, 0)
            else:
                buf[r:r + p2.size(0), :p2.size(1)].copy_(p2)
            r += p2.size(0)
    _PACK_CACHE[key] = (ver, buf)
    return buf


def pack_cols(key, params: Sequence[torch.Tensor], masks: Sequence[Optional[torch.Tensor]],
              dtype: torch.dtype) -> torch.Tensor:
    """Concatenate [F, c_i] parameters column-wise, each multiplied by its mask (MaskedLinear:
    weight * mask [* dynamic mask], nn/layercomponents.py:74-79), into [F, pad8(sum c_i)]."""
    ver = (_versions(params), _versions(masks), dtype)
    hit = _PACK_CACHE.get(key)
    if hit is not None and hit[0] == ver:
        return hit[1]
    F = params[0].size(0)
    ld = pad8(sum(p.size(1) for p in params))
    buf32 = torch.zeros(F, ld, dtype=torch.float32, device=params[0].device)
    c = 0
    with torch.no_grad():
        for p, m in zip(params, masks):
            dst = buf32[:, c:c + p.size(1)]
            if m is None:
                dst.copy_(p.detach())
            else:
                torch.mul(p.detach(), m, out=dst)
            c += p.size(1)
        if dtype == torch.bfloat16:
            buf = torch.empty(F, ld, dtype=torch.bfloat16, device=buf32.device)
            call("nc_cast_bf16", buf32, ld, None, 0, F, ld, buf, ld, 0)
        else:
            buf = buf32
    _PACK_CACHE[key] = (ver, buf)
    return buf


def _cat_f32(ts: Sequence[torch.Tensor], pad_to: Optional[int] = None) -> torch.Tensor:
    v = arena_view(ts, "p")
    if v is not None:
        return v
    flat = torch.cat([t.detach().reshape(-1).float() for t in ts])
    if pad_to is not None and flat.numel() < pad_to:
        flat = torch.cat([flat, flat.new_zeros(pad_to - flat.numel())])
    return flat.contiguous()


def _col_sum(x: torch.Tensor, n_cols: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """out[c] += sum_r x[r, c] (``out`` given: accumulate into it, e.g. a gradient arena view)."""
    if out is None:
        out = torch.zeros(n_cols, dtype=torch.float32, device=x.device)
    if x.size(0) > 0:
        call("nc_col_sum", x, dtype_code(x.dtype), x.stride(0), x.size(0), n_cols, out)
    return out


# --------------------------------------------------------------------------- GATv2 encoder
# ``W1`` value of the encoder functions meaning "``x_enc`` is the hidden tensor itself" (fc_l2 + BatchNorm in
# front of the convs, nn/encoders.py:180-182): no ReLU between the input and the message passing.
HIDDEN_INPUT = "hidden"


_CONV_SEG = {}


def _conv_segments(offs, Cs, out_offsets, dev):
    """(host, device) int32 [NC, 4] segment table of the stacked convs, built once per layout: no host-to-device
    copy inside a training step (a copy from pageable memory cannot be captured into a CUDA graph)."""
    key = (offs, Cs, out_offsets, str(dev))
    hit = _CONV_SEG.get(key)
    if hit is None:
        host = torch.tensor([[offs[c], Cs[c], out_offsets[c], 0] for c in range(len(Cs))], dtype=torch.int32)
        hit = _CONV_SEG[key] = (host, host.to(dev))
    return hit


class EncoderGATv2Fn(torch.autograd.Function):
    """fc_l1 + ReLU + the four GATv2Conv heads of nn/encoders.py:144-208 (n_fc_layers=1,
    n_layers=1) for the first ``graph.n_dst`` nodes: returns [n_dst, 2P] = (mu | logstd) with
    mu = conv_mu ++ addon_conv_mu and logstd = conv_logstd ++ addon_conv_logstd.

    ``convs`` order in ``tensors``: for each of the NC convs (lin_l.weight, lin_l.bias,
    lin_r.weight, lin_r.bias, att, bias); ``out_offsets[c]`` is the output column of conv c."""

    @staticmethod
    def forward(ctx, x_enc, graph: SubgraphCSR, n_feat: int, heads: int, out_offsets, out_width: int,
                cache_key, act, W1, b1, *tensors):
        """``W1 is None``: second conv layer of the 2-layer encoder -- ``x_enc`` is the fp32 output of
        ``conv_l1`` [n_sub, Hd] and hidden = relu(x_enc) (nn/encoders.py:191-194)."""
        NC = len(tensors) // 6
        conv = [tensors[6 * c:6 * c + 6] for c in range(NC)]
        ident = isinstance(W1, str)
        pre = W1 is None or ident
        ctx.ident = ident
        dev = x_enc.device
        n_sub, n_dst = graph.n_nodes, graph.n_dst
        assert x_enc.size(0) == n_sub and (pre or x_enc.dtype == act)
        Hd = x_enc.size(1) if pre else W1.size(0)
        Cs = [cv[5].numel() for cv in conv]
        widths = [heads * C for C in Cs]
        offs = [sum(widths[:c]) for c in range(NC)]
        NL = sum(widths)
        ldl = pad8(NL)
        seg_host, seg = _conv_segments(tuple(offs), tuple(Cs), tuple(out_offsets), dev)
        W1p = None if pre else pack_rows((cache_key, "W1"), [W1], act, W1.size(1))
        Wl = pack_rows((cache_key, "Wl"), [cv[0] for cv in conv], act, Hd)
        Wr = pack_rows((cache_key, "Wr"), [cv[2] for cv in conv], act, Hd)
        bl, br = _cat_f32([cv[1] for cv in conv]), _cat_f32([cv[3] for cv in conv])
        att = _cat_f32([cv[4] for cv in conv], pad_to=ldl)
        tight = list(out_offsets) == [sum(Cs[:c]) for c in range(NC)] and sum(Cs) == out_width
        bias_out = arena_view([cv[5] for cv in conv], "p") if tight else None
        if bias_out is None:
            bias_out = torch.zeros(out_width, dtype=torch.float32, device=dev)
            for c in range(NC):
                bias_out[out_offsets[c]:out_offsets[c] + Cs[c]] = conv[c][5].detach()
        ldh = pad8(Hd)
        Hbuf = torch.zeros(n_sub, ldh, dtype=act, device=dev) if ldh != Hd else torch.empty(n_sub, ldh, dtype=act, device=dev)
        if pre:
            xin = x_enc.detach().float().contiguous()
            if ident:
                Hbuf[:, :Hd].copy_(xin)
            else:
                call("nc_relu_cast", xin, xin.stride(0), n_sub, Hd, Hbuf, dtype_code(act), ldh)
        else:
            ops.gemm(x_enc, W1p, bias=b1.detach().float(), relu=True, out=Hbuf[:, :Hd], M=n_sub, N=Hd, K=n_feat)
        XL = torch.empty(n_sub, ldl, dtype=act, device=dev)
        XR = torch.empty(max(n_dst, 1), ldl, dtype=act, device=dev)
        ops.gemm(Hbuf, Wl, bias=bl, out=XL[:, :NL], M=n_sub, N=NL, K=Hd)
        ops.gemm(Hbuf, Wr, bias=br, out=XR[:, :NL], M=n_dst, N=NL, K=Hd)
        out = torch.empty(n_dst, out_width, dtype=torch.float32, device=dev)
        alpha = torch.empty(max(graph.n_slots, 1) * NC * heads, dtype=torch.float32, device=dev)
        # per-head outputs of the ring kernels (csrc/gatv2_ring.cu); untouched on the generic path
        head_ws = torch.empty(max(n_dst, 1), ldl, dtype=torch.float32, device=dev) if act == torch.bfloat16 else None
        call("nc_gatv2_fwd", XL, XR, dtype_code(act), ldl, att, bias_out, graph.rowptr, graph.col, n_dst,
             seg, HostPtr(seg_host), NC, heads, graph.max_deg, 0.2, out, out.stride(0), alpha, head_ws)
        ctx.save_for_backward(None if pre else x_enc, Hbuf, XL, XR, alpha, att, seg, W1p, Wl, Wr, head_ws)
        ctx.meta = (graph, n_feat, heads, out_offsets, out_width, NC, Cs, widths, offs, NL, ldl, Hd, seg_host,
                    [tuple(t.shape) for t in tensors], None if pre else tuple(W1.shape), act)
        ctx.params = (None if pre else W1, b1, conv, tight)          # arena residents receive their gradients in place
        return out

    @staticmethod
    def backward(ctx, dout):
        x_enc, Hbuf, XL, XR, alpha, att, seg, W1p, Wl, Wr, head_ws = ctx.saved_tensors
        (graph, n_feat, heads, out_offsets, out_width, NC, Cs, widths, offs, NL, ldl, Hd, seg_host,
         shapes, w1_shape, act) = ctx.meta
        pre = W1p is None
        dev = Hbuf.device
        n_sub, n_dst = graph.n_nodes, graph.n_dst
        dout = dout.contiguous().float()
        rptr, rdst, rslot = graph.reverse()
        dXL = torch.empty(n_sub, ldl, dtype=act, device=dev)
        dXR = torch.empty(max(n_dst, 1), ldl, dtype=act, device=dev)
        if ldl != NL:
            dXL[:, NL:].zero_()
            dXR[:, NL:].zero_()
        W1, b1, conv, tight = ctx.params
        need = ctx.needs_input_grad          # (x_enc, graph, n_feat, heads, offs, width, key, act, W1, b1, *tensors)
        need_c = [need[10 + 6 * c:16 + 6 * c] for c in range(NC)]

        def gv(slot):
            return grad_view([cv[slot] for cv in conv], [nc[slot] for nc in need_c])
        g_bl, g_br, g_att = gv(1), gv(3), gv(4)
        g_Wl, g_Wr = (gv(0), gv(2)) if Hd % 4 == 0 else (None, None)       # GEMM C rows must stay 16-byte aligned
        g_bias = gv(5) if tight else None
        g_W1 = grad_view([W1], [need[8]]) if (not pre and n_feat == W1.size(1) and n_feat % 4 == 0) else None
        g_b1 = None if pre else grad_view([b1], [need[9]])
        datt = g_att if g_att is not None else torch.zeros(ldl, dtype=torch.float32, device=dev)
        dbias = g_bias if g_bias is not None else torch.zeros(out_width, dtype=torch.float32, device=dev)
        ds = torch.empty_like(alpha)
        bwd_ws, ws_bytes = None, 0
        if head_ws is not None:
            ws_bytes = ops.gatv2_bwd_ws_bytes(graph.n_slots, n_dst, NC, heads)
            bwd_ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        call("nc_gatv2_bwd", XL, XR, dtype_code(act), ldl, att, graph.rowptr, graph.col, n_dst, n_sub,
             rptr, rdst, rslot, seg, HostPtr(seg_host), NC, heads, graph.max_deg, 0.2, dout, dout.stride(0), alpha, ds,
             dXL, dXR, datt, dbias, head_ws, bwd_ws, ws_bytes, graph.n_slots)
        # weight gradients of lin_l / lin_r (all convs in one GEMM each), bias gradients
        if g_Wl is not None:
            ops.gemm(dXL, Hbuf, trans_a=True, trans_b=True, out=g_Wl.view(NL, Hd), accumulate=True, M=NL, N=Hd, K=n_sub)
            dWl = None
        else:
            dWl = ops.gemm(dXL, Hbuf, trans_a=True, trans_b=True, out_dtype=torch.float32, M=NL, N=Hd, K=n_sub)
        if g_Wr is not None:
            if n_dst > 0:
                ops.gemm(dXR, Hbuf, trans_a=True, trans_b=True, out=g_Wr.view(NL, Hd), accumulate=True,
                         M=NL, N=Hd, K=n_dst)
            dWr = None
        else:
            dWr = ops.gemm(dXR, Hbuf, trans_a=True, trans_b=True, out_dtype=torch.float32, M=NL, N=Hd, K=n_dst) \
                if n_dst > 0 else torch.zeros(NL, Hd, device=dev)
        dbl = _col_sum(dXL, NL, out=g_bl)
        dbr = _col_sum(dXR[:n_dst], NL, out=g_br)
        # dH = dXL @ Wl (+ dXR @ Wr on the seed rows)
        dH = torch.empty(n_sub, pad8(Hd), dtype=torch.float32, device=dev)
        ops.gemm(dXL, Wl, trans_b=True, out=dH[:, :Hd], M=n_sub, N=Hd, K=NL)
        if n_dst > 0:
            ops.gemm(dXR, Wr, trans_b=True, out=dH[:n_dst, :Hd], accumulate=True, M=n_dst, N=Hd, K=NL)
        dHp = None
        if not ctx.ident:
            dHp = torch.empty(n_sub, Hbuf.size(1), dtype=act, device=dev)
            call("nc_relu_bwd", dH, dH.stride(0), Hbuf, dtype_code(act), Hbuf.stride(0), n_sub, Hd, dHp, dHp.stride(0))
        if pre:
            dW1 = db1 = None
        elif g_W1 is not None:
            ops.gemm(dHp, x_enc, trans_a=True, trans_b=True, out=g_W1.view(Hd, n_feat), accumulate=True,
                     M=Hd, N=n_feat, K=n_sub)
            dW1 = None
        else:
            dW1 = ops.gemm(dHp, x_enc, trans_a=True, trans_b=True, out_dtype=torch.float32, M=Hd, N=n_feat, K=n_sub)
        if not pre:
            db1 = _col_sum(dHp, Hd, out=g_b1)
        grads: List[Optional[torch.Tensor]] = []
        for c in range(NC):
            sl = slice(offs[c], offs[c] + widths[c])
            grads += [None if g_Wl is not None else dWl[sl].reshape(shapes[6 * c]),
                      None if g_bl is not None else dbl[sl],
                      None if g_Wr is not None else dWr[sl].reshape(shapes[6 * c + 2]),
                      None if g_br is not None else dbr[sl],
                      None if g_att is not None else datt[sl].reshape(shapes[6 * c + 4]),
                      None if g_bias is not None else dbias[out_offsets[c]:out_offsets[c] + Cs[c]]]
        if g_b1 is not None:
            db1 = None
        if dW1 is not None and dW1.shape != w1_shape:
            dW1 = dW1.contiguous()
        dx = None
        if pre:
            if ctx.needs_input_grad[0]:      # d relu(x): gradient of conv_l1's output (or of the hidden tensor itself)
                dx = dH[:, :Hd].contiguous() if ctx.ident else dHp[:, :Hd].float()
        elif ctx.needs_input_grad[0]:
            # only the trailing (embedding) columns of the encoder input can require grad
            dxf = ops.gemm(dHp, W1p, trans_b=True, out_dtype=torch.float32, M=n_sub, N=n_feat, K=Hd)
            dx = torch.zeros_like(x_enc, dtype=torch.float32)
            dx[:, :n_feat] = dxf
            dx = dx.to(x_enc.dtype)
        return (dx, None, None, None, None, None, None, None, dW1, db1, *grads)


class LinearActFn(torch.autograd.Function):
    """``relu?(x @ W.T + b)`` as fp32 [n, N] for the fully connected layers in front of the convs when the
    encoder is not the single fused pass (n_fc_layers == 2 or dropout, nn/encoders.py:179-182).  ``x``:
    [n, >= K] in the compute dtype (padding columns ignored) or fp32 (cast here)."""

    @staticmethod
    def forward(ctx, x, W, b, act, relu, cache_key):
        n, (N, K) = x.size(0), W.shape
        xa = x.detach()
        if xa.dtype != act or xa.stride(1) != 1 or xa.stride(0) % 8:
            buf = torch.zeros(n, pad8(K), dtype=act, device=x.device)
            buf[:, :K] = xa[:, :K]
            xa = buf
        Wp = pack_rows((cache_key, "W"), [W], act, K)
        out = ops.gemm(xa, Wp, bias=b.detach().float(), relu=relu, out_dtype=torch.float32, M=n, N=N, K=K)
        ctx.save_for_backward(xa, Wp, out if relu else None)
        ctx.meta = (act, K, N, x.dtype, x.size(1))
        return out

    @staticmethod
    def backward(ctx, dout):
        xa, Wp, out = ctx.saved_tensors
        act, K, N, x_dtype, x_cols = ctx.meta
        n = xa.size(0)
        d = dout.float()
        if out is not None:
            d = d * (out > 0)
        da = torch.zeros(n, pad8(N), dtype=act, device=d.device)
        da[:, :N] = d
        dW = ops.gemm(da, xa, trans_a=True, trans_b=True, out_dtype=torch.float32, M=N, N=K, K=n) \
            if ctx.needs_input_grad[1] else None
        db = d.sum(0) if ctx.needs_input_grad[2] else None
        dx = None
        if ctx.needs_input_grad[0]:
            dxf = ops.gemm(da, Wp, trans_b=True, out_dtype=torch.float32, M=n, N=K, K=N)
            dx = torch.zeros(n, x_cols, dtype=torch.float32, device=d.device)
            dx[:, :K] = dxf
            dx = dx.to(x_dtype)
        return dx, (None if dW is None else dW.contiguous()), db, None, None, None


# --------------------------------------------------------------------------- latent plumbing
class GatherRowsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, src, idx):
        src = src.contiguous()
        out = torch.empty(idx.numel(), src.size(1), dtype=torch.float32, device=src.device)
        call("nc_gather_rows_f32", src, src.stride(0), idx, idx.numel(), src.size(1), out, out.stride(0))
        ctx.save_for_backward(idx)
        ctx.n_src = src.size(0)
        return out

    @staticmethod
    def backward(ctx, dout):
        (idx,) = ctx.saved_tensors
        dout = dout.contiguous()
        dsrc = torch.zeros(ctx.n_src, dout.size(1), dtype=torch.float32, device=dout.device)
        call("nc_scatter_add_rows_f32", dout, dout.stride(0), idx, idx.numel(), dout.size(1), dsrc, dsrc.stride(0))
        return dsrc, None


class SplitLatentFn(torch.autograd.Function):
    """(mu | logstd) [R, 2P] -> mu, logstd views.  Autograd's own slicing would materialise two
    zero-filled [R, 2P] gradients and add them; here the backward is one concatenation."""

    @staticmethod
    def forward(ctx, ml, P):
        ctx.P, ctx.shape = P, tuple(ml.shape)
        return ml[:, :P], ml[:, P:]

    @staticmethod
    def backward(ctx, dmu, dls):
        R, W = ctx.shape
        if dmu is None and dls is None:
            return None, None
        ref = dmu if dmu is not None else dls
        if dmu is None:
            dmu = ref.new_zeros(R, ctx.P)
        if dls is None:
            dls = ref.new_zeros(R, W - ctx.P)
        return torch.cat((dmu, dls), dim=1), None


class ReparamFn(torch.autograd.Function):
    """z = (mu + eps * exp(logstd)) with inactive gene programs zeroed
    (modules/vgaemodulemixin.py:34-41, modules/vgpgae.py:632-637)."""

    @staticmethod
    def forward(ctx, mu, logstd, eps, active):
        R, P = mu.shape
        z = torch.empty(R, P, dtype=torch.float32, device=mu.device)
        call("nc_reparam_fwd", mu, mu.stride(0), logstd, logstd.stride(0), eps, eps.stride(0) if eps is not None else 0,
             active, R, P, z, z.stride(0), None, 0)
        ctx.save_for_backward(logstd, eps, active)
        return z

    @staticmethod
    def backward(ctx, dz):
        logstd, eps, active = ctx.saved_tensors
        dz = dz.contiguous()
        R, P = dz.shape
        dmu = torch.empty_like(dz)
        dls = torch.empty_like(dz)
        call("nc_reparam_bwd", logstd, logstd.stride(0), eps, eps.stride(0) if eps is not None else 0, active, R, P,
             dz, dz.stride(0), dmu, dmu.stride(0), dls, dls.stride(0))
        return dmu, dls, None, None


class KLFn(torch.autograd.Function):
    """modules/losses.py:255-284."""

    @staticmethod
    def forward(ctx, mu, logstd):
        R, P = mu.shape
        acc = torch.zeros(1, dtype=torch.float64, device=mu.device)
        dmu = torch.empty(R, P, dtype=torch.float32, device=mu.device)
        dls = torch.empty(R, P, dtype=torch.float32, device=mu.device)
        call("nc_kl_fwd_bwd", mu, mu.stride(0), logstd, logstd.stride(0), R, P, 1.0, acc, dmu, P, dls, P)
        ctx.save_for_backward(dmu, dls)
        return acc.float().reshape(())

    @staticmethod
    def backward(ctx, g):
        dmu, dls = ctx.saved_tensors
        return dmu * g, dls * g


class CosineFn(torch.autograd.Function):
    """nn/decoders.py:56-62 + nn/utils.py:8-36: rows [0, n) of z against rows [n, 2n)."""

    @staticmethod
    def forward(ctx, z):
        z = z.contiguous()
        n = z.size(0) // 2
        logits = torch.empty(n, dtype=torch.float32, device=z.device)
        norms = torch.empty(2 * n, dtype=torch.float32, device=z.device)
        call("nc_cosine_fwd", z, z.stride(0), z.size(1), n, logits, norms)
        ctx.save_for_backward(z, logits, norms)
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        z, logits, norms = ctx.saved_tensors
        n = z.size(0) // 2
        dz = torch.empty_like(z)
        call("nc_cosine_bwd", z, z.stride(0), z.size(1), n, logits, norms, dlogits.contiguous(), dz, dz.stride(0))
        return dz


class BCEFn(torch.autograd.Function):
    """modules/losses.py:123-166 (pos_weight = #neg / #pos over the included edges)."""

    @staticmethod
    def forward(ctx, logits, labels, incl, use_pos_weight=True):
        n = logits.numel()
        loss = torch.empty(1, dtype=torch.float32, device=logits.device)
        dlog = torch.empty(n, dtype=torch.float32, device=logits.device)
        call("nc_bce_fwd_bwd", logits.contiguous(), labels.contiguous().float(), incl, n, 1.0, loss, dlog,
             int(bool(use_pos_weight)))
        ctx.save_for_backward(dlog)
        return loss.reshape(())

    @staticmethod
    def backward(ctx, g):
        (dlog,) = ctx.saved_tensors
        return dlog * g, None, None, None


# --------------------------------------------------------------------------- masked decoder
def _block_layout(widths: Sequence[int]):
    """Column offsets of the (prior | add-on | embedding) blocks in the packed operands: every
    block starts on a multiple of 8 elements so that each is a legal TMA operand on its own."""
    offs, c = [], 0
    for w in widths:
        offs.append(c)
        c += pad8(w)
    return offs, c


def pack_decoder_weight(key, params, masks, dtype, offs, Kp) -> torch.Tensor:
    """[F, Kp] packed ``weight * mask [* dynamic mask]`` (nn/layercomponents.py:74-79), refreshed
    only when a parameter / mask changed."""
    ver = (_versions(params), _versions(masks), dtype)
    hit = _PACK_CACHE.get(key)
    if hit is not None and hit[0] == ver:
        return hit[1]
    F = params[0].size(0)
    buf = hit[1] if (hit is not None and hit[1].shape == (F, Kp) and hit[1].dtype == dtype) else \
        torch.zeros(F, Kp, dtype=dtype, device=params[0].device)
    with torch.no_grad():
        for p_, m, o in zip(params, masks, offs):
            w = p_.size(1)
            pd = p_.detach()
            if dtype == torch.bfloat16:
                call("nc_cast_bf16", pd, pd.stride(0), m, m.stride(0) if m is not None else 0, F, w,
                     buf[:, o:], Kp, pad8(w))
            elif m is None:
                buf[:, o:o + w].copy_(pd)
            else:
                torch.mul(pd, m, out=buf[:, o:o + w])
    _PACK_CACHE[key] = (ver, buf)
    return buf


_MASK_INFO = {}
_ZA_CACHE = {}


def mask_info(mask: Optional[torch.Tensor]):
    """Structure of a MaskedLinear mask [F, P] (cached per tensor version; one host read-back when a
    mask is first seen; the entry pins the tensor so that a freed mask's address cannot alias a new
    one): ("none",) for an absent / all-ones mask, ("sparse", rowptr, col, val) with the
    CSR-by-feature form when fewer than 10 % of the entries are set, else ("dense",)."""
    if mask is None:
        return ("none",)
    key = (mask.data_ptr(), mask._version, tuple(mask.shape))
    hit = _MASK_INFO.pop(key, None)
    if hit is not None:
        _MASK_INFO[key] = hit                  # most recently used last
        return hit[0]
    while len(_MASK_INFO) >= 256:              # least recently used first; never the masks of the running step
        _MASK_INFO.pop(next(iter(_MASK_INFO)))
    with torch.no_grad():
        nz = mask != 0
        nnz = int(nz.sum())
        if nnz == mask.numel() and bool((mask == 1).all()):
            info = ("none",)
        elif nnz < 0.1 * mask.numel():
            idx = nz.nonzero()                                    # row-major: sorted by feature, then program
            rowptr = torch.zeros(mask.size(0) + 1, dtype=torch.int32, device=mask.device)
            rowptr[1:] = torch.cumsum(torch.bincount(idx[:, 0], minlength=mask.size(0)), 0)
            info = ("sparse", rowptr, idx[:, 1].to(torch.int32).contiguous(),
                    mask[idx[:, 0], idx[:, 1]].float().contiguous())
        else:
            info = ("dense",)
    _MASK_INFO[key] = (info, mask)            # holding the tensor keeps its pointer from being recycled
    return info


_EFF_MASK = {}


def _effective_mask(m: torch.Tensor, d: torch.Tensor) -> torch.Tensor:
    """static mask * dynamic mask, cached per pair of tensor versions (a new product every call would also defeat
    ``mask_info``'s cache: one host read-back and one CSR build per decoder and iteration)."""
    key = (m.data_ptr(), m._version, d.data_ptr(), d._version, tuple(m.shape))
    hit = _EFF_MASK.pop(key, None)
    if hit is None:
        while len(_EFF_MASK) >= 16:
            _EFF_MASK.pop(next(iter(_EFF_MASK)))
        hit = (m * d, m, d)                    # operands kept: their addresses cannot be recycled while the entry lives
    _EFF_MASK[key] = hit
    return hit[0]


class MaskedDecoderFn(torch.autograd.Function):
    """logits = cat(z, embed) @ (W * mask)^T for the decoder weight blocks
    [F, P_prior | P_addon | n_embed] (nn/layers.py:83-142, nn/layercomponents.py:53-81) as ONE
    GEMM over the packed operands.  The GP mask is re-applied to each block's weight gradient in
    the GEMM epilogue; arena-resident weights receive it in place."""

    @staticmethod
    def forward(ctx, zcat, act_dtype, cache_key, dyn_masks, *params_and_masks):
        n = len(params_and_masks) // 2
        params, masks = params_and_masks[:n], params_and_masks[n:]
        eff_masks = []
        for m, d in zip(masks, dyn_masks):
            if m is not None and d is not None:
                m = _effective_mask(m, d)
            elif d is not None:
                m = d
            eff_masks.append(m)
        widths = [p.size(1) for p in params]
        offs, Kp = _block_layout(widths)
        Wp = pack_decoder_weight(cache_key, params, eff_masks, act_dtype, offs, Kp)
        F = params[0].size(0)
        nb = zcat.size(0)
        # the target and the source decoder contract the same z: pack it once
        zkey = (zcat.data_ptr(), zcat._version, act_dtype, tuple(widths), tuple(zcat.shape))
        if _ZA_CACHE.get("key") == zkey:
            za = _ZA_CACHE["za"]
        else:
            za = torch.zeros(nb, Kp, dtype=act_dtype, device=zcat.device)
            zd, c = zcat.detach(), 0
            for w, o in zip(widths, offs):
                za[:, o:o + w] = zd[:, c:c + w]
                c += w
            _ZA_CACHE.update(key=zkey, za=za, src=zcat)           # `src` keeps the pointer from being recycled
        logits = ops.gemm(za, Wp, out_dtype=torch.float32, M=nb, N=F, K=Kp)
        ctx.save_for_backward(za, Wp, *[m for m in eff_masks if m is not None])
        ctx.meta = (F, widths, offs, Kp, [m is not None for m in eff_masks], act_dtype)
        ctx.params = params
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        za, Wp = ctx.saved_tensors[:2]
        saved_masks = list(ctx.saved_tensors[2:])
        F, widths, offs, Kp, has_mask, act = ctx.meta
        nb = za.size(0)
        n = len(widths)
        if act == torch.float32 and dlogits.stride(1) == 1 and dlogits.dtype == torch.float32:
            dl = dlogits
        else:
            dl = torch.zeros(nb, pad8(F), dtype=act, device=za.device)
            dl[:, :F] = dlogits
        dzp = ops.gemm(dl, Wp, trans_b=True, out_dtype=torch.float32, M=nb, N=Kp, K=F)
        dz = dzp[:, :widths[0]] if n == 1 else torch.cat([dzp[:, o:o + w] for w, o in zip(widths, offs)], dim=1)
        grads = []
        dlT = None
        for i, (w, o) in enumerate(zip(widths, offs)):
            m = saved_masks.pop(0) if has_mask[i] else None
            if not ctx.needs_input_grad[4 + i]:
                grads.append(None)
                continue
            info = mask_info(m)
            if info[0] == "none":
                m = None
            zb = za[:, o:o + pad8(w)]
            g = grad_view([ctx.params[i]], [True]) if w % 4 == 0 else None
            if info[0] == "sparse":
                # gene-program masks are ~0.5 % dense: only the connected (feature, program) pairs
                if dlT is None:
                    dlT = dlogits.float().t().contiguous()                       # [F, nb]
                zT = zb[:, :w].float().t().contiguous()                           # [w, nb]
                dst = g.view(F, w) if g is not None else torch.zeros(F, w, dtype=torch.float32, device=za.device)
                call("nc_masked_sddmm", dlT, dlT.stride(0), zT, zT.stride(0), nb, info[1], info[2], info[3], F,
                     dst, dst.stride(0))
                grads.append(None if g is not None else dst)
            elif g is not None:
                ops.gemm(dl, zb, trans_a=True, trans_b=True, out=g.view(F, w), mask=m, accumulate=True, M=F, N=w, K=nb)
                grads.append(None)
            else:
                grads.append(ops.gemm(dl, zb, trans_a=True, trans_b=True, out_dtype=torch.float32, mask=m,
                                      M=F, N=w, K=nb))
        return (dz, None, None, None, *grads, *([None] * n))


class NBLossFusedFn(torch.autograd.Function):
    """softmax over all features -> exp(log_lib) * softmax -> reconstructed-column select ->
    mean NB negative log-likelihood, value and gradients in one kernel
    (nn/layers.py:143, nn/decoders.py:168, modules/vgpgae.py:714-749, modules/losses.py:169-212)."""

    @staticmethod
    def forward(ctx, logits, theta_param, x_full, sel):
        nb, F = logits.shape
        dev = logits.device
        acc = torch.zeros(1, dtype=torch.float64, device=dev)
        dlog = torch.empty(nb, F, dtype=torch.float32, device=dev)
        dth = torch.zeros(F, dtype=torch.float32, device=dev)
        means = torch.empty(nb, F, dtype=torch.float32, device=dev)
        dx = torch.empty(nb, F, dtype=torch.float32, device=dev) if ctx.needs_input_grad[2] else None
        logits = logits.contiguous()
        xd = x_full.detach()
        call("nc_nb_loss_fused", logits, logits.stride(0), xd, xd.stride(0), sel, theta_param.detach(),
             nb, F, 1.0 / nb, means, F, acc, dlog, F, dth, dx, F if dx is not None else 0)
        if dx is not None:
            ctx.save_for_backward(dlog, dth, dx)
        else:
            ctx.save_for_backward(dlog, dth)
        ctx.mark_non_differentiable(means)
        return acc.float().reshape(()), means

    @staticmethod
    def backward(ctx, g, _gm):
        dlog, dth = ctx.saved_tensors[:2]
        dx = ctx.saved_tensors[2] * g if len(ctx.saved_tensors) > 2 else None
        return dlog * g, dth * g, dx, None


class NBLossMeansFn(torch.autograd.Function):
    """modules/losses.py:169-212 on precomputed means: mean over rows of the summed NB negative log-likelihood."""

    @staticmethod
    def forward(ctx, x, mu, theta, eps):
        R, F = mu.shape
        xd, md, td = x.detach().float().contiguous(), mu.detach().float().contiguous(), theta.detach().float().contiguous()
        acc = torch.zeros(1, dtype=torch.float64, device=mu.device)
        dmu = torch.empty(R, F, dtype=torch.float32, device=mu.device) if ctx.needs_input_grad[1] else None
        dth = torch.zeros(F, dtype=torch.float32, device=mu.device) if ctx.needs_input_grad[2] else None
        call("nc_nb_loss_means", xd, xd.stride(0), md, md.stride(0), td, R, F, float(eps), 1.0 / R, acc, dmu,
             F if dmu is not None else 0, dth)
        ctx.save_for_backward(dmu, dth)
        return acc.float().reshape(())

    @staticmethod
    def backward(ctx, g):
        dmu, dth = ctx.saved_tensors
        return None, (dmu * g if dmu is not None else None), (dth * g if dth is not None else None), None


# --------------------------------------------------------------------------- weight regularisers
class WeightRegFn(torch.autograd.Function):
    """sum over the given decoder weights of  l1 * sum |W * l1mask|  +  gl * sum_c sqrt(cnt_c) ||W[:,c]||_2
    (modules/losses.py:287-339 and :215-252).  ``l1masks[i]``: uint8 [F, P] or None (all);
    ``gl_masks[i]``: the MaskedLinear mask buffer [F, P] (count of non-masked weights) or None."""

    @staticmethod
    def forward(ctx, l1_scale, gl_scale, l1masks, gl_masks, *weights):
        total = torch.zeros((), dtype=torch.float32, device=weights[0].device)
        coefs = []
        for W, lm, gm in zip(weights, l1masks, gl_masks):
            F, P = W.shape
            dev = W.device
            abs_sum = torch.zeros(P, dtype=torch.float32, device=dev) if l1_scale != 0 else None
            sq = torch.zeros(P, dtype=torch.float32, device=dev) if gl_scale != 0 else None
            cnt = torch.zeros(P, dtype=torch.float32, device=dev) if gl_scale != 0 else None
            Wd = W.detach()
            call("nc_weight_col_stats", Wd, Wd.stride(0), F, P, lm, gm if gl_scale != 0 else None,
                 gm.stride(0) if (gm is not None and gl_scale != 0) else 0, abs_sum, sq, None, cnt)
            coef = None
            if l1_scale != 0:
                total = total + l1_scale * abs_sum.sum()
            if gl_scale != 0:
                norm = sq.sqrt()
                root = cnt.sqrt()
                total = total + gl_scale * (root * norm).sum()
                coef = torch.where(norm > 0, gl_scale * root / norm.clamp_min(1e-30), torch.zeros_like(norm))
            coefs.append(coef)
        ctx.l1_scale = l1_scale
        ctx.l1masks = l1masks
        ctx.coefs = coefs
        ctx.params = weights
        ctx.save_for_backward(*weights)
        return total

    @staticmethod
    def backward(ctx, g):
        grads = []
        gs = g.detach().float().contiguous()
        for i, (W, lm, coef) in enumerate(zip(ctx.saved_tensors, ctx.l1masks, ctx.coefs)):
            if not ctx.needs_input_grad[4 + i]:
                grads.append(None)
                continue
            Wd = W.detach()
            gv = grad_view([ctx.params[i]], [True])
            dW = gv.view_as(W) if gv is not None else torch.zeros_like(W, dtype=torch.float32)
            call("nc_weight_reg_grad", Wd, Wd.stride(0), W.size(0), W.size(1), float(ctx.l1_scale), lm, coef, gs,
                 dW, dW.stride(0))
            grads.append(None if gv is not None else dW)
        return (None, None, None, None, *grads)


# --------------------------------------------------------------------------- attention aggregator
_SEG_CACHE = {}


def _agg_seg(F: int, device) -> torch.Tensor:
    key = (F, str(device))
    if key not in _SEG_CACHE:
        _SEG_CACHE[key] = torch.tensor([0, F, 0, 0], dtype=torch.int32, device=device)
    return _SEG_CACHE[key]


class AttentionAggregatorFn(torch.autograd.Function):
    """nn/aggregators.py:113-156: g_l = linear_l_l(x), g_r = linear_r_l(x) (tcgen05 /
    SIMT GEMMs), alpha = segment softmax of attn . leaky_relu(g_r[i] + g_l[j]) over the incoming edges
    exactly as sampled (GATv2 score kernels without self-loop re-normalisation), out[i] = mean_h sum_j
    alpha_ijh x[j] = CSR SpMM of the raw rows with the head-averaged coefficients.  Returns
    (out [n_dst, F] fp32, alpha [E_dst, H] in CSR entry order)."""

    @staticmethod
    def forward(ctx, x, graph: SubgraphCSR, heads, slope, act_dtype, cache_key, want_agg, w_l, w_r, attn):
        dev = x.device
        n_sub, n_dst = graph.n_nodes, graph.n_dst
        F = x.size(1)
        assert w_l.shape == (heads * F, F) and x.size(0) == n_sub
        x = x.contiguous().float()
        ldx = pad8(F)
        xa = torch.zeros(n_sub, ldx, dtype=act_dtype, device=dev)
        xa[:, :F] = x
        Wl = pack_rows((cache_key, "Wl"), [w_l], act_dtype, F)
        Wr = pack_rows((cache_key, "Wr"), [w_r], act_dtype, F)
        NL = heads * F
        ldg = pad8(NL)
        Gl = torch.empty(n_sub, ldg, dtype=act_dtype, device=dev)
        Gr = torch.empty(max(n_dst, 1), ldg, dtype=act_dtype, device=dev)
        ops.gemm(xa, Wl, out=Gl[:, :NL], M=n_sub, N=NL, K=F)
        ops.gemm(xa, Wr, out=Gr[:, :NL], M=n_dst, N=NL, K=F)
        att = attn.detach().reshape(-1).float().contiguous()
        n_slots = graph.n_slots
        alpha = torch.empty(max(n_slots, 1) * heads, dtype=torch.float32, device=dev)
        w = torch.empty(max(graph.col.numel(), 1), dtype=torch.float32, device=dev)
        seg = _agg_seg(F, dev)
        call("nc_attn_agg_fwd", Gl, Gr, dtype_code(act_dtype), ldg, att, graph.rowptr, graph.col, n_dst, seg, heads, F,
             graph.max_deg, float(slope), alpha, w)
        out = ops.spmm_csr(x, graph.rowptr, graph.col, w, None, False, n_dst)
        ctx.save_for_backward(x, xa, Gl, Gr, alpha, att, Wl, Wr)
        ctx.meta = (graph, heads, float(slope), F, NL, ldg, act_dtype)
        ctx.params = (w_l, w_r, attn)
        # per-entry, per-head coefficients in CSR entry order (slot of entry e of row i is e + i)
        E = graph.col.numel() if want_agg else 0
        if E > 0:
            rows = torch.searchsorted(graph.rowptr[1:n_dst + 1].contiguous(),
                                      torch.arange(E, device=dev, dtype=torch.int32), right=True)
            slots = torch.arange(E, device=dev) + rows
            agg = alpha.view(-1, heads)[slots.clamp_(max=max(n_slots - 1, 0))]
        else:
            agg = alpha.new_zeros(0, heads)
        ctx.mark_non_differentiable(agg)
        return out, agg

    @staticmethod
    def backward(ctx, dout, _dagg):
        x, xa, Gl, Gr, alpha, att, Wl, Wr = ctx.saved_tensors
        graph, heads, slope, F, NL, ldg, act = ctx.meta
        w_l, w_r, attn = ctx.params
        dev = x.device
        n_sub, n_dst = graph.n_nodes, graph.n_dst
        dout = dout.contiguous().float()
        rev = ops.reverse_csr(graph.rowptr, graph.col, n_sub, n_dst, 2, graph.n_slots)
        dGl = torch.empty(n_sub, ldg, dtype=act, device=dev)
        dGr = torch.empty(max(n_dst, 1), ldg, dtype=act, device=dev)
        if ldg != NL:
            dGl[:, NL:].zero_()
            dGr[:, NL:].zero_()
        need = ctx.needs_input_grad
        g_att = grad_view([attn], [need[9]])
        datt = g_att if g_att is not None else torch.zeros(NL, dtype=torch.float32, device=dev)
        dal = torch.empty_like(alpha)
        ds = torch.empty_like(alpha)
        seg = _agg_seg(F, dev)
        call("nc_attn_agg_bwd", Gl, Gr, dtype_code(act), ldg, att, graph.rowptr, graph.col, n_dst, n_sub,
             rev[0], rev[1], rev[2], seg, heads, F, graph.max_deg, slope, dout, dout.stride(0), x, x.stride(0),
             alpha, dal, ds, dGl, dGr, datt)
        g_wl = grad_view([w_l], [need[7]]) if F % 4 == 0 else None
        g_wr = grad_view([w_r], [need[8]]) if F % 4 == 0 else None
        if g_wl is not None:
            ops.gemm(dGl, xa, trans_a=True, trans_b=True, out=g_wl.view(NL, F), accumulate=True, M=NL, N=F, K=n_sub)
            dWl = None
        else:
            dWl = ops.gemm(dGl, xa, trans_a=True, trans_b=True, out_dtype=torch.float32, M=NL, N=F, K=n_sub)
        if g_wr is not None:
            if n_dst > 0:
                ops.gemm(dGr, xa, trans_a=True, trans_b=True, out=g_wr.view(NL, F), accumulate=True, M=NL, N=F, K=n_dst)
            dWr = None
        else:
            dWr = ops.gemm(dGr, xa, trans_a=True, trans_b=True, out_dtype=torch.float32, M=NL, N=F, K=n_dst) \
                if n_dst > 0 else torch.zeros(NL, F, device=dev)
        d_attn = None if g_att is not None else datt.reshape(attn.shape)
        return (None, None, None, None, None, None, None,
                dWl.contiguous() if dWl is not None and dWl.stride(0) != F else dWl,
                dWr.contiguous() if dWr is not None and dWr.stride(0) != F else dWr, d_attn)


# --------------------------------------------------------------------------- GCN encoder
class EncoderGCNFn(torch.autograd.Function):
    """fc_l1 + ReLU + the GCNConv heads of nn/encoders.py:105-117: one
    packed GEMM for the four ``lin`` weights, then the symmetric-normalised CSR SpMM with
    subgraph-local degrees and exactly one self loop per node.  ``tensors``: (lin.weight, bias)
    per conv in output-column order."""

    @staticmethod
    def forward(ctx, x_enc, graph: SubgraphCSR, n_feat: int, out_offsets, out_width: int, cache_key, act, W1, b1,
                *tensors):
        """``W1 is None``: second conv layer of the 2-layer encoder (see ``EncoderGATv2Fn.forward``)."""
        NC = len(tensors) // 2
        Ws, bs = [tensors[2 * c] for c in range(NC)], [tensors[2 * c + 1] for c in range(NC)]
        ident = isinstance(W1, str)
        pre = W1 is None or ident
        ctx.ident = ident
        dev = x_enc.device
        n_sub, n_dst = graph.n_nodes, graph.n_dst
        Hd = x_enc.size(1) if pre else W1.size(0)
        Cs = [w.size(0) for w in Ws]
        assert list(out_offsets) == [sum(Cs[:c]) for c in range(NC)] and sum(Cs) == out_width
        W1p = None if pre else pack_rows((cache_key, "W1"), [W1], act, W1.size(1))
        Wall = pack_rows((cache_key, "Wgcn"), Ws, act, Hd)
        bias_out = _cat_f32(bs)
        ldh = pad8(Hd)
        Hbuf = torch.zeros(n_sub, ldh, dtype=act, device=dev)
        if pre:
            xin = x_enc.detach().float().contiguous()
            if ident:
                Hbuf[:, :Hd].copy_(xin)
            else:
                call("nc_relu_cast", xin, xin.stride(0), n_sub, Hd, Hbuf, dtype_code(act), ldh)
        else:
            ops.gemm(x_enc, W1p, bias=b1.detach().float(), relu=True, out=Hbuf[:, :Hd], M=n_sub, N=Hd, K=n_feat)
        ldw = pad8(out_width)
        XW = torch.zeros(n_sub, ldw, dtype=act, device=dev)
        ops.gemm(Hbuf, Wall, out=XW[:, :out_width], M=n_sub, N=out_width, K=Hd)
        w, wself = graph.gcn_weights(use_out_degree=False)
        out = ops.spmm_csr(XW, graph.rowptr, graph.col, w, wself, True, n_dst, bias=bias_out, n_cols=out_width)
        ctx.save_for_backward(None if pre else x_enc, Hbuf, W1p, Wall, w, wself)
        ctx.meta = (graph, n_feat, out_width, NC, Cs, Hd, [tuple(t.shape) for t in tensors], act)
        return out

    @staticmethod
    def backward(ctx, dout):
        x_enc, Hbuf, W1p, Wall, w, wself = ctx.saved_tensors
        graph, n_feat, out_width, NC, Cs, Hd, shapes, act = ctx.meta
        pre = W1p is None
        dev = Hbuf.device
        n_sub, n_dst = graph.n_nodes, graph.n_dst
        dout = dout.contiguous().float()
        rptr, rdst, rslot = graph.reverse()
        rd = rdst.long()
        is_self = rslot == (graph.rowptr[rd + 1] + rdst)
        e_idx = (rslot - rdst).long().clamp_(0, max(w.numel() - 1, 0))
        rw = torch.where(is_self, wself[rd.clamp_(0, max(n_dst - 1, 0))], w[e_idx]).contiguous()
        ldw = pad8(out_width)
        dXW = torch.zeros(n_sub, ldw, dtype=act, device=dev)
        ops.spmm_csr(dout, rptr, rdst, rw, None, False, n_sub, out=dXW[:, :out_width], n_cols=out_width)
        dbias = _col_sum(dout, out_width)
        dWall = ops.gemm(dXW, Hbuf, trans_a=True, trans_b=True, out_dtype=torch.float32, M=out_width, N=Hd, K=n_sub)
        dH = torch.empty(n_sub, pad8(Hd), dtype=torch.float32, device=dev)
        ops.gemm(dXW, Wall, trans_b=True, out=dH[:, :Hd], M=n_sub, N=Hd, K=out_width)
        dHp = None
        if not ctx.ident:
            dHp = torch.empty(n_sub, Hbuf.size(1), dtype=act, device=dev)
            call("nc_relu_bwd", dH, dH.stride(0), Hbuf, dtype_code(act), Hbuf.stride(0), n_sub, Hd, dHp, dHp.stride(0))
        dW1 = db1 = None
        if not pre:
            dW1 = ops.gemm(dHp, x_enc, trans_a=True, trans_b=True, out_dtype=torch.float32, M=Hd, N=n_feat, K=n_sub)
            db1 = _col_sum(dHp, Hd)
        grads, r = [], 0
        for c in range(NC):
            grads += [dWall[r:r + Cs[c]], dbias[r:r + Cs[c]]]
            r += Cs[c]
        dx = None
        if pre:
            if ctx.needs_input_grad[0]:
                dx = dH[:, :Hd].contiguous() if ctx.ident else dHp[:, :Hd].float()
        elif ctx.needs_input_grad[0]:
            dxf = ops.gemm(dHp, W1p, trans_b=True, out_dtype=torch.float32, M=n_sub, N=n_feat, K=Hd)
            dx = torch.zeros_like(x_enc, dtype=torch.float32)
            dx[:, :n_feat] = dxf
            dx = dx.to(x_enc.dtype)
        return (dx, None, None, None, None, None, None, dW1, db1, *grads)
