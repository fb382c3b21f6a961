"""
Full-graph latent inference, layer-wise and partitioned over the GPUs of one box.

Reference: ``NicheCompass.get_latent_representation`` (models/nichecompass.py:1264-1381) walks a
``NeighborLoader`` over blocks of ``node_batch_size`` nodes and calls
``VGPGAE.get_latent_representation`` (modules/vgpgae.py:1368-1437) per block, so every node's
``fc_l1`` / ``lin_l`` rows are recomputed once per block that contains it as a neighbour
((1 + deg) x redundant encoder work).

With the GATv2 encoder (1 layer) and all neighbours sampled, a node's ``mu`` depends only on its own
``hidden`` row and those of its neighbours, so the same values are obtained layer by layer:

  1. partition   nodes are ordered along a space-filling (Morton) curve of ``obsm['spatial']`` (or by
                 reverse Cuthill-McKee of the graph when there are no coordinates) and cut into W
                 contiguous tiles, one per rank: tiles are spatially compact, the halo (neighbours owned
                 by another rank) is the tile boundary only;
  2. hidden      every rank computes ``relu(fc_l1(log1p x))`` ONCE for its own nodes (CSR gather ->
                 tcgen05 GEMM with bias + ReLU epilogue);
  3. halo        the ``hidden`` rows of boundary nodes are exchanged: one ``all_to_all_single`` of the
                 requested ids, one of the packed rows (``Hd`` values per halo node instead of the
                 ``H * P`` of ``x_l``: at C2 1600 vs 6400);
  4. x_l         ``lin_l`` of the stacked convs for own + halo rows, once per node;
  5. attention   per tile of targets: ``lin_r`` GEMM + ``nc_gatv2_fwd_tile`` (gather -> LeakyReLU ->
                 segment softmax -> weighted sum -> head mean) straight over the rank-local CSR;
  6. gather      ``all_gather`` of the per-rank results, scattered back to the caller's row order.

The index plumbing (permutation, local CSR, halo lists) is torch tensor code and also runs on CPU
tensors under ``gloo`` (tests/test_host_logic.py); all arithmetic is in the CUDA library.
"""
import os
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from .. import ops
from .._lib import HostPtr, call, dtype_code
from ..functional import _cat_f32, pack_rows
from ..ops import pad8


# ----------------------------------------------------------------------------- partition
def morton_order(coords: np.ndarray, bits: int = 16) -> np.ndarray:
    """new position -> old id: nodes sorted by the Z-order code of their quantised 2-D coordinates."""
    c = np.asarray(coords, dtype=np.float64)[:, :2]
    lo, hi = c.min(0), c.max(0)
    q = np.minimum(((c - lo) / np.maximum(hi - lo, 1e-300) * (2 ** bits - 1)).astype(np.uint64), 2 ** bits - 1)

    def spread(v):                                     # interleave zeros between the low 16 bits
        v = v & np.uint64(0xFFFF)
        v = (v | (v << np.uint64(8))) & np.uint64(0x00FF00FF)
        v = (v | (v << np.uint64(4))) & np.uint64(0x0F0F0F0F)
        v = (v | (v << np.uint64(2))) & np.uint64(0x33333333)
        v = (v | (v << np.uint64(1))) & np.uint64(0x55555555)
        return v
    code = spread(q[:, 0]) | (spread(q[:, 1]) << np.uint64(1))
    return np.argsort(code, kind="stable").astype(np.int64)


def partition_order(n_obs: int, coords: Optional[np.ndarray], adj=None, method: str = "auto") -> np.ndarray:
    """Permutation (new position -> old id) under which contiguous ranges are compact tiles."""
    if method == "auto":
        method = "spatial" if coords is not None else ("rcm" if adj is not None else "none")
    if method == "spatial":
        if coords is None:
            raise ValueError("partition='spatial' needs adata.obsm['spatial']")
        return morton_order(coords)
    if method == "rcm":
        from scipy.sparse.csgraph import reverse_cuthill_mckee
        import scipy.sparse as sp
        return np.asarray(reverse_cuthill_mckee(sp.csr_matrix(adj), symmetric_mode=True), dtype=np.int64)
    if method == "none":
        return np.arange(n_obs, dtype=np.int64)
    raise ValueError(f"unknown partition method {method}")


def tile_bounds(n: int, world: int) -> List[int]:
    """Contiguous, near-equal ranges of the permuted node order: rank r owns [b[r], b[r+1])."""
    return [(n * r) // world for r in range(world + 1)]


# ----------------------------------------------------------------------------- rank-local graph
class LocalGraph:
    """CSR (incoming neighbours) of the targets a rank owns, in rank-local ids: own nodes are
    [0, n_own) in permuted order, halo nodes follow as n_own + k (k = rank in ``halo_pos``)."""

    def __init__(self, rowptr, col, halo_pos, n_own, max_deg, lo, hi):
        self.rowptr, self.col, self.halo_pos = rowptr, col, halo_pos
        self.n_own, self.max_deg, self.lo, self.hi = int(n_own), int(max_deg), int(lo), int(hi)

    @property
    def n_halo(self) -> int:
        return int(self.halo_pos.numel())


def build_local_graph(rowptr: torch.Tensor, col: torch.Tensor, perm: torch.Tensor, inv: torch.Tensor,
                      lo: int, hi: int) -> LocalGraph:
    """``rowptr`` / ``col``: CSC of the full graph by original ids (any int dtype, one device);
    ``perm``: new position -> old id, ``inv``: old id -> new position (int64, same device).
    The rank owns new positions [lo, hi)."""
    dev = rowptr.device
    n_own = hi - lo
    own_old = perm[lo:hi]
    rp = rowptr.long()
    start = rp[own_old]
    deg = rp[own_old + 1] - start
    rowptr_loc = torch.zeros(n_own + 1, dtype=torch.int64, device=dev)
    torch.cumsum(deg, 0, out=rowptr_loc[1:])
    total = int(rowptr_loc[-1]) if n_own > 0 else 0
    if total >= 2 ** 31:
        raise ValueError("rank-local graph has more than 2^31 edges; use more ranks")
    within = torch.arange(total, device=dev) - torch.repeat_interleave(rowptr_loc[:-1], deg)
    src_new = inv[col.long()[torch.repeat_interleave(start, deg) + within]]
    own = (src_new >= lo) & (src_new < hi)
    halo_pos = torch.unique(src_new[~own])                       # sorted new positions owned elsewhere
    col_loc = torch.where(own, src_new - lo, n_own + torch.searchsorted(halo_pos, src_new))
    max_deg = int(deg.max()) if n_own > 0 else 0
    return LocalGraph(rowptr_loc.to(torch.int32), col_loc.to(torch.int32).contiguous(), halo_pos, n_own,
                      max_deg, lo, hi)


# ----------------------------------------------------------------------------- halo exchange
def exchange_halo_rows(own_rows: torch.Tensor, lg: LocalGraph, bounds: Sequence[int], group=None,
                       out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Rows of the halo nodes of ``lg`` (owned by other ranks), in ``lg.halo_pos`` order.
    ``own_rows`` [n_own, ld]: this rank's rows in permuted order.  Two collectives: the requested
    positions (int64) and the packed rows.  With one rank (or no process group) the halo is empty."""
    import torch.distributed as dist
    n_halo = lg.n_halo
    ld = own_rows.size(1)
    if out is None:
        out = torch.empty(n_halo, ld, dtype=own_rows.dtype, device=own_rows.device)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        assert n_halo == 0, "halo nodes without a process group"
        return out
    world = dist.get_world_size(group)
    dev = own_rows.device
    b = torch.tensor(list(bounds), dtype=torch.int64, device=dev)
    owner = torch.searchsorted(b, lg.halo_pos, right=True) - 1             # ascending, like halo_pos
    want = torch.bincount(owner, minlength=world)                          # rows I request from rank r
    give = torch.empty_like(want)
    dist.all_to_all_single(give, want, group=group)                        # rows rank r requests from me
    want_l, give_l = want.tolist(), give.tolist()
    req = torch.empty(int(sum(give_l)), dtype=torch.int64, device=dev)
    dist.all_to_all_single(req, lg.halo_pos.contiguous(), output_split_sizes=give_l, input_split_sizes=want_l,
                           group=group)
    send = own_rows.index_select(0, req - lg.lo).contiguous()
    dist.all_to_all_single(out, send, output_split_sizes=want_l, input_split_sizes=give_l, group=group)
    return out


def gather_halo_rows_peer(peer, n_own_rows: int, ld: int, dtype: torch.dtype, lg: LocalGraph, bounds: Sequence[int],
                          out: torch.Tensor, group=None) -> torch.Tensor:
    """The same rows without packing or a collective: every rank's ``hidden`` buffer lives in peer-mapped memory
    (``train/optim.py: PeerBuffer``) and the halo rows are read where they are, over NVLink, by one kernel
    (``nc_halo_gather_rows``).  ``peer``: this rank's PeerBuffer holding [>= n_own_rows, ld] rows of ``dtype``."""
    import torch.distributed as dist
    if lg.n_halo == 0:
        dist.barrier(group=group)
        return out
    dev = out.device
    b = torch.tensor(list(bounds), dtype=torch.int64, device=dev)
    owner = torch.searchsorted(b, lg.halo_pos, right=True) - 1
    pos = (lg.halo_pos - b[owner]).contiguous()                            # row index inside the owner's buffer
    esz = torch.empty(0, dtype=dtype).element_size()
    torch.cuda.synchronize()
    dist.barrier(group=group)                                              # every owner has written its rows
    call("nc_halo_gather_rows", peer.ptr_array, len(peer.ptrs), owner.to(torch.int32).contiguous(), pos, lg.n_halo,
         ld * esz, ld * esz, out, out.stride(0) * esz)
    torch.cuda.synchronize()
    dist.barrier(group=group)                                              # nobody frees / rewrites while peers read
    return out


# ----------------------------------------------------------------------------- layer-wise encoder
def supports(module) -> bool:
    """The layer-wise path reproduces the loader path exactly only where a node's output is
    independent of the batch it is sampled in: GATv2 convs, one conv layer, one fc layer."""
    enc = module.encoder
    return enc.conv_layer == "gatv2conv" and enc.n_layers == 1 and enc.n_fc_layers == 1


@torch.no_grad()
def encode_partition(module, store, lg: LocalGraph, own_old: torch.Tensor, bounds: Sequence[int],
                     cats_dev: Optional[torch.Tensor], want_logstd: bool, group=None,
                     chunk_rows: int = 32768, tile_rows: int = 32768) -> torch.Tensor:
    """(mu | logstd) rows [n_own, P or 2P] fp32 of the nodes this rank owns (permuted order).
    ``own_old``: their original ids (int64, device)."""
    enc = module.encoder
    act = enc.compute_dtype
    dev = own_old.device
    n_own, n_halo = lg.n_own, lg.n_halo
    n_loc = n_own + n_halo
    Hd = enc.fc_l1.weight.size(0)
    ldh = pad8(Hd)
    ldx = pad8(enc.n_input)
    P = enc.n_gps
    inject = (module.n_cat_covariates_ > 0 and "encoder" in module.cat_covariates_embeds_injection_
              and cats_dev is not None)

    # ---- 2. hidden = relu(fc_l1(log1p x)) for the owned nodes
    W1 = pack_rows((enc._key, "W1"), [enc.fc_l1.weight], act, enc.fc_l1.weight.size(1))
    b1 = enc.fc_l1.bias.detach().float()
    import torch.distributed as dist
    multi = dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1
    peer = None
    if multi and os.environ.get("NC_HALO_EXCHANGE", "peer") == "peer":
        # hidden rows in peer-mapped memory: the halo is gathered straight from the owners' HBM (no pack, no collective)
        from ..train.optim import PeerBuffer
        try:
            peer = PeerBuffer(max(n_loc, 1) * ldh * torch.empty(0, dtype=act).element_size(), group)
        except Exception:                                                  # noqa: BLE001 -- e.g. no cudaIpc between the devices
            peer = None
        ok = torch.tensor([1 if peer is not None else 0], device=dev)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=group)
        if int(ok.item()) == 0 and peer is not None:
            peer.close()
            peer = None
    if peer is not None:
        Hall = peer.tensor(act)[:n_loc * ldh].view(n_loc, ldh)
    else:
        Hall = torch.zeros(n_loc, ldh, dtype=act, device=dev) if ldh != Hd else \
            torch.empty(n_loc, ldh, dtype=act, device=dev)
    ids32 = own_old.to(torch.int32)
    for c0 in range(0, n_own, chunk_rows):
        c1 = min(n_own, c0 + chunk_rows)
        ids = ids32[c0:c1].contiguous()
        if inject:
            xd = ops.csr_gather_rows(store, ids, c1 - c0, module.log_variational_, torch.float32)
            x_enc = enc.pack_input(xd, module._embed(cats_dev[own_old[c0:c1]]))
        else:
            x_enc = ops.csr_gather_rows(store, ids, c1 - c0, module.log_variational_, act, ldx)
        ops.gemm(x_enc, W1, bias=b1, relu=True, out=Hall[c0:c1, :Hd], M=c1 - c0, N=Hd, K=enc.n_input)

    # ---- 3. halo rows of hidden from their owners
    if peer is not None:
        gather_halo_rows_peer(peer, n_own, ldh, act, lg, bounds, Hall[n_own:], group=group)
    else:
        exchange_halo_rows(Hall[:n_own], lg, bounds, group=group, out=Hall[n_own:])

    # ---- 4. x_l = lin_l(hidden) of the stacked convs, own + halo rows
    convs = [(c, o) for c, o in enc._convs() if want_logstd or o < P]
    heads = enc.n_attention_heads
    tens = [c.param_tuple() for c, _ in convs]
    Cs = [t[5].numel() for t in tens]
    widths = [heads * C for C in Cs]
    offs = [sum(widths[:k]) for k in range(len(convs))]
    NL = sum(widths)
    ldl = pad8(NL)
    out_width = 2 * P if want_logstd else P
    tag = "all" if want_logstd else "mu"
    Wl = pack_rows((enc._key, "Wl", tag), [t[0] for t in tens], act, Hd)
    Wr = pack_rows((enc._key, "Wr", tag), [t[2] for t in tens], act, Hd)
    bl, br = _cat_f32([t[1] for t in tens]), _cat_f32([t[3] for t in tens])
    att = _cat_f32([t[4] for t in tens], pad_to=ldl)
    bias_out = torch.zeros(out_width, dtype=torch.float32, device=dev)
    for (cv, o), C, t in zip(convs, Cs, tens):
        bias_out[o:o + C] = t[5].detach()
    seg_host = torch.tensor([[offs[k], Cs[k], convs[k][1], 0] for k in range(len(convs))], dtype=torch.int32)
    seg = seg_host.to(dev)
    XL = torch.empty(n_loc, ldl, dtype=act, device=dev)
    if ldl != NL:
        XL[:, NL:].zero_()
    for c0 in range(0, n_loc, chunk_rows):
        c1 = min(n_loc, c0 + chunk_rows)
        ops.gemm(Hall[c0:c1], Wl, bias=bl, out=XL[c0:c1, :NL], M=c1 - c0, N=NL, K=Hd)

    # ---- 5. per tile of targets: x_r = lin_r(hidden), fused attention
    out = torch.empty(n_own, out_width, dtype=torch.float32, device=dev)
    XR = torch.empty(min(tile_rows, max(n_own, 1)), ldl, dtype=act, device=dev)
    if ldl != NL:
        XR[:, NL:].zero_()
    # per-head outputs of the ring attention kernels (csrc/gatv2_ring.cu), scratch of one tile
    head_ws = torch.empty(XR.size(0), ldl, dtype=torch.float32, device=dev) if act == torch.bfloat16 else None
    for t0 in range(0, n_own, tile_rows):
        t1 = min(n_own, t0 + tile_rows)
        ops.gemm(Hall[t0:t1], Wr, bias=br, out=XR[:t1 - t0, :NL], M=t1 - t0, N=NL, K=Hd)
        o = out[t0:t1]
        call("nc_gatv2_fwd_tile", XL, XR, dtype_code(act), ldl, att, bias_out, lg.rowptr, lg.col, t0, t1 - t0,
             seg, HostPtr(seg_host), len(convs), heads, lg.max_deg, 0.2, o, o.stride(0), head_ws)
    if peer is not None:
        del Hall
        dist.barrier(group=group)
        peer.close()
    return out
