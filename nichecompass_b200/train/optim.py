"""
Flat-buffer Adam on the ``nc_adam_step`` kernel (reference: train/trainer.py:339-342,414-421 --
``torch.optim.Adam(params, lr, weight_decay)``, ``zero_grad``, optional ``clip_grad_value_``,
``step``).

All trainable parameters are re-homed as views into ONE contiguous fp32 buffer ("arena"), their
gradients as views into a second one, so that
  * ``zero_grad`` costs nothing (the Adam kernel clears the gradient it has just consumed),
  * the data-parallel gradient exchange is one NCCL all-reduce on the flat gradient buffer,
  * the optimiser step (clip + Adam + bias correction + bf16 shadow refresh) is one launch,
  * parameters that a kernel consumes stacked (the four ``lin_l`` weights of the encoder, ...)
    are laid out back to back (``groups``), so the stacked matrix, its bf16 shadow and its
    gradient are plain VIEWS of the arenas: the GEMMs read the shadow and accumulate the weight
    gradients in place, with no per-step cast / cat / add kernels
    (``nichecompass_b200.functional.arena_view``).
Update rule, bias correction and epsilon placement are torch.optim.Adam's.  One documented
deviation: a parameter that never receives a gradient still takes part in the update with a
zero gradient (torch skips it), which only matters together with ``weight_decay != 0``.
"""
from typing import Iterable, List, Optional, Sequence

import torch

from .. import functional as Fn
from .._lib import call

ALIGN = 8          # elements: group starts are 32-byte (fp32) / 16-byte (bf16 shadow) aligned


class _RawCuda:
    """``__cuda_array_interface__`` over a raw device pointer (memory owned by ``PeerBuffer``)."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.owner = owner
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class PeerBuffer:
    """Device memory the other ranks of the box can map: ``nc_peer_alloc`` (cudaMalloc, zero-filled) + a cudaIpc
    handle exchanged through ``torch.distributed``; ``ptrs[r]`` is rank r's buffer as seen from this process
    (own pointer for the own rank).  Backs the arenas of ``FlatAdam(peer_group=...)``."""

    def __init__(self, nbytes: int, group=None):
        import ctypes
        import torch.distributed as dist
        from .._lib import NicheCompassCudaError, lib
        self._lib, self._ct = lib(), ctypes
        self.nbytes = int(nbytes)
        p = ctypes.c_void_p()
        if self._lib.nc_peer_alloc(ctypes.c_int64(self.nbytes), ctypes.byref(p)) != 0:
            raise NicheCompassCudaError(self._lib.nc_last_error().decode())
        self.ptr = p.value
        h = (ctypes.c_ubyte * 64)()
        if self._lib.nc_peer_get_handle(ctypes.c_void_p(self.ptr), h) != 0:
            raise NicheCompassCudaError(self._lib.nc_last_error().decode())
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        handles = [None] * world
        dist.all_gather_object(handles, bytes(h), group=group)
        self.ptrs, self._opened = [], []
        for r, hb in enumerate(handles):
            if r == rank:
                self.ptrs.append(self.ptr)
                continue
            q = ctypes.c_void_p()
            buf = (ctypes.c_ubyte * 64).from_buffer_copy(hb)
            if self._lib.nc_peer_open(buf, ctypes.byref(q)) != 0:
                raise NicheCompassCudaError(f"rank {rank}: cannot map rank {r}'s buffer: " +
                                            self._lib.nc_last_error().decode())
            self.ptrs.append(q.value)
            self._opened.append(q.value)
        self.ptr_array = (ctypes.c_void_p * world)(*self.ptrs)
        dist.barrier(group=group)

    def tensor(self, dtype: torch.dtype) -> torch.Tensor:
        t = torch.as_tensor(_RawCuda(self.ptr, self.nbytes, self), device=torch.device("cuda", torch.cuda.current_device()))
        return t.view(dtype)

    def close(self):
        """Unmap the peers' buffers and free the own one (all ranks must have stopped using them)."""
        torch.cuda.synchronize()
        for q in self._opened:
            self._lib.nc_peer_close(self._ct.c_void_p(q))
        self._opened = []
        if self.ptr:
            self._lib.nc_peer_free(self._ct.c_void_p(self.ptr))
            self.ptr = None


class FlatAdam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 0.,
                 groups: Optional[Sequence[Sequence[torch.nn.Parameter]]] = None,
                 bf16_shadow: bool = False, peer_group=None, peer_ctas: int = 0):
        """``peer_group``: a ``torch.distributed`` process group of ranks on ONE NVLink box (``True``: the world).
        The arenas then live in peer-mapped memory and ``step()`` is the fused exchange + Adam kernel
        (``nc_dp_adam_step``, csrc/dp_exchange.cu): no NCCL all-reduce, the moments exist for the owned shard only."""
        wanted = [p for p in params if p.requires_grad]
        if not wanted:
            raise ValueError("optimizer got an empty parameter list")
        dev = wanted[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatAdam needs CUDA parameters (nichecompass_b200 has no CPU path)")
        ids = {id(p) for p in wanted}
        # arena order: the hinted groups (tight inside, every member trainable), then the rest
        layout: List[List[torch.nn.Parameter]] = []
        placed = set()
        for grp in (groups or []):
            grp = list(grp)
            if grp and all(id(p) in ids and id(p) not in placed for p in grp):
                layout.append(grp)
                placed.update(id(p) for p in grp)
        for p in wanted:
            if id(p) not in placed:
                layout.append([p])
                placed.add(id(p))
        self.params: List[torch.nn.Parameter] = []
        self.offsets: List[int] = []
        n = 0
        for grp in layout:
            n = (n + ALIGN - 1) // ALIGN * ALIGN
            for p in grp:
                self.params.append(p)
                self.offsets.append(n)
                n += p.numel()
        self.numel = n = (n + ALIGN - 1) // ALIGN * ALIGN
        self.peer = None
        if peer_group is not None and peer_group is not False:
            import torch.distributed as dist
            grp = None if peer_group is True else peer_group
            world, rank = dist.get_world_size(grp), dist.get_rank(grp)
            if world > 1:
                bufs = {"p": PeerBuffer(4 * n, grp), "g": PeerBuffer(4 * n, grp),
                        "flags": PeerBuffer(max(8 * world, 256), grp)}
                if bf16_shadow:
                    bufs["sh"] = PeerBuffer(2 * n, grp)
                shard = ((n + world - 1) // world + 3) // 4 * 4
                self.peer = {"bufs": bufs, "world": world, "rank": rank, "group": grp, "shard": shard,
                             "ctas": int(peer_ctas),
                             "done": torch.zeros(2, dtype=torch.int32, device=dev), "aux": None}
        if self.peer is None:
            self.flat_p = torch.zeros(n, dtype=torch.float32, device=dev)
            self.flat_g = torch.zeros(n, dtype=torch.float32, device=dev)
            self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
            self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
            self.flat_bf16 = torch.zeros(n, dtype=torch.bfloat16, device=dev) if bf16_shadow else None
        else:
            b = self.peer["bufs"]
            self.flat_p, self.flat_g = b["p"].tensor(torch.float32), b["g"].tensor(torch.float32)
            self.flat_bf16 = b["sh"].tensor(torch.bfloat16) if bf16_shadow else None
            self.exp_avg = torch.zeros(self.peer["shard"], dtype=torch.float32, device=dev)
            self.exp_avg_sq = torch.zeros(self.peer["shard"], dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.flat_p[o:o + p.numel()].view_as(p)
                view.copy_(p.data)
                p.data = view
                p.grad = self.flat_g[o:o + p.numel()].view_as(p)
                p._nc_arena = (self, o)
        self._shadow_versions = None
        self.refresh_shadow()
        self._grads_clean = True
        # the one param group keeps the reference's ``param_group["lr"] *= factor`` idiom working
        self.param_groups = [{"lr": lr, "betas": betas, "eps": eps, "weight_decay": weight_decay,
                              "params": self.params}]
        self.step_count = 0
        # {learning rate, number of the NEXT update} on the device: the update kernels read them there, so a captured
        # training step can be replayed without any host-side argument going stale (train/trainer.py graph buckets)
        self.dyn = torch.tensor([float(lr), 1.0], dtype=torch.float32, device=dev)
        self._lr_on_device = float(lr)

    def attach_aux(self, buf: torch.Tensor) -> torch.Tensor:
        """Peer mode: a small fp32 statistic (the module's ``running_mean_abs_mu``) that every step replaces by its
        mean over the ranks, inside the exchange kernel.  Returns the peer-mapped tensor that must take the place
        of ``buf`` (same values)."""
        if self.peer is None:
            return buf
        n = buf.numel()
        pb = PeerBuffer(max(4 * n, 256), self.peer["group"])
        t = pb.tensor(torch.float32)[:n]
        t.copy_(buf.detach().float().reshape(-1))
        self.peer["aux"] = (pb, torch.zeros(n, dtype=torch.float32, device=buf.device), n)
        return t.view_as(buf)

    def close(self):
        """Peer mode: unmap / free the peer buffers (collective: every rank calls it once its last step is queued)."""
        if getattr(self, "peer", None) is None:
            return
        import torch.distributed as dist
        torch.cuda.synchronize()
        dist.barrier(group=self.peer["group"])
        bufs = list(self.peer["bufs"].values()) + ([self.peer["aux"][0]] if self.peer["aux"] else [])
        self.flat_p = self.flat_g = self.flat_bf16 = None
        for b in bufs:
            b.close()
        dist.barrier(group=self.peer["group"])
        self.peer = None

    def sync_lr(self):
        """Push ``param_groups[0]["lr"]`` (the reference's ``param_group["lr"] *= factor`` idiom) to the device."""
        lr = float(self.param_groups[0]["lr"])
        if lr != self._lr_on_device:
            self.dyn[0:1].fill_(lr)
            self._lr_on_device = lr

    def note_replayed_step(self):
        """Book-keeping for one update that ran from a replayed CUDA graph (no Python ``step()`` call)."""
        self.step_count += 1
        self._grads_clean = True
        Fn.bump_param_epoch()

    # ------------------------------------------------------------------ bf16 shadow
    def _versions(self):
        return tuple(p._version for p in self.params)

    def refresh_shadow(self):
        """Re-cast the whole arena (after load_state_dict / best-state reload / manual edits)."""
        if self.flat_bf16 is not None:
            call("nc_cast_bf16", self.flat_p, self.numel, None, 0, 1, self.numel, self.flat_bf16, self.numel, 0)
        self._shadow_versions = self._versions()

    def shadow_is_current(self) -> bool:
        """The Adam kernel keeps the shadow in step with the arena; in-place edits made through
        torch (``copy_``, ``mul_``...) bump the parameters' version counters instead."""
        return self._shadow_versions == self._versions()

    # ------------------------------------------------------------------ torch.optim idiom
    def zero_grad(self, set_to_none: bool = False):
        if self._grads_clean:
            self._grads_clean = False            # cleared by the previous step(): nothing to do
        else:
            self.flat_g.zero_()
        for p, o in zip(self.params, self.offsets):
            if p.grad is None or p.grad.data_ptr() != self.flat_g.data_ptr() + 4 * o:
                p.grad = self.flat_g[o:o + p.numel()].view_as(p)

    def all_reduce_grads(self, group=None):
        """Data-parallel exchange: mean of the flat gradient buffer over the ranks (a no-op with peer arenas: the
        exchange is part of ``step``)."""
        if getattr(self, "peer", None) is not None:
            return
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(self.flat_g, op=dist.ReduceOp.AVG if self.flat_g.is_cuda else dist.ReduceOp.SUM,
                            group=group)
            if not self.flat_g.is_cuda:
                self.flat_g.div_(dist.get_world_size(group))

    @torch.no_grad()
    def step(self, clip_value: float = 0.):
        g = self.param_groups[0]
        self.step_count += 1
        self.sync_lr()
        if self.flat_bf16 is not None and not self.shadow_is_current():
            # parameters were edited in place since the last step: the Adam kernel refreshes the shadow only where it
            # changes a value (chunks that never saw a gradient are skipped), so bring all of it up to date first
            self.refresh_shadow()
        if self.peer is not None:
            pr, b = self.peer, self.peer["bufs"]
            call("nc_dp_adam_step", b["p"].ptr_array, b["g"].ptr_array, b["sh"].ptr_array if "sh" in b else None,
                 b["flags"].ptr_array, self.exp_avg, self.exp_avg_sq, self.numel, pr["rank"], pr["world"],
                 float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                 float(g["weight_decay"]), self.step_count, float(clip_value), pr["done"], pr["ctas"], self.dyn,
                 pr["aux"][0].ptr_array if pr["aux"] else None, pr["aux"][1] if pr["aux"] else None,
                 pr["aux"][2] if pr["aux"] else 0)
            self._grads_clean = True
            Fn.bump_param_epoch()
            return
        call("nc_adam_step_dyn", self.flat_p, self.flat_g, self.exp_avg, self.exp_avg_sq, self.numel, self.dyn,
             float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]),
             float(clip_value), self.flat_bf16, 1)
        self._grads_clean = True
        Fn.bump_param_epoch()

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "lr": self.param_groups[0]["lr"]}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.dyn[1:2].fill_(float(self.step_count + 1))
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.param_groups[0]["lr"] = sd["lr"]
