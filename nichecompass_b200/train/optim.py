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


class FlatAdam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 0.,
                 groups: Optional[Sequence[Sequence[torch.nn.Parameter]]] = None,
                 bf16_shadow: bool = False):
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
        self.flat_p = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_bf16 = torch.zeros(n, dtype=torch.bfloat16, device=dev) if bf16_shadow else None
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
        """Data-parallel exchange: mean of the flat gradient buffer over the ranks."""
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
        if self.flat_bf16 is not None and not self.shadow_is_current():
            self._shadow_versions = self._versions()       # the kernel below rewrites the whole shadow
        call("nc_adam_step", self.flat_p, self.flat_g, self.exp_avg, self.exp_avg_sq, self.numel,
             float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
             float(g["weight_decay"]), self.step_count, float(clip_value), self.flat_bf16, 1)
        self._grads_clean = True
        Fn.bump_param_epoch()

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "lr": self.param_groups[0]["lr"]}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.param_groups[0]["lr"] = sd["lr"]
