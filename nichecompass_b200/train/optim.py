"""
Flat-buffer Adam on the ``nc_adam_step`` kernel (reference: train/trainer.py:339-342,414-421 --
``torch.optim.Adam(params, lr, weight_decay)``, ``zero_grad``, optional ``clip_grad_value_``,
``step``).

All trainable parameters are re-homed as views into ONE contiguous fp32 buffer, their gradients
as views into a second one, so that
  * ``zero_grad`` is one memset,
  * the data-parallel gradient exchange is one NCCL all-reduce on the flat gradient buffer,
  * the optimiser step (clip + Adam + bias correction) is one kernel launch.
Update rule, bias correction and epsilon placement are torch.optim.Adam's.  One documented
deviation: a parameter that never receives a gradient still takes part in the update with a
zero gradient (torch skips it), which only matters together with ``weight_decay != 0``.
"""
from typing import Iterable, List

import torch

from .. import functional as Fn
from .._lib import call


class FlatAdam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 0.):
        self.params: List[torch.nn.Parameter] = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatAdam needs CUDA parameters (nichecompass_b200 has no CPU path)")
        # 16-byte aligned segments (4 floats)
        self.offsets, n = [], 0
        for p in self.params:
            self.offsets.append(n)
            n += (p.numel() + 3) // 4 * 4
        self.numel = n
        self.flat_p = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg = torch.zeros(n, dtype=torch.float32, device=dev)
        self.exp_avg_sq = torch.zeros(n, dtype=torch.float32, device=dev)
        with torch.no_grad():
            for p, o in zip(self.params, self.offsets):
                view = self.flat_p[o:o + p.numel()].view_as(p)
                view.copy_(p.data)
                p.data = view
                p.grad = self.flat_g[o:o + p.numel()].view_as(p)
        # the one param group keeps the reference's ``param_group["lr"] *= factor`` idiom working
        self.param_groups = [{"lr": lr, "betas": betas, "eps": eps, "weight_decay": weight_decay,
                              "params": self.params}]
        self.step_count = 0

    def zero_grad(self, set_to_none: bool = False):
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
        call("nc_adam_step", self.flat_p, self.flat_g, self.exp_avg, self.exp_avg_sq, self.numel,
             float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
             float(g["weight_decay"]), self.step_count, float(clip_value), None)
        Fn.bump_param_epoch()

    def state_dict(self):
        return {"step": self.step_count, "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(),
                "lr": self.param_groups[0]["lr"]}

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
        self.param_groups[0]["lr"] = sd["lr"]
