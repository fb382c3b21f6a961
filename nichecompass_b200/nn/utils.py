"""Helpers of the ``nn`` subpackage (reference: nn/utils.py:8-36)."""
import torch

from ..functional import CosineFn


def compute_cosine_similarity(tensor1: torch.Tensor, tensor2: torch.Tensor, eps: float = 1e-8) -> torch.Tensor:
    """Row-wise cosine similarity of two [n_obs, n_features] tensors with norms clamped at
    ``eps`` (fixed at 1e-8 in the kernel, the reference default).  Differentiable; runs the
    fused cosine kernel on the stacked rows."""
    if eps != 1e-8:
        raise ValueError("the CUDA cosine kernel implements the reference default eps=1e-8 only")
    return CosineFn.apply(torch.cat((tensor1, tensor2), dim=0).float())
