"""
Masked linear layer component (reference: nn/layercomponents.py:12-81).
"""
from typing import Optional

import torch
import torch.nn as nn

from ..functional import MaskedDecoderFn, new_cache_key


class MaskedLinear(nn.Linear):
    """Linear layer whose connections are restricted by a static binary mask [n_input, n_output]
    and an optional dynamic mask.  The contraction runs on the library GEMM kernels with
    ``weight * mask`` packed once per parameter version; the mask is re-applied to the weight
    gradient in the GEMM epilogue, so masked weights never receive gradient (they are zeroed at
    construction exactly as in the reference, nn/layercomponents.py:51)."""

    def __init__(self, n_input: int, n_output: int, mask: torch.Tensor, bias=False):
        if n_input != mask.shape[0] or n_output != mask.shape[1]:
            raise ValueError("Incorrect shape of the mask. Mask should have dim (n_input x n_output). "
                             f"Please provide a mask with dimensions ({n_input} x {n_output}).")
        if bias:
            raise NotImplementedError("the reference never builds a MaskedLinear with bias")
        super().__init__(n_input, n_output, bias)
        self.register_buffer("mask", mask.t().to(self.weight.dtype).contiguous())
        self.weight.data *= self.mask
        self.compute_dtype = torch.float32
        self._key = new_cache_key(self)

    def apply_dynamic_mask_(self, dynamic_mask: torch.Tensor) -> torch.Tensor:
        """``weight.data *= dynamic_mask.t()`` -- the permanent side effect of the reference
        forward (nn/layercomponents.py:74-77).  Returns the transposed float mask."""
        # Applying the same mask again is the identity (masked weights are zero, their gradient is masked, Adam leaves
        # them at zero): the [n_output, n_input] float copy and the in-place product are redone only when the mask
        # changed.  At config 4 (50 k peaks x 1.6 k programs) that was 0.6 GB of element-wise traffic per decoder and
        # iteration, plus a refresh of the whole bf16 shadow because the in-place product bumps the weight's version.
        key = (dynamic_mask.data_ptr(), dynamic_mask._version, tuple(dynamic_mask.shape), str(self.weight.device))
        hit = getattr(self, "_dyn_cache", None)
        # (the weight's version counter sees torch-side edits such as load_state_dict; the Adam kernel writes through
        # the raw pointer and cannot un-mask a weight)
        if hit is not None and hit[0] == key and hit[3] == self.weight._version:
            return hit[1]
        dm = dynamic_mask.t().to(self.weight.device, self.weight.dtype).contiguous()
        with torch.no_grad():
            self.weight.mul_(dm)
        # the mask tensor is kept so that its address cannot be reused while the entry lives
        self._dyn_cache = (key, dm, dynamic_mask, self.weight._version)
        return dm

    def forward(self, input: torch.Tensor, dynamic_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        dm = self.apply_dynamic_mask_(dynamic_mask) if dynamic_mask is not None else None
        return MaskedDecoderFn.apply(input, self.compute_dtype, (self._key, "w"), [dm], self.weight, self.mask)
