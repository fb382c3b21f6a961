"""
Add-on masked layer (reference: nn/layers.py:13-144): prior-GP masked linear + add-on-GP masked
linear + categorical-covariate embedding linear, summed and passed through an activation.
Here the three contractions are one packed GEMM (``MaskedDecoderFn``).
"""
from typing import List, Optional

import torch
import torch.nn as nn

from ..functional import MaskedDecoderFn, new_cache_key
from .layercomponents import MaskedLinear


class AddOnMaskedLayer(nn.Module):
    def __init__(self,
                 n_input: int,
                 n_output: int,
                 mask: torch.Tensor,
                 addon_mask: Optional[torch.Tensor],
                 masked_features_idx: List,
                 bias: bool = False,
                 n_addon_input: int = 0,
                 n_cat_covariates_embed_input: int = 0,
                 activation: nn.Module = nn.Softmax(dim=-1)):
        super().__init__()
        self.n_input = n_input
        self.n_output = n_output
        self.n_addon_input = n_addon_input
        self.n_cat_covariates_embed_input = n_cat_covariates_embed_input
        self.masked_features_idx = masked_features_idx
        self.masked_l = MaskedLinear(n_input=n_input, n_output=n_output, mask=mask, bias=bias)
        if n_addon_input != 0:
            self.addon_l = MaskedLinear(n_input=n_addon_input, n_output=n_output, mask=addon_mask, bias=False)
        if n_cat_covariates_embed_input != 0:
            self.cat_covariates_embed_l = nn.Linear(n_cat_covariates_embed_input, n_output, bias=False)
        self.activation = activation
        self.compute_dtype = torch.float32
        self._embed_mask = None
        self._key = new_cache_key(self)

    def _embed_col_mask(self) -> Optional[torch.Tensor]:
        """Without add-on GPs the embedding only feeds the masked features (nn/layers.py:134-142)."""
        if self.n_addon_input != 0:
            return None
        w = self.cat_covariates_embed_l.weight
        if self._embed_mask is None or self._embed_mask.device != w.device:
            m = torch.zeros(self.n_output, 1, device=w.device, dtype=w.dtype)
            m[torch.as_tensor(list(self.masked_features_idx), dtype=torch.long, device=w.device)] = 1.
            self._embed_mask = m.expand(-1, self.n_cat_covariates_embed_input).contiguous()
        return self._embed_mask

    def logits(self, input: torch.Tensor, dynamic_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Pre-activation output [n, n_output]: input columns are (prior GPs | add-on GPs |
        covariate embedding), exactly the split of nn/layers.py:102-121."""
        params, masks, dyn = [self.masked_l.weight], [self.masked_l.mask], [None]
        if dynamic_mask is not None:
            dyn[0] = self.masked_l.apply_dynamic_mask_(dynamic_mask[:self.n_input, :])
        if self.n_addon_input != 0:
            params.append(self.addon_l.weight)
            masks.append(self.addon_l.mask)
            dyn.append(self.addon_l.apply_dynamic_mask_(dynamic_mask[self.n_input:, :])
                       if dynamic_mask is not None else None)
        if self.n_cat_covariates_embed_input != 0:
            params.append(self.cat_covariates_embed_l.weight)
            masks.append(self._embed_col_mask())
            dyn.append(None)
        return MaskedDecoderFn.apply(input, self.compute_dtype, (self._key, "packed"), dyn, *params, *masks)

    def forward(self, input: torch.Tensor, dynamic_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Reference signature.  The activation here is the module passed at construction
        (``nn.Softmax``): the training path never takes this route -- ``VGPGAE`` consumes
        ``logits()`` and the fused softmax / negative-binomial kernel."""
        return self.activation(self.logits(input, dynamic_mask))
