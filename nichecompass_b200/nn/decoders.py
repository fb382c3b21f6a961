"""
Decoders of the VGPGAE (reference: nn/decoders.py): cosine-similarity graph decoder and the
masked omics feature decoder (``modality`` "rna" for gene expression, "atac" for chromatin
accessibility).
"""
from typing import List, Literal, Optional

import torch
import torch.nn as nn

from ..functional import CosineFn
from .layers import AddOnMaskedLayer


class CosineSimGraphDecoder(nn.Module):
    """Edge logits = cosine similarity between the first and the second half of the rows of
    ``z`` (nn/decoders.py:14-62)."""

    def __init__(self, dropout_rate: float = 0.):
        super().__init__()
        print(f"COSINE SIM GRAPH DECODER -> dropout_rate: {dropout_rate}")
        self.dropout_rate = dropout_rate
        self.dropout = nn.Dropout(dropout_rate)

    def forward(self, z: torch.Tensor) -> torch.Tensor:
        if self.dropout_rate != 0.:
            z = self.dropout(z)                     # nn/decoders.py:56 (identity in eval mode)
        return CosineFn.apply(z.float())


class MaskedOmicsFeatureDecoder(nn.Module):
    """Gene-program -> omics feature decoder (nn/decoders.py:65-169).  ``forward`` keeps the
    reference semantics: negative-binomial means = exp(log library size) * softmax over ALL
    output features.  ``logits`` exposes the pre-softmax contraction that ``VGPGAE`` feeds to
    the fused softmax + NB-loss kernel."""

    def __init__(self,
                 modality: Literal["rna", "atac"],
                 entity: Literal["target", "source"],
                 n_prior_gp_input: int,
                 n_addon_gp_input: int,
                 n_cat_covariates_embed_input: int,
                 n_output: int,
                 mask: torch.Tensor,
                 addon_mask: Optional[torch.Tensor],
                 masked_features_idx: List,
                 recon_loss: Literal["nb"]):
        super().__init__()
        print(f"MASKED {entity.upper()} {modality.upper()} DECODER -> "
              f"n_prior_gp_input: {n_prior_gp_input}, n_addon_gp_input: {n_addon_gp_input}, "
              f"n_cat_covariates_embed_input: {n_cat_covariates_embed_input}, n_output: {n_output}")
        self.masked_features_idx = masked_features_idx
        self.recon_loss = recon_loss
        self.nb_means_normalized_decoder = AddOnMaskedLayer(
            n_input=n_prior_gp_input,
            n_addon_input=n_addon_gp_input,
            n_cat_covariates_embed_input=n_cat_covariates_embed_input,
            n_output=n_output,
            bias=False,
            mask=mask,
            addon_mask=addon_mask,
            masked_features_idx=masked_features_idx,
            activation=nn.Softmax(dim=-1))

    def logits(self, z: torch.Tensor, cat_covariates_embed: Optional[torch.Tensor] = None,
               dynamic_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        if cat_covariates_embed is not None:
            z = torch.cat((z, cat_covariates_embed), dim=-1)
        return self.nb_means_normalized_decoder.logits(z, dynamic_mask)

    def forward(self, z: torch.Tensor, log_library_size: torch.Tensor,
                cat_covariates_embed: Optional[torch.Tensor] = None,
                dynamic_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        if cat_covariates_embed is not None:
            z = torch.cat((z, cat_covariates_embed), dim=-1)
        normalized = self.nb_means_normalized_decoder(input=z, dynamic_mask=dynamic_mask)
        return torch.exp(log_library_size) * normalized
