"""
Loss functions of the VGPGAE (reference: modules/losses.py), same names and argument meaning,
evaluated by the CUDA kernels behind ``nichecompass_b200.functional``.
"""
from typing import List, Literal, Optional

import torch
import torch.nn as nn

from ..functional import BCEFn, KLFn, NBLossFusedFn, WeightRegFn


def compute_kl_reg_loss(mu: torch.Tensor, logstd: torch.Tensor) -> torch.Tensor:
    """-0.5 * mean_rows sum_cols (1 + 2 logstd - mu^2 - exp(logstd)^2)  (modules/losses.py:255-284)."""
    return KLFn.apply(mu, logstd)


def compute_edge_recon_loss(edge_recon_logits: torch.Tensor,
                            edge_recon_labels: torch.Tensor,
                            edge_incl: Optional[torch.Tensor] = None) -> torch.Tensor:
    """BCE-with-logits with pos_weight = #neg / #pos over the included edges
    (modules/losses.py:123-166)."""
    incl = None if edge_incl is None else edge_incl.to(torch.uint8).contiguous()
    return BCEFn.apply(edge_recon_logits, edge_recon_labels, incl)


def compute_omics_recon_nb_loss_from_logits(logits: torch.Tensor, x_full: torch.Tensor,
                                            theta_param: torch.Tensor, reconstructed: torch.Tensor):
    """Fused form used by ``VGPGAE``: softmax over all features of ``logits`` -> library-size
    scaling -> selection of the ``reconstructed`` (uint8 [F]) columns -> mean NB negative
    log-likelihood with theta = exp(theta_param).  Equals
    ``compute_omics_recon_nb_loss(x[:, idx], (exp(log_lib) * softmax(logits))[:, idx], exp(theta_param[idx]))``
    of the reference (modules/vgpgae.py:714-749,1065-1077; modules/losses.py:169-212).
    Returns (loss, nb_means over all features)."""
    return NBLossFusedFn.apply(logits, theta_param, x_full, reconstructed)


def compute_omics_recon_nb_loss(x: torch.Tensor, mu: torch.Tensor, theta: torch.Tensor, eps: float = 1e-8):
    """Mean over rows of the summed negative-binomial negative log-likelihood for given means ``mu`` [R, F] and
    inverse dispersions ``theta`` [F] (modules/losses.py:169-212), differentiable w.r.t. ``mu`` and ``theta``.
    ``VGPGAE`` itself uses ``compute_omics_recon_nb_loss_from_logits`` (the same arithmetic fused with the decoder
    softmax and the column selection)."""
    from ..functional import NBLossMeansFn
    if theta.dim() != 1 or theta.numel() != mu.size(-1):
        theta = theta.expand(mu.size(-1)) if theta.numel() == 1 else theta.reshape(-1)
    return NBLossMeansFn.apply(x, mu, theta, eps)


def _rna_decoder_weights(model: nn.Module, layer_name: str):
    out = []
    for name, p in model.named_parameters():
        if f"rna_decoder.nb_means_normalized_decoder.{layer_name}" in name:
            out.append((name, p))
    return out


def compute_gp_group_lasso_reg_loss(model: nn.Module) -> torch.Tensor:
    """sum over prior gene programs of sqrt(#unmasked weights) * ||W[:, gp]||_2 for the target and
    source rna decoders (modules/losses.py:215-252)."""
    items = _rna_decoder_weights(model, "masked_l")
    weights = [p for _, p in items]
    masks = []
    for name, _ in items:
        mod = model
        for part in name.split(".")[:3]:
            mod = getattr(mod, part)
        masks.append(mod.mask)
    return WeightRegFn.apply(0.0, 1.0, [None] * len(weights), masks, *weights)


def compute_gp_l1_reg_loss(model: nn.Module,
                           gp_type: Literal["prior", "addon"],
                           l1_targets_mask: Optional[torch.Tensor] = None,
                           l1_sources_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """sum |W[mask]| over the target and source rna decoder weights of the prior (``masked_l``)
    or add-on (``addon_l``) gene programs (modules/losses.py:287-339); a ``None`` mask selects
    every entry."""
    layer_name = "masked_l" if gp_type == "prior" else "addon_l"
    items = _rna_decoder_weights(model, layer_name)
    weights = [p for _, p in items]
    masks = []
    for name, p in items:
        m = l1_targets_mask if "target" in name else l1_sources_mask
        masks.append(None if m is None else m.to(p.device, torch.uint8).contiguous())
    return WeightRegFn.apply(1.0, 0.0, masks, [None] * len(weights), *weights)


def compute_cat_covariates_contrastive_loss(edge_recon_logits: torch.Tensor,
                                            edge_recon_labels: torch.Tensor,
                                            edge_same_cat_covariates_cat: Optional[List[torch.Tensor]] = None,
                                            contrastive_logits_pos_ratio: float = 0.,
                                            contrastive_logits_neg_ratio: float = 0.) -> torch.Tensor:
    """modules/losses.py:15-120.  Returns 0 exactly like the reference when no covariates are
    given or both ratios are 0 (the default of ``NicheCompass.train``).  Selection of the
    top / bottom logits among the cross-category pairs is index plumbing (``torch.topk``); the loss
    and its gradient are the BCE kernel without the positive weight."""
    if edge_same_cat_covariates_cat is None or (contrastive_logits_pos_ratio == 0 and
                                                contrastive_logits_neg_ratio == 0):
        return torch.tensor(0., device=edge_recon_logits.device)
    import math
    total = torch.zeros((), dtype=torch.float32, device=edge_recon_logits.device)
    logits = edge_recon_logits
    for same in edge_same_cat_covariates_cat:
        diff = ~same
        diff_logits = logits.detach()[diff]
        n_diff = int(diff_logits.numel())                    # host read-back (the reference's len())
        n_pos = math.ceil(contrastive_logits_pos_ratio * n_diff)
        n_neg = math.ceil(contrastive_logits_neg_ratio * n_diff)
        pos_thresh = torch.topk(diff_logits, n_pos).values[-1] if n_pos > 0 else float("inf")
        neg_thresh = torch.topk(diff_logits, n_neg, largest=False).values[-1] if n_neg > 0 else float("-inf")
        pos_ex = diff & (logits.detach() >= pos_thresh)
        neg_ex = diff & (logits.detach() <= neg_thresh)
        # the reference relabels IN PLACE (modules/losses.py:96-101): later covariates see the new labels
        edge_recon_labels[pos_ex] = 1
        edge_recon_labels[neg_ex] = 0
        sel = (pos_ex | neg_ex).to(torch.uint8).contiguous()
        total = total + BCEFn.apply(logits, edge_recon_labels.clone(), sel, False)
    return total
