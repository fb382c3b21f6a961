"""
Validation metrics of the trainer (reference: train/metrics.py:14-135).  Host side, once per
epoch, outside the hot path: AUROC / AUPRC through scikit-learn as in the reference, the two
threshold sweeps (accuracy and F1 over 0.01, 0.015, ..., 0.995) as one vectorised comparison
instead of 2 x 198 scikit-learn calls.
"""
from typing import List, Optional, Union

import numpy as np
import sklearn.metrics as skm
import torch

_THRESHOLDS = np.arange(0.01, 1, 0.005)


def _np(a):
    return a.detach().cpu().numpy() if isinstance(a, torch.Tensor) else a


def eval_metrics(edge_recon_probs: Union[torch.Tensor, np.ndarray],
                 edge_labels: Union[torch.Tensor, np.ndarray],
                 edge_same_cat_covariates_cat: Optional[List[Union[torch.Tensor, np.ndarray]]] = None,
                 edge_incl: Optional[Union[torch.Tensor, np.ndarray]] = None,
                 omics_pred_dict: Optional[dict] = None) -> dict:
    """``auroc_score``, ``auprc_score``, ``best_acc_score`` (+ ``best_acc_threshold``),
    ``best_f1_score`` over a balanced sample of positive / negative edges, per-covariate
    ``cat_covariate{i}_mean_sim_diff`` over the negative edges, and ``{entity}_{modality}_mse_score``
    when predictions are given."""
    out = {}
    probs, labels, incl = _np(edge_recon_probs), _np(edge_labels), _np(edge_incl)

    if omics_pred_dict is not None:
        for mod in ("rna", "atac"):
            if f"target_{mod}_preds" in omics_pred_dict:
                for ent in ("target", "source"):
                    truth = np.asarray(_np(omics_pred_dict[f"{ent}_{mod}"]), dtype=np.float64)
                    pred = np.asarray(_np(omics_pred_dict[f"{ent}_{mod}_preds"]), dtype=np.float64)
                    out[f"{ent}_{mod}_mse_score"] = float(np.mean((truth - pred) ** 2))

    if edge_same_cat_covariates_cat is not None:
        neg = labels == 0
        for i, same in enumerate(edge_same_cat_covariates_cat):
            same = _np(same)[neg]
            p = probs[neg]
            out[f"cat_covariate{i}_mean_sim_diff"] = np.mean(p[same == 0]) - np.mean(p[same == 1])

    if incl is not None:
        incl = incl.astype(bool)
        probs, labels = probs[incl], labels[incl]

    out["auroc_score"] = skm.roc_auc_score(labels, probs)
    out["auprc_score"] = skm.average_precision_score(labels, probs)

    # threshold sweeps: pred[t, e] = probs[e] > thr[t]
    pos = labels == 1
    n = labels.shape[0]
    order = np.sort(probs[pos])
    order_neg = np.sort(probs[~pos])
    # number of positives / negatives predicted positive at each threshold
    tp = order.shape[0] - np.searchsorted(order, _THRESHOLDS, side="right")
    fp = order_neg.shape[0] - np.searchsorted(order_neg, _THRESHOLDS, side="right")
    tn = order_neg.shape[0] - fp
    acc = (tp + tn) / max(n, 1)
    best = int(np.argmax(acc))                     # first maximum, as the strict '>' sweep
    out["best_acc_score"] = float(acc[best]) if acc[best] > 0 else 0
    out["best_acc_threshold"] = float(_THRESHOLDS[best]) if acc[best] > 0 else 0
    fn = order.shape[0] - tp
    denom = 2 * tp + fp + fn
    f1 = np.where(denom > 0, 2 * tp / np.maximum(denom, 1), 0.0)
    out["best_f1_score"] = float(f1.max()) if f1.max() > 0 else 0
    return out
