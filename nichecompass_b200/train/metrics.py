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


# ------------------------------------------------------------------------------------------------ device
_THR_T = {}


@torch.no_grad()
def eval_metrics_device(edge_recon_probs: torch.Tensor, edge_labels: torch.Tensor,
                        edge_same_cat_covariates_cat: Optional[List[torch.Tensor]] = None,
                        edge_incl: Optional[torch.Tensor] = None,
                        omics_pred_dict: Optional[dict] = None) -> dict:
    """``eval_metrics`` evaluated where the validation outputs already are: one descending sort
    of the edge probabilities, prefix sums over it and ONE read-back of the finished scalars, instead of copying
    every probability to the host for scikit-learn (train/metrics.py:107-134 in the reference).
      AUROC  rank statistic with average ranks for tied probabilities (= sklearn's trapezoid over distinct thresholds)
      AUPRC  sum over distinct thresholds of (recall_k - recall_k-1) * precision_k (= average_precision_score)
      sweeps accuracy / F1 at 0.01, 0.015, ..., 0.995 from two searches in the sorted probabilities
    fp64 on the device; results equal the host function to rounding (tests/test_eval_gpu.py)."""
    dev = edge_recon_probs.device
    out_t, names = [], []

    def put(name, v):
        names.append(name)
        out_t.append(v.reshape(()).double())
    probs, labels = edge_recon_probs.double(), edge_labels
    if omics_pred_dict is not None:
        for mod in ("rna", "atac"):
            if f"target_{mod}_preds" in omics_pred_dict:
                for ent in ("target", "source"):
                    d = omics_pred_dict[f"{ent}_{mod}"].double() - omics_pred_dict[f"{ent}_{mod}_preds"].double()
                    put(f"{ent}_{mod}_mse_score", (d * d).mean())
    if edge_same_cat_covariates_cat is not None:
        neg = labels == 0
        for i, same in enumerate(edge_same_cat_covariates_cat):
            s, p = same[neg].bool(), probs[neg]
            put(f"cat_covariate{i}_mean_sim_diff", p[~s].mean() - p[s].mean())
    if edge_incl is not None:
        keep = edge_incl.bool()
        probs, labels = probs[keep], labels[keep]
    pos = labels == 1
    n = probs.numel()
    n_pos = pos.sum().double()
    n_neg = n - n_pos
    order = torch.argsort(probs, descending=True, stable=True)
    ps, ys = probs[order], pos[order].double()
    tp = torch.cumsum(ys, 0)
    fp = torch.cumsum(1.0 - ys, 0)
    last = torch.ones(n, dtype=torch.bool, device=dev)                 # last element of each group of tied probabilities
    last[:-1] = ps[1:] != ps[:-1]
    tp_k, fp_k = tp[last], fp[last]
    zero = torch.zeros(1, dtype=torch.float64, device=dev)
    tp_prev, fp_prev = torch.cat((zero, tp_k[:-1])), torch.cat((zero, fp_k[:-1]))
    # ROC: trapezoids between consecutive distinct thresholds
    put("auroc_score", ((fp_k - fp_prev) * (tp_k + tp_prev)).sum() / (2.0 * n_pos * n_neg))
    put("auprc_score", ((tp_k - tp_prev) / n_pos * tp_k / (tp_k + fp_k)).sum())
    thr = _THR_T.get(dev)
    if thr is None:
        thr = _THR_T[dev] = torch.from_numpy(_THRESHOLDS).to(dev)
    asc_pos = torch.sort(probs[pos]).values
    asc_neg = torch.sort(probs[~pos]).values
    tpt = asc_pos.numel() - torch.searchsorted(asc_pos, thr, right=True)          # predicted positive: prob > thr
    fpt = asc_neg.numel() - torch.searchsorted(asc_neg, thr, right=True)
    tnt = asc_neg.numel() - fpt
    acc = (tpt + tnt).double() / max(n, 1)
    best = torch.argmax(acc)                                             # first maximum, as the reference's strict '>'
    put("best_acc_score", acc[best])
    put("best_acc_threshold", thr[best])
    fnt = asc_pos.numel() - tpt
    denom = (2 * tpt + fpt + fnt).double()
    f1 = torch.where(denom > 0, 2.0 * tpt.double() / denom.clamp_min(1.0), torch.zeros_like(denom))
    put("best_f1_score", f1.max())
    vals = torch.stack(out_t).cpu().tolist()                             # the one read-back
    out = dict(zip(names, vals))
    if out["best_acc_score"] <= 0:
        out["best_acc_score"] = out["best_acc_threshold"] = 0
    if out["best_f1_score"] <= 0:
        out["best_f1_score"] = 0
    return out
