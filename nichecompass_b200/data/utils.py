"""Helpers of the ``data`` subpackage (reference: data/utils.py:12-49)."""
import numpy as np


def encode_labels(adata, label_encoder: dict, label_key: str = "str") -> np.ndarray:
    """Integer codes of ``adata.obs[label_key]`` under ``label_encoder``; labels unknown to the
    encoder get -1 (with the reference's warning)."""
    labels = np.asarray(adata.obs[label_key])
    out = np.zeros(adata.shape[0])
    known = set(label_encoder.keys())
    unknown = [u for u in np.unique(labels) if u not in known]
    if unknown:
        print(f"Warning: Labels in adata.obs[{label_key}] are not a subset of the label encoder!")
        print("Therefore integer value of those labels is set to '-1'.")
        for u in unknown:
            out[labels == u] = -1
    for label, code in label_encoder.items():
        out[labels == label] = code
    return out
