"""
Helpers of the ``models`` subpackage (reference: models/utils.py:18-149): reading a saved model
directory and re-instantiating a model from its stored constructor arguments.
"""
import logging
import os
import pickle
from collections import OrderedDict
from typing import Optional, Tuple

import numpy as np
import torch

logger = logging.getLogger(__name__)


def _read_adata(path: str):
    try:
        import anndata as ad
    except ImportError as e:                                     # anndata is not part of this image
        raise ImportError("reading a stored .h5ad needs the `anndata` package; pass `adata` explicitly") from e
    return ad.read_h5ad(path)


def load_saved_files(dir_path: str,
                     load_adata: bool,
                     adata_file_name: str = "adata.h5ad",
                     load_adata_atac: bool = False,
                     adata_atac_file_name: str = "adata_atac.h5ad",
                     map_location=None) -> Tuple[OrderedDict, np.ndarray, dict, object, object]:
    """(state dict, var names, attribute dict, adata | None, adata_atac | None) from the files
    ``model_params.pt`` / ``var_names.csv`` / ``attr.pkl`` (+ optional ``.h5ad``) of ``dir_path``."""
    adata = adata_atac = None
    if load_adata:
        path = os.path.join(dir_path, adata_file_name)
        if not os.path.exists(path):
            raise ValueError("Dir path contains no saved anndata and no adata was passed.")
        adata = _read_adata(path)
    if load_adata_atac:
        path = os.path.join(dir_path, adata_atac_file_name)
        if not os.path.exists(path):
            raise ValueError("Dir path contains no saved 'adata_atac' and no 'adata_atac' was passed.")
        adata_atac = _read_adata(path)
    model_state_dict = torch.load(os.path.join(dir_path, "model_params.pt"), map_location=map_location)
    var_names = np.genfromtxt(os.path.join(dir_path, "var_names.csv"), delimiter=",", dtype=str)
    with open(os.path.join(dir_path, "attr.pkl"), "rb") as fh:
        attr_dict = pickle.load(fh)
    return model_state_dict, var_names, attr_dict, adata, adata_atac


def validate_var_names(adata, source_var_names) -> bool:
    user_var_names = np.asarray(adata.var_names.astype(str))
    ok = np.array_equal(np.asarray(source_var_names), user_var_names)
    if not ok:
        logger.warning("The ´var_names´ of the passed in adata do not match the ´var_names´ of the adata used "
                       "to train the model. For valid results, the var_names need to be the same and in the "
                       "same order as the adata used to train the model.")
    return ok


def initialize_model(cls, adata, attr_dict: dict, adata_atac=None):
    """Instantiate ``cls`` from the ``init_params_`` stored with a saved model (dict-valued
    entries are the expanded ``**kwargs`` of the original call)."""
    if "init_params_" not in attr_dict:
        raise ValueError("No init_params_ were saved by the model.")
    init_params = attr_dict.pop("init_params_")
    plain = {k: v for k, v in init_params.items() if not isinstance(v, dict)}
    extra = {k: v for d in init_params.values() if isinstance(d, dict) for k, v in d.items()}
    if adata_atac is None:
        return cls(adata=adata, **plain, **extra)
    return cls(adata=adata, adata_atac=adata_atac, **plain, **extra)
