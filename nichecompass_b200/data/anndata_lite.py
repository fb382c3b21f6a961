"""
Minimal AnnData duck type.

The reference consumes ``anndata.AnnData`` purely through attribute access
(/root/reference/src/nichecompass/models/nichecompass.py:301-545, data/datasets.py:54-110:
``.X .layers .obsp .obsm .varm .uns .obs .var .var_names .n_vars .n_obs .shape``).  The
``anndata`` wheel is not available in this image, so synthetic data and tests use this class;
a real ``anndata.AnnData`` works unchanged wherever an ``AnnDataLite`` is accepted.
"""
from typing import Optional

import numpy as np
import pandas as pd
import scipy.sparse as sp


class AnnDataLite:
    def __init__(self, X, obs: Optional[pd.DataFrame] = None, var: Optional[pd.DataFrame] = None):
        self.X = X
        n_obs, n_vars = X.shape
        self.obs = obs if obs is not None else pd.DataFrame(index=[f"cell{i}" for i in range(n_obs)])
        self.var = var if var is not None else pd.DataFrame(index=[f"gene{i}" for i in range(n_vars)])
        self.layers, self.obsp, self.obsm, self.varm, self.uns = {}, {}, {}, {}, {}

    @property
    def shape(self):
        return self.X.shape

    @property
    def n_obs(self):
        return self.X.shape[0]

    @property
    def n_vars(self):
        return self.X.shape[1]

    @property
    def var_names(self):
        return self.var.index

    @property
    def obs_names(self):
        return self.obs.index

    def copy(self):
        out = AnnDataLite(self.X.copy(), self.obs.copy(), self.var.copy())
        for name in ("layers", "obsp", "obsm", "varm", "uns"):
            setattr(out, name, {k: (v.copy() if hasattr(v, "copy") else v)
                                for k, v in getattr(self, name).items()})
        return out

    def __repr__(self):
        return f"AnnDataLite(n_obs={self.n_obs}, n_vars={self.n_vars}, nnz={sp.csr_matrix(self.X).nnz})"
