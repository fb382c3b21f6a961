"""
Generic model functionality: public-attribute bookkeeping, ``save`` / ``load`` with the
reference's on-disk layout and the freeze / unfreeze flags of the reference-mapping (query
fine-tuning) workflow (reference: models/basemodelmixin.py:20-322).
"""
import inspect
import os
import pickle
import warnings
from typing import Optional

import numpy as np
import scipy.sparse as sp
import torch

from .utils import initialize_model, load_saved_files, validate_var_names


def _is_adata(v) -> bool:
    return hasattr(v, "obsp") and hasattr(v, "varm") and hasattr(v, "n_vars")


class BaseModelMixin:
    def _get_user_attributes(self) -> list:
        members = inspect.getmembers(self, lambda m: not inspect.isroutine(m))
        return [(k, v) for k, v in members if not (k.startswith("__") and k.endswith("__"))]

    def _get_public_attributes(self) -> dict:
        """Attributes whose name ends with an underscore are public by convention."""
        return {k: v for k, v in self._get_user_attributes() if k[-1] == "_"}

    def _get_init_params(self, locals: dict) -> dict:
        """Constructor arguments with their passed values (AnnData objects excluded)."""
        names = list(inspect.signature(self.__init__).parameters)
        return {k: v for k, v in locals.items() if k in names and not _is_adata(v)}

    # ------------------------------------------------------------------ save
    def save(self,
             dir_path: str,
             overwrite: bool = False,
             save_adata: bool = False,
             adata_file_name: str = "adata.h5ad",
             save_adata_atac: bool = False,
             adata_atac_file_name: str = "adata_atac.h5ad",
             **anndata_write_kwargs):
        """Writes ``model_params.pt`` (module state_dict, reference key names), ``attr.pkl`` (all
        public attributes incl. ``init_params_``) and ``var_names.csv``; the optimiser state is not
        saved (as in the reference)."""
        if not os.path.exists(dir_path) or overwrite:
            os.makedirs(dir_path, exist_ok=overwrite)
        else:
            raise ValueError(f"Directory '{dir_path}' already exists."
                             "Please provide another directory for saving.")
        if save_adata:
            if not hasattr(self.adata, "write"):
                raise NotImplementedError("save_adata needs a real anndata.AnnData (``.write``)")
            if self.adata.obsp["spatial_connectivities"] is not None:
                self.adata.obsp["spatial_connectivities"] = sp.csr_matrix(
                    self.adata.obsp["spatial_connectivities"])
            self.adata.write(os.path.join(dir_path, adata_file_name), **anndata_write_kwargs)
        if save_adata_atac:
            self.adata_atac.write(os.path.join(dir_path, adata_atac_file_name))
        var_names = self.adata.var_names.astype(str).to_numpy()
        public_attributes = self._get_public_attributes()
        state = {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()}
        torch.save(state, os.path.join(dir_path, "model_params.pt"))
        np.savetxt(os.path.join(dir_path, "var_names.csv"), var_names, fmt="%s")
        with open(os.path.join(dir_path, "attr.pkl"), "wb") as fh:
            pickle.dump(public_attributes, fh)

    # ------------------------------------------------------------------ load
    @classmethod
    def load(cls,
             dir_path: str,
             adata=None,
             adata_atac=None,
             adata_file_name: str = "adata.h5ad",
             adata_atac_file_name: Optional[str] = None,
             use_cuda: bool = False,
             n_addon_gps: int = 0,
             gp_names_key: Optional[str] = None,
             genes_idx_key: Optional[str] = None,
             unfreeze_all_weights: bool = False,
             unfreeze_addon_gp_weights: bool = False,
             unfreeze_cat_covariates_embedder_weights: bool = False):
        """Instantiate a model from a saved directory; optionally grow it by ``n_addon_gps`` new
        add-on gene programs and by covariate categories unseen at training time, then freeze all
        weights and selectively unfreeze (query fine-tuning)."""
        load_adata = adata is None
        load_adata_atac = (adata_atac is None) and (adata_atac_file_name is not None)
        use_cuda = use_cuda and torch.cuda.is_available()
        map_location = torch.device("cpu") if not use_cuda else None
        state_dict, var_names, attr_dict, new_adata, new_adata_atac = load_saved_files(
            dir_path, load_adata, adata_file_name, load_adata_atac, adata_atac_file_name, map_location)
        adata = new_adata if new_adata is not None else adata
        adata_atac = new_adata_atac if new_adata_atac is not None else adata_atac
        validate_var_names(adata, var_names)

        if n_addon_gps != 0:
            if genes_idx_key not in adata.uns:
                raise ValueError("Please specifiy a valid 'genes_idx_key' if 'n_addon_gps' > 0, so that all "
                                 "genes can be included in the genes idx.")
            adata.uns[genes_idx_key] = np.arange(adata.n_vars * 2)

        # categories of the query data that the stored model has not seen
        cats = attr_dict["cat_covariates_cats_"]
        keys = attr_dict["init_params_"]["cat_covariates_keys"]
        new_cats = []
        if keys is not None:
            for i, key in enumerate(keys):
                for c in adata.obs[key].unique().tolist():
                    if c not in cats[i]:
                        new_cats.append(c)
                        cats[i].append(c)
        attr_dict["init_params_"]["cat_covariates_cats"] = cats

        if n_addon_gps != 0:
            # (the reference indexes "n_addon_gps_" / init_params_["n_addon_gps"], names its own
            # constructor never stores -- models/basemodelmixin.py:241-242; the stored names are used here)
            attr_dict["n_addon_gp_"] += n_addon_gps
            attr_dict["init_params_"]["n_addon_gp"] += n_addon_gps
            if gp_names_key is None:
                raise ValueError("Please specify 'gp_names_key' so that addon gps can be added to the gene "
                                 "program list.")
            gps = list(adata.uns[gp_names_key])
            start = int(gps[-1][-1]) + 1 if any("addon_GP_" in gp for gp in gps) else 0
            adata.uns[gp_names_key] = np.array(gps + [f"addon_GP_{start + i}" for i in range(n_addon_gps)])

        model = initialize_model(cls, adata, attr_dict, adata_atac)
        for attr, val in attr_dict.items():
            setattr(model, attr, val)
        if use_cuda:
            model.model.cuda()
        if n_addon_gps != 0 or len(new_cats) > 0:
            model.model.load_and_expand_state_dict(state_dict)
        else:
            model.model.load_state_dict(state_dict)
        model.model.eval()

        for _, p in model.model.named_parameters():
            p.requires_grad = False
        model.freeze_ = True
        if unfreeze_all_weights:
            for _, p in model.model.named_parameters():
                p.requires_grad = True
            model.freeze_ = False
        if unfreeze_addon_gp_weights:
            for name, p in model.model.named_parameters():
                if "addon" in name or "theta" in name or "aggregator" in name:
                    p.requires_grad = True
        if unfreeze_cat_covariates_embedder_weights:
            for name, p in model.model.named_parameters():
                if ("cat_covariate" in name) and ("embedder" in name):
                    p.requires_grad = True
        if model.freeze_ and not model.is_trained_:
            raise ValueError("The model has not been pre-trained and therefore weights should not be frozen.")
        return model

    def _check_if_trained(self, warn: bool = True):
        message = ("Trying to query inferred values from an untrained model. Please train the model first.")
        if not self.is_trained_:
            if warn:
                warnings.warn(message)
            else:
                raise RuntimeError(message)
