from .basemodulemixin import BaseModuleMixin
from .losses import (compute_cat_covariates_contrastive_loss,
                     compute_edge_recon_loss,
                     compute_gp_group_lasso_reg_loss,
                     compute_gp_l1_reg_loss,
                     compute_kl_reg_loss,
                     compute_omics_recon_nb_loss)
from .vgaemodulemixin import VGAEModuleMixin
from .vgpgae import VGPGAE

__all__ = ["BaseModuleMixin",
           "compute_cat_covariates_contrastive_loss",
           "compute_edge_recon_loss",
           "compute_gp_group_lasso_reg_loss",
           "compute_gp_l1_reg_loss",
           "compute_kl_reg_loss",
           "compute_omics_recon_nb_loss",
           "VGAEModuleMixin",
           "VGPGAE"]
