"""
NicheCompass model (reference: models/nichecompass.py:24-1986): validates the AnnData object,
builds gene-program masks and feature index lists, owns the VGPGAE module, and exposes
``train`` / ``get_latent_representation`` / ``get_active_gps`` / ``save`` / ``load`` with the
reference's signatures and AnnData key schema.  The post-hoc analysis helpers of the reference
(differential GP tests, GP summaries, importances, plots) are out of scope of this build and
fail loudly.
"""
from typing import List, Literal, Optional, Tuple, Union

import numpy as np
import scipy.sparse as sp
import torch

from .. import ops
from ..data import initialize_dataloaders, prepare_data
from ..modules import VGPGAE
from ..train import Trainer
from .basemodelmixin import BaseModelMixin


def _complement(idx, n: int) -> List[int]:
    """[i for i in range(n) if i not in idx] without the reference's O(n * len(idx)) scan
    (models/nichecompass.py:389-402)."""
    keep = np.ones(n, dtype=bool)
    ii = np.asarray(idx, dtype=np.int64)
    keep[ii[(ii >= 0) & (ii < n)]] = False
    return np.nonzero(keep)[0].tolist()


def _sparse_bool_to_dense(m) -> torch.Tensor:
    return torch.from_numpy(np.asarray(sp.csr_matrix(m).astype(bool).todense()))


class NicheCompass(BaseModelMixin):
    def __init__(self,
                 adata,
                 adata_atac=None,
                 counts_key: Optional[str] = "counts",
                 adj_key: str = "spatial_connectivities",
                 gp_names_key: str = "nichecompass_gp_names",
                 active_gp_names_key: str = "nichecompass_active_gp_names",
                 gp_targets_mask_key: str = "nichecompass_gp_targets",
                 gp_targets_categories_mask_key: str = "nichecompass_gp_targets_categories",
                 targets_categories_label_encoder_key: str = "nichecompass_targets_categories_label_encoder",
                 gp_sources_mask_key: str = "nichecompass_gp_sources",
                 gp_sources_categories_mask_key: str = "nichecompass_gp_sources_categories",
                 sources_categories_label_encoder_key: str = "nichecompass_sources_categories_label_encoder",
                 ca_targets_mask_key: Optional[str] = "nichecompass_ca_targets",
                 ca_sources_mask_key: Optional[str] = "nichecompass_ca_sources",
                 latent_key: str = "nichecompass_latent",
                 cat_covariates_embeds_keys: Optional[List[str]] = None,
                 cat_covariates_embeds_injection: Optional[List[
                     Literal["encoder", "gene_expr_decoder", "chrom_access_decoder"]]] = [
                         "gene_expr_decoder", "chrom_access_decoder"],
                 cat_covariates_keys: Optional[List[str]] = None,
                 cat_covariates_no_edges: Optional[List[bool]] = None,
                 genes_idx_key: str = "nichecompass_genes_idx",
                 target_genes_idx_key: str = "nichecompass_target_genes_idx",
                 source_genes_idx_key: str = "nichecompass_source_genes_idx",
                 peaks_idx_key: str = "nichecompass_peaks_idx",
                 target_peaks_idx_key: str = "nichecompass_target_peaks_idx",
                 source_peaks_idx_key: str = "nichecompass_source_peaks_idx",
                 gene_peaks_mask_key: str = "nichecompass_gene_peaks",
                 recon_adj_key: Optional[str] = "nichecompass_recon_connectivities",
                 agg_weights_key: Optional[str] = "nichecompass_agg_weights",
                 include_edge_recon_loss: bool = True,
                 include_gene_expr_recon_loss: bool = True,
                 include_chrom_access_recon_loss: Optional[bool] = True,
                 include_cat_covariates_contrastive_loss: bool = False,
                 gene_expr_recon_dist: Literal["nb"] = "nb",
                 log_variational: bool = True,
                 node_label_method: Literal["one-hop-sum", "one-hop-norm", "one-hop-attention"] = "one-hop-norm",
                 active_gp_thresh_ratio: float = 0.01,
                 active_gp_type: Literal["mixed", "separate"] = "separate",
                 n_fc_layers_encoder: int = 1,
                 n_layers_encoder: int = 1,
                 n_hidden_encoder: Optional[int] = None,
                 conv_layer_encoder: Literal["gcnconv", "gatv2conv"] = "gatv2conv",
                 encoder_n_attention_heads: Optional[int] = 4,
                 encoder_use_bn: bool = False,
                 dropout_rate_encoder: float = 0.,
                 dropout_rate_graph_decoder: float = 0.,
                 cat_covariates_cats: Optional[List[List]] = None,
                 n_addon_gp: int = 100,
                 cat_covariates_embeds_nums: Optional[List[int]] = None,
                 include_edge_kl_loss: bool = True,
                 use_cuda_if_available: bool = True,
                 seed: int = 0,
                 **kwargs):
        init_locals = dict(locals())
        self.adata = adata
        self.adata_atac = adata_atac
        self.counts_key_ = counts_key
        self.adj_key_ = adj_key
        self.gp_names_key_ = gp_names_key
        self.active_gp_names_key_ = active_gp_names_key
        self.gp_targets_mask_key_ = gp_targets_mask_key
        self.gp_targets_categories_mask_key_ = gp_targets_categories_mask_key
        self.targets_categories_label_encoder_key_ = targets_categories_label_encoder_key
        self.gp_sources_mask_key_ = gp_sources_mask_key
        self.gp_sources_categories_mask_key_ = gp_sources_categories_mask_key
        self.sources_categories_label_encoder_key_ = sources_categories_label_encoder_key
        self.ca_targets_mask_key_ = ca_targets_mask_key
        self.ca_sources_mask_key_ = ca_sources_mask_key
        self.latent_key_ = latent_key
        self.cat_covariates_embeds_keys_ = cat_covariates_embeds_keys
        self.cat_covariates_embeds_injection_ = cat_covariates_embeds_injection
        self.cat_covariates_keys_ = cat_covariates_keys
        self.genes_idx_key_ = genes_idx_key
        self.target_genes_idx_key_ = target_genes_idx_key
        self.source_genes_idx_key_ = source_genes_idx_key
        self.peaks_idx_key_ = peaks_idx_key
        self.target_peaks_idx_key_ = target_peaks_idx_key
        self.source_peaks_idx_key_ = source_peaks_idx_key
        self.gene_peaks_mask_key_ = gene_peaks_mask_key
        self.recon_adj_key_ = recon_adj_key
        self.agg_weights_key_ = agg_weights_key
        self.include_edge_recon_loss_ = include_edge_recon_loss
        self.include_gene_expr_recon_loss_ = include_gene_expr_recon_loss
        self.include_chrom_access_recon_loss_ = include_chrom_access_recon_loss
        self.include_cat_covariates_contrastive_loss_ = include_cat_covariates_contrastive_loss
        self.gene_expr_recon_dist_ = gene_expr_recon_dist
        self.log_variational_ = log_variational
        self.node_label_method_ = node_label_method
        self.active_gp_thresh_ratio_ = active_gp_thresh_ratio
        self.active_gp_type_ = active_gp_type
        self.include_edge_kl_loss_ = include_edge_kl_loss
        self.seed_ = seed

        np.random.seed(self.seed_)
        if use_cuda_if_available and torch.cuda.is_available():
            torch.cuda.manual_seed(self.seed_)
        torch.manual_seed(self.seed_)

        # ---- gene-program masks [n_prior_gp, n_genes]
        for key, what in ((gp_targets_mask_key, "targets"), (gp_sources_mask_key, "sources")):
            if key not in adata.varm:
                raise ValueError(f"Please specify an adequate ´gp_{what}_mask_key´ for your adata object. The "
                                 f"{what} mask needs to be stored in ´adata.varm[gp_{what}_mask_key]´. If you do "
                                 "not want to mask gene expression reconstruction, you can create a mask of 1s "
                                 "that allows all gene program latent nodes to reconstruct all genes.")
        self.gp_targets_mask_ = torch.tensor(np.asarray(adata.varm[gp_targets_mask_key]).T, dtype=torch.bool)
        self.gp_sources_mask_ = torch.tensor(np.asarray(adata.varm[gp_sources_mask_key]).T, dtype=torch.bool)

        col_sums = torch.tensor(np.asarray(adata.X.sum(0)).reshape(-1))
        self.features_scale_factors_ = torch.concat((col_sums, col_sums))

        # ---- chromatin-accessibility masks
        if adata_atac is None:
            self.ca_targets_mask_ = None
            self.ca_sources_mask_ = None
            gene_peaks_mask = None
        else:
            gene_peaks_mask = _sparse_bool_to_dense(adata.varm[gene_peaks_mask_key])
            for key, what in ((ca_targets_mask_key, "targets"), (ca_sources_mask_key, "sources")):
                if key not in adata_atac.varm:
                    raise ValueError(f"Please specify an adequate ´ca_{what}_mask_key´ for your adata_atac object."
                                     f" The {what} mask needs to be stored in ´adata_atac.varm[ca_{what}_mask_key]´.")
            self.ca_targets_mask_ = _sparse_bool_to_dense(adata_atac.varm[ca_targets_mask_key].T)
            self.ca_sources_mask_ = _sparse_bool_to_dense(adata_atac.varm[ca_sources_mask_key].T)

        # ---- feature index lists
        n_genes = len(adata.var_names)
        fd = {}
        fd["masked_rna_idx"] = adata.uns[genes_idx_key]
        fd["unmasked_rna_idx"] = _complement(fd["masked_rna_idx"], n_genes)
        fd["target_masked_rna_idx"] = list(adata.uns[target_genes_idx_key])
        fd["target_unmasked_rna_idx"] = _complement(fd["target_masked_rna_idx"], n_genes)
        fd["source_masked_rna_idx"] = list(adata.uns[source_genes_idx_key])
        fd["source_unmasked_rna_idx"] = _complement(fd["source_masked_rna_idx"], n_genes)
        if adata_atac is not None:
            n_peaks = len(adata_atac.var_names)
            self.peaks_idx_ = adata_atac.uns[peaks_idx_key]
            self.target_peaks_idx_ = adata_atac.uns[target_peaks_idx_key]
            self.source_peaks_idx_ = adata_atac.uns[source_peaks_idx_key]
            fd["masked_atac_idx"] = adata_atac.uns[peaks_idx_key]
            fd["unmasked_atac_idx"] = _complement(fd["masked_atac_idx"], n_peaks)
            fd["target_masked_atac_idx"] = list(adata_atac.uns[target_peaks_idx_key])
            fd["target_unmasked_atac_idx"] = _complement(fd["target_masked_atac_idx"], n_peaks)
            fd["source_masked_atac_idx"] = list(adata_atac.uns[source_peaks_idx_key])
            fd["source_unmasked_atac_idx"] = _complement(fd["source_masked_atac_idx"], n_peaks)
        self.features_idx_dict_ = fd

        # ---- module sizes
        self.n_input_ = adata.n_vars
        self.n_output_genes_ = adata.n_vars
        if adata_atac is not None:
            self.modalities_ = ["rna", "atac"]
            if not np.all(adata.obs.index == adata_atac.obs.index):
                raise ValueError("Please make sure that 'adata' and 'adata_atac' contain the same observations "
                                 "in the same order.")
            self.n_input_ += adata_atac.n_vars
            self.n_output_peaks_ = adata_atac.n_vars
        else:
            self.modalities_ = ["rna"]
            self.n_output_peaks_ = 0
        self.n_fc_layers_encoder_ = n_fc_layers_encoder
        self.n_layers_encoder_ = n_layers_encoder
        self.conv_layer_encoder_ = conv_layer_encoder
        self.encoder_n_attention_heads_ = encoder_n_attention_heads if conv_layer_encoder == "gatv2conv" else 0
        self.encoder_use_bn_ = encoder_use_bn
        self.dropout_rate_encoder_ = dropout_rate_encoder
        self.dropout_rate_graph_decoder_ = dropout_rate_graph_decoder
        self.n_prior_gp_ = len(self.gp_targets_mask_)
        self.n_addon_gp_ = n_addon_gp

        # ---- validation of keys (before the adata object is touched)
        if counts_key is not None and counts_key not in adata.layers:
            raise ValueError("Please specify an adequate ´counts_key´. By default the counts are assumed to be "
                             "stored in data.layers['counts'].")
        if include_gene_expr_recon_loss and log_variational:
            x = adata.X if counts_key is None else adata.layers[counts_key]
            if (x < 0).sum() > 0:
                raise ValueError("Please make sure that ´adata.layers[counts_key]´ contains the raw counts (not log"
                                 " library size normalized) if ´include_gene_expr_recon_loss´ is ´True´ and "
                                 "´log_variational´ is ´True´. If you want to use log library size normalized "
                                 "counts, make sure that ´log_variational´ is ´False´.")
        if adj_key not in adata.obsp:
            raise ValueError("Please specify an adequate ´adj_key´. By default the adjacency matrix is assumed to "
                             "be stored in adata.obsm['spatial_connectivities'].")
        if gp_names_key not in adata.uns:
            raise ValueError("Please specify an adequate ´gp_names_key´. By default the gene program names are "
                             "assumed to be stored in adata.uns['nichecompass_gp_names'].")
        if cat_covariates_keys is not None:
            for key in cat_covariates_keys:
                if key not in adata.obs:
                    raise ValueError("Please specify adequate ´cat_covariates_keys´. "
                                     f"The key {key} was not found in adata.")

        # ---- add-on gene-program names
        if n_addon_gp > 0:
            gp_list = list(self.adata.uns[self.gp_names_key_])
            present = set(gp_list)
            gp_list += [f"Add-on_{i}_GP" for i in range(n_addon_gp) if f"Add-on_{i}_GP" not in present]
            self.adata.uns[self.gp_names_key_] = np.array(gp_list)
        else:
            names = np.asarray(self.adata.uns[gp_names_key])
            self.adata.uns[gp_names_key] = names[np.array(["Add-on" not in str(g) for g in names], dtype=bool)]

        # ---- categorical covariates
        if cat_covariates_cats is None:
            self.cat_covariates_cats_ = ([adata.obs[k].unique().tolist() for k in cat_covariates_keys]
                                         if cat_covariates_keys is not None else [])
        else:
            self.cat_covariates_cats_ = cat_covariates_cats
        if cat_covariates_embeds_nums is None:
            cat_covariates_embeds_nums = [len(c) for c in self.cat_covariates_cats_]
        self.cat_covariates_embeds_nums_ = cat_covariates_embeds_nums
        if n_hidden_encoder is None:
            n_hidden_encoder = min(len(adata.var), self.n_prior_gp_ + self.n_addon_gp_)
        self.n_hidden_encoder_ = n_hidden_encoder
        if cat_covariates_no_edges is None and len(self.cat_covariates_cats_) > 0:
            self.cat_covariates_no_edges_ = [True] * len(self.cat_covariates_cats_)
        else:
            self.cat_covariates_no_edges_ = cat_covariates_no_edges

        self.model = VGPGAE(
            n_input=self.n_input_,
            n_fc_layers_encoder=self.n_fc_layers_encoder_,
            n_layers_encoder=self.n_layers_encoder_,
            n_hidden_encoder=self.n_hidden_encoder_,
            n_prior_gp=self.n_prior_gp_,
            n_addon_gp=self.n_addon_gp_,
            cat_covariates_embeds_nums=self.cat_covariates_embeds_nums_,
            n_output_genes=self.n_output_genes_,
            n_output_peaks=self.n_output_peaks_,
            target_rna_decoder_mask=self.gp_targets_mask_,
            source_rna_decoder_mask=self.gp_sources_mask_,
            target_atac_decoder_mask=self.ca_targets_mask_,
            source_atac_decoder_mask=self.ca_sources_mask_,
            features_idx_dict=self.features_idx_dict_,
            features_scale_factors=self.features_scale_factors_,
            gene_peaks_mask=gene_peaks_mask,
            cat_covariates_cats=self.cat_covariates_cats_,
            cat_covariates_no_edges=self.cat_covariates_no_edges_ or [],
            conv_layer_encoder=self.conv_layer_encoder_,
            encoder_n_attention_heads=self.encoder_n_attention_heads_,
            encoder_use_bn=self.encoder_use_bn_,
            dropout_rate_encoder=self.dropout_rate_encoder_,
            dropout_rate_graph_decoder=self.dropout_rate_graph_decoder_,
            include_edge_recon_loss=self.include_edge_recon_loss_,
            include_gene_expr_recon_loss=self.include_gene_expr_recon_loss_,
            include_chrom_access_recon_loss=self.include_chrom_access_recon_loss_,
            include_cat_covariates_contrastive_loss=self.include_cat_covariates_contrastive_loss_,
            rna_recon_loss=self.gene_expr_recon_dist_,
            node_label_method=self.node_label_method_,
            active_gp_thresh_ratio=self.active_gp_thresh_ratio_,
            active_gp_type=self.active_gp_type_,
            log_variational=self.log_variational_,
            cat_covariates_embeds_injection=self.cat_covariates_embeds_injection_,
            include_edge_kl_loss=self.include_edge_kl_loss_)

        self.is_trained_ = False
        self.init_params_ = self._get_init_params(init_locals)

    # ------------------------------------------------------------------ training
    def train(self,
              n_epochs: int = 100,
              n_epochs_all_gps: int = 25,
              n_epochs_no_edge_recon: int = 0,
              n_epochs_no_cat_covariates_contrastive: int = 5,
              lr: float = 0.001,
              weight_decay: float = 0.,
              lambda_edge_recon: Optional[float] = 500000.,
              lambda_gene_expr_recon: float = 300.,
              lambda_chrom_access_recon: float = 100.,
              lambda_cat_covariates_contrastive: float = 0.,
              contrastive_logits_pos_ratio: float = 0.,
              contrastive_logits_neg_ratio: float = 0.,
              lambda_group_lasso: float = 0.,
              lambda_l1_masked: float = 0.,
              l1_targets_categories: Optional[list] = ["target_gene"],
              l1_sources_categories: Optional[list] = None,
              lambda_l1_addon: float = 30.,
              edge_val_ratio: float = 0.1,
              node_val_ratio: float = 0.1,
              edge_batch_size: int = 256,
              node_batch_size: Optional[int] = None,
              mlflow_experiment_id: Optional[str] = None,
              retrieve_cat_covariates_embeds: bool = False,
              retrieve_recon_edge_probs: bool = False,
              retrieve_agg_weights: bool = False,
              use_cuda_if_available: bool = True,
              n_sampled_neighbors: int = -1,
              latent_dtype: type = np.float64,
              **trainer_kwargs):
        """Train the model and write the latent representation and the active gene-program names
        into ``adata`` (reference: models/nichecompass.py:590-828).  Extra ``trainer_kwargs`` of
        this build: ``compute_dtype`` (torch.float32 parity mode / torch.bfloat16 tensor-core
        mode), ``store_location`` ("device" | "host"), ``max_iters_per_epoch``."""
        if retrieve_recon_edge_probs or retrieve_agg_weights:
            raise NotImplementedError("retrieve_recon_edge_probs / retrieve_agg_weights call analysis helpers "
                                      "that are out of scope of this build (and broken in the reference)")
        max_iters = trainer_kwargs.pop("max_iters_per_epoch", None)
        self.trainer = Trainer(
            adata=self.adata,
            adata_atac=self.adata_atac,
            model=self.model,
            counts_key=self.counts_key_,
            adj_key=self.adj_key_,
            gp_targets_mask_key=self.gp_targets_mask_key_,
            gp_sources_mask_key=self.gp_sources_mask_key_,
            cat_covariates_keys=self.cat_covariates_keys_,
            edge_val_ratio=edge_val_ratio,
            node_val_ratio=node_val_ratio,
            edge_batch_size=edge_batch_size,
            node_batch_size=node_batch_size,
            use_cuda_if_available=use_cuda_if_available,
            n_sampled_neighbors=n_sampled_neighbors,
            latent_dtype=latent_dtype,
            **trainer_kwargs)

        if lambda_l1_masked > 0.:
            def encoded(categories, enc_key):
                enc = self.adata.uns[enc_key]
                if categories is None:
                    return list(enc.values())
                return [enc[c] for c in categories if c in enc]
            l1_targets_mask = torch.from_numpy(np.isin(
                self.adata.varm[self.gp_targets_categories_mask_key_],
                encoded(l1_targets_categories, self.targets_categories_label_encoder_key_)))
            l1_sources_mask = torch.from_numpy(np.isin(
                self.adata.varm[self.gp_sources_categories_mask_key_],
                encoded(l1_sources_categories, self.sources_categories_label_encoder_key_)))
        else:
            l1_targets_mask = None
            l1_sources_mask = None

        self.trainer.train(
            n_epochs=n_epochs,
            n_epochs_no_edge_recon=n_epochs_no_edge_recon,
            n_epochs_no_cat_covariates_contrastive=n_epochs_no_cat_covariates_contrastive,
            n_epochs_all_gps=n_epochs_all_gps,
            lr=lr,
            weight_decay=weight_decay,
            lambda_edge_recon=lambda_edge_recon,
            lambda_gene_expr_recon=lambda_gene_expr_recon,
            lambda_chrom_access_recon=lambda_chrom_access_recon,
            lambda_cat_covariates_contrastive=lambda_cat_covariates_contrastive,
            contrastive_logits_pos_ratio=contrastive_logits_pos_ratio,
            contrastive_logits_neg_ratio=contrastive_logits_neg_ratio,
            lambda_group_lasso=lambda_group_lasso,
            lambda_l1_masked=lambda_l1_masked,
            l1_targets_mask=l1_targets_mask,
            l1_sources_mask=l1_sources_mask,
            lambda_l1_addon=lambda_l1_addon,
            mlflow_experiment_id=mlflow_experiment_id,
            max_iters_per_epoch=max_iters)

        self.node_batch_size_ = self.trainer.node_batch_size_
        self.is_trained_ = True
        self.model.eval()

        self.adata.obsm[self.latent_key_], _ = self.get_latent_representation(
            adata=self.adata,
            counts_key=self.counts_key_,
            adj_key=self.adj_key_,
            cat_covariates_keys=self.cat_covariates_keys_,
            only_active_gps=True,
            return_mu_std=True,
            node_batch_size=self.node_batch_size_,
            dtype=latent_dtype)
        self.adata.uns[self.active_gp_names_key_] = self.get_active_gps()

        if len(self.cat_covariates_cats_) > 0 and retrieve_cat_covariates_embeds:
            for key, embed in zip(self.cat_covariates_embeds_keys_, self.get_cat_covariates_embeds()):
                self.adata.uns[key] = embed

    # ------------------------------------------------------------------ retrieval
    def get_gp_data(self, selected_gps: Optional[Union[str, list]] = None
                    ) -> Tuple[np.ndarray, np.ndarray, Optional[np.ndarray]]:
        """(indices, rna decoder weights [2 * n_genes, n_sel], atac decoder weights | None) of the
        selected gene programs (models/nichecompass.py:1166-1216)."""
        self._check_if_trained(warn=True)
        all_gps = list(self.adata.uns[self.gp_names_key_])
        if selected_gps is None:
            selected_gps = all_gps
        elif isinstance(selected_gps, str):
            selected_gps = [selected_gps]
        pos = {g: i for i, g in enumerate(all_gps)}
        idx = np.array([pos[g] for g in selected_gps])
        weights = self.model.get_gp_weights()
        rna = weights[0][:, idx].cpu().detach().numpy()
        atac = weights[1][:, idx].cpu().detach().numpy() if "atac" in self.modalities_ else None
        return idx, rna, atac

    def get_cat_covariates_embeds(self) -> List[np.ndarray]:
        self._check_if_trained(warn=True)
        return [e.weight.cpu().detach().numpy() for e in self.model.cat_covariates_embedders]

    def get_active_gps(self) -> np.ndarray:
        """Names of the active gene programs (models/nichecompass.py:1235-1262)."""
        self._check_if_trained(warn=True)
        mask = self.model.get_active_gp_mask().detach().cpu().numpy()
        return np.asarray(self.adata.uns[self.gp_names_key_])[mask]

    @torch.no_grad()
    def get_latent_representation(self,
                                  adata=None,
                                  adata_atac=None,
                                  counts_key: Optional[str] = "counts",
                                  adj_key: str = "spatial_connectivities",
                                  cat_covariates_keys: Optional[List[str]] = None,
                                  only_active_gps: bool = True,
                                  return_mu_std: bool = False,
                                  node_batch_size: int = 64,
                                  dtype: type = np.float64,
                                  layerwise: Optional[bool] = None,
                                  partition: str = "auto",
                                  ) -> Union[np.ndarray, Tuple[np.ndarray, np.ndarray]]:
        """Latent representation (gene-program scores) of every observation
        (models/nichecompass.py:1264-1381).

        ``layerwise`` / ``partition`` are extensions of this build: with the 1-layer GATv2 encoder the
        default is the layer-wise full-graph path of ``models/tiled_inference.py`` (every node's
        ``fc_l1`` / ``lin_l`` rows computed once; with ``torch.distributed`` initialised the nodes are
        split into spatial tiles, one per rank, and boundary ``hidden`` rows are exchanged over NCCL;
        every rank returns the full array).  ``layerwise=False`` forces the loader walk below.

        The reference walks a ``NeighborLoader`` over consecutive blocks of ``node_batch_size``
        nodes.  With the GATv2 encoder and all neighbours sampled, a node's output does not depend
        on which other seeds share its batch, so the blocks are widened to tiles of >= 8192 nodes
        (same values, ~100x fewer launches); with the GCN encoder the sub-graph-local degree
        normalisation makes outputs depend on the batch composition, so the
        reference block size is kept.  Results stay on the device until one final copy."""
        self._check_if_trained(warn=False)
        device = next(self.model.parameters()).device
        if adata is None:
            adata = self.adata
        if adata_atac is None and hasattr(self, "adata_atac"):
            adata_atac = self.adata_atac
        data_dict = prepare_data(
            adata=adata,
            cat_covariates_label_encoders=self.model.cat_covariates_label_encoders_,
            adata_atac=adata_atac,
            counts_key=counts_key,
            adj_key=adj_key,
            cat_covariates_keys=cat_covariates_keys,
            edge_val_ratio=0.,
            edge_test_ratio=0.,
            node_val_ratio=0.,
            node_test_ratio=0.,
            device=device)
        from . import tiled_inference as TI
        if layerwise is None:
            layerwise = TI.supports(self.model)
        elif layerwise and not TI.supports(self.model):
            raise ValueError("layerwise inference needs the 1-layer GATv2 encoder (outputs of the GCN encoder "
                             "depend on the sampled batch: SURVEY.md App. B-6)")
        if layerwise:
            return self._latent_layerwise(adata, data_dict["node_masked_data"], only_active_gps, return_mu_std,
                                          dtype, partition, adj_key)
        tile = node_batch_size
        if self.model.conv_layer_encoder_ == "gatv2conv" and self.model.n_layers_encoder_ == 1:
            tile = max(node_batch_size, 8192)
        loader = initialize_dataloaders(node_masked_data=data_dict["node_masked_data"], edge_train_data=None,
                                        edge_val_data=None, edge_batch_size=None, node_batch_size=tile,
                                        shuffle=False)["node_train_loader"]
        was_training = self.model.training
        self.model.eval()
        n_gps = (int(self.model.get_active_gp_mask().sum().item()) if only_active_gps
                 else self.n_prior_gp_ + self.n_addon_gp_)
        n_obs = adata.shape[0]
        tdt = {np.float16: torch.float16, np.float32: torch.float32, np.float64: torch.float64}.get(dtype)
        if tdt is None:
            tdt = torch.from_numpy(np.empty(0, dtype=dtype)).dtype
        outs = [torch.empty(n_obs, n_gps, dtype=tdt, device=device) for _ in range(2 if return_mu_std else 1)]
        lo = 0
        for node_batch in loader:
            hi = lo + node_batch.batch_size
            res = self.model.get_latent_representation(node_batch=node_batch, only_active_gps=only_active_gps,
                                                       return_mu_std=return_mu_std)
            if return_mu_std:
                outs[0][lo:hi] = res[0]
                outs[1][lo:hi] = res[1]
            else:
                outs[0][lo:hi] = res
            lo = hi
        self.model.train(was_training)
        host = [o.cpu().numpy() for o in outs]
        return (host[0], host[1]) if return_mu_std else host[0]

    def _latent_layerwise(self, adata, data, only_active_gps, return_mu_std, dtype, partition, adj_key):
        """Layer-wise full-graph inference over spatial tiles (models/tiled_inference.py)."""
        import torch.distributed as dist
        from . import tiled_inference as TI
        device = next(self.model.parameters()).device
        dist_on = dist.is_available() and dist.is_initialized()
        rank, world = (dist.get_rank(), dist.get_world_size()) if dist_on else (0, 1)
        n_obs = adata.shape[0]
        if partition == "auto" and world == 1:
            partition = "none"
        obsm = getattr(adata, "obsm", None)
        coords = obsm["spatial"] if (obsm is not None and "spatial" in obsm) else None
        perm_np = TI.partition_order(n_obs, coords, adata.obsp[adj_key], partition)
        bounds = TI.tile_bounds(n_obs, world)
        perm = torch.from_numpy(perm_np).to(device)
        inv = torch.empty_like(perm)
        inv[perm] = torch.arange(n_obs, device=device)
        g = data.device_graph()
        lo, hi = bounds[rank], bounds[rank + 1]
        lg = TI.build_local_graph(g.rowptr, g.col, perm, inv, lo, hi)
        was_training = self.model.training
        self.model.eval()
        ml = TI.encode_partition(self.model, data.store, lg, perm[lo:hi], bounds, data.cats_device(),
                                 want_logstd=return_mu_std)
        P = self.n_prior_gp_ + self.n_addon_gp_
        parts = [ml[:, :P]] + ([torch.exp(ml[:, P:])] if return_mu_std else [])
        if only_active_gps:
            a = self.model.get_active_gp_mask()
            parts = [p_[:, a] for p_ in parts]
        self.model.train(was_training)
        tdt = {np.float16: torch.float16, np.float32: torch.float32, np.float64: torch.float64}.get(dtype)
        if tdt is None:
            tdt = torch.from_numpy(np.empty(0, dtype=dtype)).dtype
        self.last_inference_stats_ = {"n_own": lg.n_own, "n_halo": lg.n_halo, "rank": rank, "world": world}
        host = []
        for part in parts:
            part = part.to(tdt).contiguous()
            full = torch.empty(n_obs, part.size(1), dtype=tdt, device=device)
            if world == 1:
                full[perm] = part
            else:
                n_max = max(bounds[r + 1] - bounds[r] for r in range(world))
                mine = torch.zeros(n_max, part.size(1), dtype=tdt, device=device)
                mine[:part.size(0)] = part
                everyone = torch.empty(world * n_max, part.size(1), dtype=tdt, device=device)
                dist.all_gather_into_tensor(everyone, mine)
                for r in range(world):
                    full[perm[bounds[r]:bounds[r + 1]]] = everyone[r * n_max:r * n_max + bounds[r + 1] - bounds[r]]
            host.append(full.cpu().numpy())
        return (host[0], host[1]) if return_mu_std else host[0]

    # ------------------------------------------------------------------ out of scope
    def _out_of_scope(self, name: str):
        raise NotImplementedError(f"NicheCompass.{name} is a post-hoc analysis helper outside the hot path this "
                                  "build replaces (SURVEY.md section 2, row 5); use the reference package for it")

    def run_differential_gp_tests(self, *a, **k):
        self._out_of_scope("run_differential_gp_tests")

    def compute_gp_gene_importances(self, *a, **k):
        self._out_of_scope("compute_gp_gene_importances")

    def compute_gp_peak_importances(self, *a, **k):
        self._out_of_scope("compute_gp_peak_importances")

    # ------------------------------------------------------------------ post-hoc retrieval on the hot-path kernels
    def _full_node_loader(self, adata, adata_atac, node_batch_size: int):
        device = next(self.model.parameters()).device
        data_dict = prepare_data(
            adata=adata, cat_covariates_label_encoders=self.model.cat_covariates_label_encoders_,
            adata_atac=adata_atac, counts_key=self.counts_key_, adj_key=self.adj_key_,
            cat_covariates_keys=self.cat_covariates_keys_, edge_val_ratio=0., edge_test_ratio=0.,
            node_val_ratio=0., node_test_ratio=0., device=device)
        return initialize_dataloaders(node_masked_data=data_dict["node_masked_data"], edge_train_data=None,
                                      edge_val_data=None, edge_batch_size=None, node_batch_size=node_batch_size,
                                      shuffle=False)["node_train_loader"]

    @torch.no_grad()
    def get_omics_decoder_outputs(self, adata=None, adata_atac=None, only_active_gps: bool = True,
                                  node_batch_size: int = 64) -> dict:
        """Negative-binomial means of the target / source decoders for every observation over all output
        features (models/nichecompass.py:1383-1470): ``{target,source}_{rna,atac}_nb_means`` as float64 arrays
        [n_obs, n_features].  The walk over consecutive blocks of ``node_batch_size`` nodes is the reference's
        (with the default one-hop-norm labels the source library size depends on the sampled block, SURVEY.md
        App. B-6); every block runs encoder GEMMs + GATv2 kernels + decoder GEMM + ``nc_nb_loss_fused``."""
        self._check_if_trained(warn=False)
        if adata is None:
            adata = self.adata
        if adata_atac is None and hasattr(self, "adata_atac"):
            adata_atac = self.adata_atac
        loader = self._full_node_loader(adata, adata_atac, node_batch_size)
        device = next(self.model.parameters()).device
        n_obs = adata.shape[0]
        keys = [f"{ent}_{mod}_nb_means" for mod in self.model.modalities_ for ent in ("target", "source")]
        width = {"rna": self.n_output_genes_, "atac": getattr(self, "n_output_peaks_", 0)}
        outs = {k: torch.empty(n_obs, width[k.split("_")[1]], dtype=torch.float32, device=device) for k in keys}
        was_training = self.model.training
        self.model.eval()
        lo = 0
        for node_batch in loader:
            hi = lo + int(node_batch.batch_size)
            res = self.model.get_omics_decoder_outputs(node_batch=node_batch, only_active_gps=only_active_gps)
            for k in keys:
                outs[k][lo:hi] = res[k]
            lo = hi
        self.model.train(was_training)
        return {k: v.cpu().numpy().astype(np.float64) for k, v in outs.items()}

    @torch.no_grad()
    def get_neighbor_importances(self, node_batch_size: Optional[int] = None) -> sp.csr_matrix:
        """Aggregation weights of the node-label aggregator as a sparse [n_obs, n_obs] matrix: entry (i, j) is the
        weight observation i gives to its neighbour j (models/nichecompass.py:1600-1686; mean over the heads for
        ``one-hop-attention``).  The reference reads ``self.model.node_label_aggregator`` (the attribute is
        ``rna_node_label_aggregator``) and pairs SparseTensor-ordered weights with loader-ordered edges; this is
        the documented intent: one weight per (target, neighbour) pair of the sampled block, the self pair
        carrying the fill_diag weight.  ``one-hop-sum`` has unit weights."""
        self._check_if_trained(warn=False)
        if node_batch_size is None:
            node_batch_size = self.node_batch_size_
        loader = self._full_node_loader(self.adata, getattr(self, "adata_atac", None), node_batch_size)
        method = self.model.node_label_method_
        agg = self.model.rna_node_label_aggregator
        gi, gj, gv = [], [], []
        for b in loader:
            nb, g = int(b.batch_size), b.graph
            n_id = b.n_id.long()
            rp = g.rowptr[:nb + 1].long()
            E = int(rp[-1])
            if E == 0:
                continue
            rows = torch.repeat_interleave(torch.arange(nb, device=rp.device), rp[1:] - rp[:-1])
            cols = g.col[:E].long()
            if method == "one-hop-norm":
                w, wself = g.gcn_weights(use_out_degree=True, n_rows=nb)
                val = torch.where(cols == rows, wself[rows], w[:E])
            elif method == "one-hop-sum":
                val = torch.ones(E, dtype=torch.float32, device=rp.device)
            else:
                x = ops.csr_gather_rows(b.store, b.n_id, g.n_nodes, False, torch.float32)[:, :agg.n_input]
                _, alpha = agg(x, g, return_agg_weights=True, n_rows=nb)
                val = alpha[:E].mean(dim=-1)
            gi.append(n_id[rows]); gj.append(n_id[cols]); gv.append(val)
        n = self.adata.shape[0]
        if not gi:
            return sp.csr_matrix((n, n))
        gi, gj, gv = (torch.cat(t).cpu().numpy() for t in (gi, gj, gv))
        return sp.csr_matrix((gv.astype(np.float64), (gi, gj)), shape=(n, n))

    @torch.no_grad()
    def get_recon_edge_probs(self, node_batch_size: int = 2048, device: Optional[str] = None,
                             edge_thresh: Optional[float] = None, n_neighbors: Optional[int] = None,
                             return_edge_probs: bool = False, compute_dtype: torch.dtype = torch.float32
                             ) -> Union[sp.csr_matrix, torch.Tensor]:
        """Reconstructed adjacency (or the dense edge-probability matrix) from the stored latent representation
        (models/nichecompass.py:1473-1597): p_ij = sigmoid(cosine(z_i, z_j)) with the inactive gene programs put
        back as zeros; an edge wherever p_ij >= ``edge_thresh``, or >= the row's k-th largest probability with
        k = ``n_neighbors`` (result symmetrised) / the row's degree in the input graph (not symmetrised).
        (The reference passes ``reduced_obs_start_idx`` to a graph decoder whose ``forward`` no longer takes it,
        nn/decoders.py:39-62, so the call fails as shipped; this is the behaviour its docstring describes.)
        One row block = one GEMM of normalised rows, ``nc_gemm_f32`` or (``compute_dtype=torch.bfloat16``) the
        tcgen05 kernel; thresholds by a device sort; only the edges found leave the device."""
        self._check_if_trained(warn=False)
        if self.latent_key_ not in self.adata.obsm:
            raise ValueError(f"Please first store the latent representations in adata.obsm['{self.latent_key_}']. "
                             "They can be retrieved via 'model.get_latent_representation()'.")
        dev = next(self.model.parameters()).device if device is None else torch.device(device)
        if dev.type != "cuda":
            raise RuntimeError("get_recon_edge_probs runs on the CUDA kernels (nichecompass_b200 has no CPU path)")
        z = torch.as_tensor(np.asarray(self.adata.obsm[self.latent_key_]), dtype=torch.float32, device=dev)
        active = self.model.get_active_gp_mask().to(dev)
        zf = torch.zeros(z.size(0), active.numel(), dtype=torch.float32, device=dev)
        zf[:, active] = z
        # nn/utils.py:8-36: rows scaled by 1 / max(|z|, eps); padded to a TMA-legal pitch
        zn = zf / zf.norm(dim=1, keepdim=True).clamp_min(1e-8)
        P = zn.size(1)
        zp = torch.zeros(zn.size(0), ops.pad8(P), dtype=compute_dtype, device=dev)
        zp[:, :P] = zn
        n = zp.size(0)

        def block_probs(lo, hi):
            logits = ops.gemm(zp[lo:hi], zp, out_dtype=torch.float32, M=hi - lo, N=n, K=zp.size(1))
            return torch.sigmoid(logits)
        if return_edge_probs:
            return torch.cat([block_probs(lo, min(lo + node_batch_size, n)) for lo in range(0, n, node_batch_size)])
        deg = None
        if edge_thresh is None and n_neighbors is None:
            deg = torch.as_tensor(np.asarray(self.adata.obsp[self.adj_key_].sum(axis=1)).flatten().astype(np.int64),
                                  device=dev)
        ri, ci = [], []
        for lo in range(0, n, node_batch_size):
            hi = min(lo + node_batch_size, n)
            probs = block_probs(lo, hi)
            if edge_thresh is not None:
                keep = probs >= edge_thresh
            else:
                k = deg[lo:hi] if deg is not None else torch.full((hi - lo,), int(n_neighbors), device=dev)
                srt = probs.sort(dim=1, descending=True).values
                thr = srt.gather(1, (k - 1).clamp_(0, n - 1)[:, None])
                keep = probs >= thr
            nz = keep.nonzero()
            ri.append(nz[:, 0] + lo); ci.append(nz[:, 1])
        ri, ci = torch.cat(ri).cpu().numpy(), torch.cat(ci).cpu().numpy()
        adj = sp.csr_matrix((np.ones(ri.shape[0]), (ri, ci)), shape=(n, n))
        if n_neighbors is not None:
            adj = adj.maximum(adj.T)
        return adj

    def get_gp_summary(self, *a, **k):
        self._out_of_scope("get_gp_summary")

    def add_active_gp_scores_to_obs(self, *a, **k):
        self._out_of_scope("add_active_gp_scores_to_obs")
