"""
Variational Gene Program Graph Autoencoder module (reference: modules/vgpgae.py:32-1628).

Same constructor, ``forward`` / ``loss`` / ``get_active_gp_mask`` / ``get_latent_representation``
signatures, output-dict keys and state_dict keys as the reference; every tensor operation of
the hot path runs in the hand-written kernels of libnichecompass_b200.so.  Differences that are
deliberate and result-preserving:
  * encoder outputs are only computed for the rows that are consumed (the seeds of the batch);
  * the softmax -> NB-means -> column-select -> NB-loss chain is one fused kernel evaluated in
    ``forward`` (its value is stashed in the output dict and weighted in ``loss``);
  * the rna dynamic decoder masks are kept as one "alive" flag per gene program (the reference
    only ever zeroes whole rows of them, modules/vgpgae.py:655-656).
Accepted batches: PyG-style objects (``x`` dense, ``edge_index``, ``batch_size`` /
``edge_label_index`` + ``edge_label``) and ``nichecompass_b200.data.SampledBatch``
(CSR subgraph + ids into the HBM count store).
"""
from typing import List, Literal, Optional, Tuple, Union

import torch
import torch.nn as nn

from .. import ops
from ..functional import GatherRowsFn, ReparamFn, SplitLatentFn, SubgraphCSR
from .._lib import call
from ..nn import (CosineSimGraphDecoder, Encoder, MaskedOmicsFeatureDecoder,
                  OneHopAttentionNodeLabelAggregator, OneHopGCNNormNodeLabelAggregator,
                  OneHopSumNodeLabelAggregator)
from ..ops import pad8
from .basemodulemixin import BaseModuleMixin
from .losses import (compute_cat_covariates_contrastive_loss, compute_edge_recon_loss,
                     compute_gp_group_lasso_reg_loss, compute_gp_l1_reg_loss, compute_kl_reg_loss,
                     compute_omics_recon_nb_loss_from_logits)
from .vgaemodulemixin import VGAEModuleMixin


class VGPGAE(nn.Module, BaseModuleMixin, VGAEModuleMixin):
    def __init__(self,
                 n_input: int,
                 n_fc_layers_encoder: int,
                 n_layers_encoder: int,
                 n_hidden_encoder: int,
                 n_prior_gp: int,
                 n_addon_gp: int,
                 cat_covariates_embeds_nums: List[int],
                 n_output_genes: int,
                 target_rna_decoder_mask: torch.Tensor,
                 source_rna_decoder_mask: torch.Tensor,
                 features_idx_dict: dict,
                 features_scale_factors: torch.Tensor,
                 n_output_peaks: int = 0,
                 target_atac_decoder_mask: Optional[torch.Tensor] = None,
                 source_atac_decoder_mask: Optional[torch.Tensor] = None,
                 gene_peaks_mask: Optional[torch.Tensor] = None,
                 cat_covariates_cats: List[List] = [],
                 cat_covariates_no_edges: List[bool] = [],
                 conv_layer_encoder: Literal["gcnconv", "gatv2conv"] = "gcnconv",
                 encoder_n_attention_heads: int = 4,
                 encoder_use_bn: bool = False,
                 dropout_rate_encoder: float = 0.,
                 dropout_rate_graph_decoder: float = 0.,
                 include_edge_recon_loss: bool = True,
                 include_gene_expr_recon_loss: bool = True,
                 include_chrom_access_recon_loss: bool = True,
                 include_cat_covariates_contrastive_loss: bool = True,
                 rna_recon_loss: Literal["nb"] = "nb",
                 atac_recon_loss: Literal["nb"] = "nb",
                 node_label_method: Literal["one-hop-norm", "one-hop-sum", "one-hop-attention"] = "one-hop-norm",
                 active_gp_thresh_ratio: float = 0.03,
                 active_gp_type: Literal["mixed", "separate"] = "separate",
                 log_variational: bool = True,
                 cat_covariates_embeds_injection: Optional[List[str]] = ["gene_expr_decoder",
                                                                         "chrom_access_decoder"],
                 use_fc_decoder: bool = False,
                 fc_decoder_n_layers: int = 2,
                 include_edge_kl_loss: bool = True):
        super().__init__()
        print("--- INITIALIZING NEW NETWORK MODULE: VARIATIONAL GENE PROGRAM GRAPH AUTOENCODER ---")
        print(f"LOSS -> include_edge_recon_loss: {include_edge_recon_loss}, "
              f"include_gene_expr_recon_loss: {include_gene_expr_recon_loss}, "
              f"rna_recon_loss: {rna_recon_loss}", end="")
        if target_atac_decoder_mask is not None:
            print(f", include_chrom_access_recon_loss: {include_chrom_access_recon_loss}, "
                  f"atac_recon_loss: {atac_recon_loss}", end=" ")
        print(f"\nNODE LABEL METHOD -> {node_label_method}")
        print(f"ACTIVE GP THRESHOLD RATIO -> {active_gp_thresh_ratio}")
        print(f"LOG VARIATIONAL -> {log_variational}")
        if len(cat_covariates_cats) != 0:
            print(f"CATEGORICAL COVARIATES EMBEDDINGS INJECTION -> {cat_covariates_embeds_injection}")
        if use_fc_decoder:
            raise NotImplementedError("use_fc_decoder is an ablation of the reference and is out of scope")
        if rna_recon_loss != "nb" or atac_recon_loss != "nb":
            raise ValueError("only the negative-binomial reconstruction loss exists ('nb')")

        self.n_input_ = n_input
        self.n_fc_layers_encoder_ = n_fc_layers_encoder
        self.n_layers_encoder_ = n_layers_encoder
        self.n_hidden_encoder_ = n_hidden_encoder
        self.n_prior_gp_ = n_prior_gp
        self.n_addon_gp_ = n_addon_gp
        self.cat_covariates_embeds_nums_ = cat_covariates_embeds_nums
        self.n_output_genes_ = n_output_genes
        self.target_rna_decoder_mask = target_rna_decoder_mask
        self.source_rna_decoder_mask = source_rna_decoder_mask
        self.n_output_peaks_ = n_output_peaks
        self.target_atac_decoder_mask = target_atac_decoder_mask
        self.source_atac_decoder_mask = source_atac_decoder_mask
        self.features_idx_dict_ = features_idx_dict
        self.features_scale_factors_ = features_scale_factors
        self.gene_peaks_mask_ = gene_peaks_mask
        self.cat_covariates_cats_ = cat_covariates_cats
        self.n_cat_covariates_ = len(cat_covariates_cats)
        self.cat_covariates_no_edges_ = cat_covariates_no_edges
        self.nums_cat_covariates_cats_ = [len(c) for c in cat_covariates_cats]
        self.cat_covariates_label_encoders_ = [{k: v for k, v in zip(c, range(len(c)))}
                                               for c in cat_covariates_cats]
        self.conv_layer_encoder_ = conv_layer_encoder
        self.encoder_n_attention_heads_ = encoder_n_attention_heads
        self.encoder_use_bn_ = encoder_use_bn
        self.dropout_rate_encoder_ = dropout_rate_encoder
        self.dropout_rate_graph_decoder_ = dropout_rate_graph_decoder
        self.include_edge_recon_loss_ = include_edge_recon_loss
        self.include_gene_expr_recon_loss_ = include_gene_expr_recon_loss
        self.include_chrom_access_recon_loss_ = include_chrom_access_recon_loss
        self.include_cat_covariates_contrastive_loss_ = include_cat_covariates_contrastive_loss
        self.include_edge_kl_loss_ = include_edge_kl_loss
        self.rna_recon_loss_ = rna_recon_loss
        self.atac_recon_loss_ = atac_recon_loss
        self.node_label_method_ = node_label_method
        self.active_gp_thresh_ratio_ = active_gp_thresh_ratio
        self.active_gp_type_ = active_gp_type
        self.log_variational_ = log_variational
        self.cat_covariates_embeds_injection_ = cat_covariates_embeds_injection or []
        self.freeze_ = False
        self.modalities_ = ["rna"] + (["atac"] if target_atac_decoder_mask is not None else [])
        P = n_prior_gp + n_addon_gp

        # categorical covariate embedders (modules/vgpgae.py:259-271)
        if self.n_cat_covariates_ > 0:
            self.cat_covariates_embedders = []
            for i in range(self.n_cat_covariates_):
                emb = nn.Embedding(self.nums_cat_covariates_cats_[i], cat_covariates_embeds_nums[i])
                self.cat_covariates_embedders.append(emb)
                setattr(self, f"cat_covariate{i}_embedder", emb)
        n_embed = sum(cat_covariates_embeds_nums) if self.n_cat_covariates_ > 0 else 0

        # node-label aggregators (modules/vgpgae.py:274-290)
        agg = {"one-hop-norm": OneHopGCNNormNodeLabelAggregator, "one-hop-sum": OneHopSumNodeLabelAggregator}
        if node_label_method == "one-hop-attention":
            self.rna_node_label_aggregator = OneHopAttentionNodeLabelAggregator(modality="rna", n_input=n_input)
        elif node_label_method in agg:
            self.rna_node_label_aggregator = agg[node_label_method](modality="rna")
        else:
            raise ValueError(f"unknown node_label_method {node_label_method}")
        if "atac" in self.modalities_ and node_label_method in agg:
            self.atac_node_label_aggregator = agg[node_label_method](modality="atac")

        self.encoder = Encoder(
            n_input=n_input,
            n_cat_covariates_embed_input=(n_embed if "encoder" in self.cat_covariates_embeds_injection_ else 0),
            n_fc_layers=n_fc_layers_encoder, n_layers=n_layers_encoder, n_hidden=n_hidden_encoder,
            n_latent=n_prior_gp, n_addon_latent=n_addon_gp, conv_layer=conv_layer_encoder,
            n_attention_heads=encoder_n_attention_heads, dropout_rate=dropout_rate_encoder,
            activation=torch.relu, use_bn=encoder_use_bn)
        self.graph_decoder = CosineSimGraphDecoder(dropout_rate=dropout_rate_graph_decoder)

        # add-on masks and decoders (modules/vgpgae.py:326-402, 463-496)
        for mod in self.modalities_:
            for ent in ("target", "source"):
                a = set(features_idx_dict[f"{ent}_masked_{mod}_idx"])
                b = set(features_idx_dict[f"{ent}_unmasked_{mod}_idx"])
                assert len(a.intersection(b)) == 0
        for ent in ("target", "source"):
            if n_addon_gp > 0:
                m = torch.zeros(n_addon_gp, n_output_genes, dtype=torch.float32)
                m[:, features_idx_dict[f"{ent}_unmasked_rna_idx"]] = 1.
                setattr(self, f"{ent}_rna_decoder_addon_mask", m)
                features_idx_dict[f"{ent}_addon_rna_idx"] = features_idx_dict[f"{ent}_unmasked_rna_idx"]
                if "atac" in self.modalities_:
                    gp = self.gene_peaks_mask_
                    gp = gp.to_dense() if gp.is_sparse else gp
                    am = (m.to(torch.float64) @ gp.to(torch.float64)) > 0
                    setattr(self, f"{ent}_atac_decoder_addon_mask", am)
                    features_idx_dict[f"{ent}_addon_atac_idx"] = torch.nonzero(
                        am.sum(dim=0) > 0).squeeze().tolist()
            else:
                for mod in self.modalities_:
                    setattr(self, f"{ent}_{mod}_decoder_addon_mask", None)
                    features_idx_dict[f"{ent}_addon_{mod}_idx"] = None
        inj = {"rna": "gene_expr_decoder", "atac": "chrom_access_decoder"}
        n_out = {"rna": n_output_genes, "atac": n_output_peaks}
        for mod in self.modalities_:
            for ent in ("target", "source"):
                dec = MaskedOmicsFeatureDecoder(
                    modality=mod, entity=ent, n_prior_gp_input=n_prior_gp, n_addon_gp_input=n_addon_gp,
                    n_cat_covariates_embed_input=(n_embed if inj[mod] in self.cat_covariates_embeds_injection_
                                                  else 0),
                    n_output=n_out[mod], mask=getattr(self, f"{ent}_{mod}_decoder_mask"),
                    addon_mask=getattr(self, f"{ent}_{mod}_decoder_addon_mask"),
                    masked_features_idx=features_idx_dict[f"{ent}_masked_{mod}_idx"], recon_loss="nb")
                setattr(self, f"{ent}_{mod}_decoder", dec)
        for mod in self.modalities_:
            for ent in ("target", "source"):
                setattr(self, f"{ent}_{mod}_theta", nn.Parameter(torch.randn(n_out[mod])))
        self.register_buffer("running_mean_abs_mu", torch.zeros(P))

        # dynamic decoder masks: rna = one alive flag per gene program, atac = full [P, Fp] bool
        self._rna_gp_alive = torch.ones(P, dtype=torch.bool)
        if "atac" in self.modalities_:
            self.target_atac_dynamic_decoder_mask = torch.ones(P, n_output_peaks, dtype=torch.bool)
            self.source_atac_dynamic_decoder_mask = torch.ones(P, n_output_peaks, dtype=torch.bool)
        self._mask_version = 0
        self._sel_cache = {}
        self._static_dev = {}
        self.compute_dtype = torch.float32
        self.materialize_outputs = True
        self._eps_queue: Optional[list] = None      # test hook: injected reparameterisation noise

    # ------------------------------------------------------------------ configuration
    def set_compute_dtype(self, dtype: torch.dtype):
        """torch.float32 (parity mode) or torch.bfloat16 (tensor-core mode)."""
        assert dtype in (torch.float32, torch.bfloat16)
        self.compute_dtype = dtype
        for m in self.modules():
            if m is not self and hasattr(m, "compute_dtype"):
                m.compute_dtype = dtype
        return self

    @property
    def device(self):
        return self.running_mean_abs_mu.device

    @property
    def target_rna_dynamic_decoder_mask(self) -> torch.Tensor:
        return self._rna_gp_alive[:, None].expand(-1, self.n_output_genes_).clone()

    @property
    def source_rna_dynamic_decoder_mask(self) -> torch.Tensor:
        return self._rna_gp_alive[:, None].expand(-1, self.n_output_genes_).clone()

    # ------------------------------------------------------------------ reconstructed features
    def _static_mask(self, key: str) -> torch.Tensor:
        """(prior ++ add-on) static mask [P, F_mod] as float on the module device (cached)."""
        dev = self.device
        hit = self._static_dev.get((key, dev))
        if hit is None:
            m = getattr(self, f"{key}_decoder_mask").to(dev, torch.float32)
            if self.n_addon_gp_ > 0:
                m = torch.cat((m, getattr(self, f"{key}_decoder_addon_mask").to(dev, torch.float32)), dim=0)
            hit = m
            self._static_dev[(key, dev)] = hit
        return hit

    def _reconstructed(self, key: str, need_index: bool):
        """uint8 [F_mod] flag (+ index tensor when ``need_index``) of the features connected to
        >= 1 live gene program (modules/vgpgae.py:667-690, 751-778).  The flag is computed on the
        device without a host sync; the index list (``features_idx_dict_[..reconstructed..]``,
        a host sync in the reference on every forward) only when outputs are materialised."""
        dev = self.device
        hit = self._sel_cache.get((key, dev))
        if hit is None or hit[0] != self._mask_version:
            static = self._static_mask(key)
            if key.endswith("rna"):
                n_per_feat = self._rna_gp_alive.to(dev, torch.float32) @ static
            else:
                dyn = getattr(self, f"{key}_dynamic_decoder_mask").to(dev, torch.float32)
                n_per_feat = (static * dyn).sum(0)
            hit = [self._mask_version, (n_per_feat != 0).to(torch.uint8).contiguous(), None]
            self._sel_cache[(key, dev)] = hit
        if need_index and hit[2] is None:
            hit[2] = torch.nonzero(hit[1]).flatten()
            ent, mod = key.split("_")
            self.features_idx_dict_[f"{ent}_reconstructed_{mod}_idx"] = hit[2].tolist()
        return hit[1], hit[2]

    def _atac_dynamic_mask_on_device(self, key: str) -> torch.Tensor:
        """The [P, n_peaks] bool dynamic mask of an ATAC decoder, moved to the module's device ONCE (it is a plain
        attribute, not a buffer, as in the reference: a ``.to(device)`` per forward was an 80 MB pageable
        host-to-device copy per decoder and iteration at config 4)."""
        name = f"{key}_dynamic_decoder_mask"
        dyn = getattr(self, name)
        if dyn.device != self.device:
            dyn = dyn.to(self.device)
            setattr(self, name, dyn)
        return dyn

    # ------------------------------------------------------------------ batch normalisation
    def _embed(self, cats: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        if self.n_cat_covariates_ == 0:
            return None
        return torch.cat([emb(cats[:, i]) for i, emb in enumerate(self.cat_covariates_embedders)], dim=1)

    def _prepare(self, batch, decoder: str, need_x_enc: bool = True):
        """-> (graph, x_enc tile, labels(n) accessor, aggregated-neighbour accessor, embed, batch rows).
        ``need_x_enc=False`` (the encoder ran on a merged batch, ``forward_joint``): no encoder tile."""
        dev = self.device
        store = getattr(batch, "store", None)
        embed = self._embed(getattr(batch, "cat_covariates_cats", None))
        enc_embed = embed if (embed is not None and "encoder" in self.cat_covariates_embeds_injection_) else None
        ld = pad8(self.encoder.n_input)
        if store is not None:
            graph = batch.graph
            n_sub = graph.n_nodes
            if not need_x_enc:
                x_enc = None
            elif enc_embed is None:
                x_enc = ops.csr_gather_rows(store, batch.n_id, n_sub, self.log_variational_, self.compute_dtype, ld)
            else:
                xd = ops.csr_gather_rows(store, batch.n_id, n_sub, self.log_variational_, torch.float32)
                x_enc = self.encoder.pack_input(xd, enc_embed)
            n_id = batch.n_id

            def target_rows(n):
                return ops.csr_gather_rows(store, n_id, n, False, torch.float32)

            def neighbor_rows(n):
                return self.rna_node_label_aggregator.forward_store(store, n_id, graph, n)
        else:
            x = batch.x
            if not x.is_cuda:
                raise RuntimeError("VGPGAE.forward needs CUDA tensors (nichecompass_b200 has no CPU path)")
            x = x.float().contiguous()
            n_sub = x.size(0)
            if decoder == "omics":
                n_dst = int(batch.batch_size)
            else:
                n_dst = int(getattr(batch, "n_seed", n_sub))
            graph = getattr(batch, "graph", None)
            if graph is None:
                graph = SubgraphCSR.from_edge_index(batch.edge_index, n_sub, n_dst)
            if enc_embed is None:
                x_enc = ops.dense_gather_rows(x, None, n_sub, self.log_variational_, self.compute_dtype, ld)
            else:
                xd = ops.dense_gather_rows(x, None, n_sub, self.log_variational_, torch.float32)
                x_enc = self.encoder.pack_input(xd, enc_embed)

            def target_rows(n):
                return x[:n]

            def neighbor_rows(n):
                if self.node_label_method_ == "one-hop-attention":
                    Fg = self.n_output_genes_
                    return self.rna_node_label_aggregator(x[:, :Fg], graph, n_rows=n)[0]
                return self.rna_node_label_aggregator(x, graph, n_rows=n)[0]
        return graph, x_enc, target_rows, neighbor_rows, embed

    def _draw_eps(self, like: torch.Tensor) -> Optional[torch.Tensor]:
        if not self.training:
            return None
        if self._eps_queue:
            return self._eps_queue.pop(0).to(like.device, torch.float32).contiguous()
        return torch.randn_like(like)

    # ------------------------------------------------------------------ forward
    def forward(self,
                data_batch,
                decoder: Literal["graph", "omics"],
                use_only_active_gps: bool = False,
                return_agg_weights: bool = False,
                update_atac_dynamic_decoder_mask: bool = False,
                _latent=None, _labels=None) -> dict:
        """``_latent`` (internal, ``forward_joint``): (encoder output of a merged batch, first row of this
        batch's seeds in it) -- the encoder is then not run here."""
        graph, x_enc, target_rows, neighbor_rows, embed = self._prepare(data_batch, decoder, _latent is None)
        P = self.n_prior_gp_ + self.n_addon_gp_
        if _latent is None:
            ml_u, row_off = self.encoder.encode(x_enc, graph), 0      # [n_dst, 2P]
        else:
            ml_u, row_off = _latent
        if decoder == "omics":
            nb = int(data_batch.batch_size)
            ml = ml_u[row_off:row_off + nb]
            batch_rows = None
        elif decoder == "graph":
            eli = data_batch.edge_label_index
            batch_rows = torch.cat((eli[0], eli[1]), 0).to(torch.int32)
            if row_off:
                batch_rows = batch_rows + row_off
            batch_rows = batch_rows.contiguous()
            ml = GatherRowsFn.apply(ml_u, batch_rows)
        else:
            raise ValueError(f"unknown decoder {decoder}")
        self.mu, self.logstd = SplitLatentFn.apply(ml, P) if ml.requires_grad else (ml[:, :P], ml[:, P:])
        output = {"mu": self.mu, "logstd": self.logstd}
        active = self.get_active_gp_mask() if use_only_active_gps else None
        active_u8 = active.to(torch.uint8) if active is not None else None
        eps = self._draw_eps(self.mu)
        if eps is None and active_u8 is None:
            z = self.mu
        else:
            z = ReparamFn.apply(self.mu, self.logstd, eps, active_u8)
            if eps is None:
                # evaluation mode: the reference's z IS mu (vgaemodulemixin.py:39-41), so zeroing the
                # inactive gene programs in z (vgpgae.py:637) also zeroes them in output["mu"]
                self.mu = z
                output["mu"] = z

        if decoder == "graph":
            output["edge_recon_labels"] = data_batch.edge_label
            if self.n_cat_covariates_ > 0:
                cats = data_batch.cat_covariates_cats
                li0, li1 = eli[0].long(), eli[1].long()
                same = [cats[li0, i] == cats[li1, i] for i in range(self.n_cat_covariates_)]
                incl = [s if no_edge else torch.ones_like(s) for s, no_edge in
                        zip(same, self.cat_covariates_no_edges_)]
                output["edge_same_cat_covariates_cat"] = same
                output["edge_incl"] = torch.all(torch.stack(incl), dim=0)
            else:
                output["edge_same_cat_covariates_cat"] = None
                output["edge_incl"] = None
            output["edge_recon_logits"] = self.graph_decoder(z=z)
            return output

        # ---- omics decoder
        with torch.no_grad():
            if self.training:
                mu_d = self.mu.detach()
                call("nc_abs_mean_ema", mu_d, mu_d.stride(0), mu_d.size(0), P, 0.1, self.running_mean_abs_mu)
                if getattr(self, "_dp_sync_ema", False):
                    # data-parallel training (train/trainer.py of this build): every rank saw a different node batch
                    import torch.distributed as dist
                    dist.all_reduce(self.running_mean_abs_mu, op=dist.ReduceOp.AVG)
            if use_only_active_gps:
                # all on the device, no host sync (the reference's boolean-mask assignments sync)
                self.running_mean_abs_mu.mul_(active)
                if self._rna_gp_alive.device != active.device:
                    self._rna_gp_alive = self._rna_gp_alive.to(active.device)
                self._rna_gp_alive.logical_and_(active)      # in place: one persistent buffer (CUDA-graph replays)
                if "atac" in self.modalities_:
                    for ent in ("target", "source"):
                        dm = getattr(self, f"{ent}_atac_dynamic_decoder_mask").to(active.device)
                        setattr(self, f"{ent}_atac_dynamic_decoder_mask", dm & active[:, None])
                self._mask_version += 1
        Fg = self.n_output_genes_
        if _labels is not None:
            x_t, x_n = _labels                                   # staged with the joint encoder input
        else:
            x_t = target_rows(nb)
            x_n = neighbor_rows(nb)
        emb_b = embed[:nb] if embed is not None else None
        if "atac" in self.modalities_ and update_atac_dynamic_decoder_mask:
            self._update_atac_dynamic_masks()
        output["node_labels"] = {}
        output["_nb_loss"] = {}
        cols = {"rna": slice(0, Fg), "atac": slice(Fg, Fg + self.n_output_peaks_)}
        inj = {"rna": "gene_expr_decoder", "atac": "chrom_access_decoder"}
        for mod in self.modalities_:
            for ent, xsrc in (("target", x_t), ("source", x_n)):
                key = f"{ent}_{mod}"
                sel, ridx = self._reconstructed(key, self.materialize_outputs)
                dec = getattr(self, f"{key}_decoder")
                dyn = self._atac_dynamic_mask_on_device(key) if mod == "atac" else None
                e = emb_b if (emb_b is not None and inj[mod] in self.cat_covariates_embeds_injection_) else None
                logits = dec.logits(z, e, dyn)
                labels_full = xsrc[:, cols[mod]]
                loss_nb, means = compute_omics_recon_nb_loss_from_logits(
                    logits, labels_full, getattr(self, f"{key}_theta"), sel)
                output["_nb_loss"][key] = loss_nb
                if self.materialize_outputs:
                    output["node_labels"][key] = labels_full.index_select(1, ridx)
                    output[f"{key}_nb_means"] = means.index_select(1, ridx)
        return output

    # ------------------------------------------------------------------ one encoder pass per iteration
    def joint_supported(self, node_batch, edge_batch) -> bool:
        """The node batch and the edge batch of one training iteration (train/trainer.py:368-389) can share
        ONE encoder pass when both are index batches over the same count store, the encoder is GATv2 (a
        target's output depends on its own neighbourhood only, so a block-diagonal union of the two
        sub-graphs gives the same rows) and no covariate embedding is injected into the encoder input."""
        store = getattr(node_batch, "store", None)
        return (store is not None and getattr(edge_batch, "store", None) is store
                and self.encoder.conv_layer == "gatv2conv" and self.encoder.n_layers == 1
                and self.encoder.n_fc_layers == 1     # BatchNorm statistics are per encoder pass in the reference
                and not (self.n_cat_covariates_ > 0 and "encoder" in self.cat_covariates_embeds_injection_))

    def _joint_input(self, node_batch, edge_batch, background: bool = False):
        """Block-diagonal union of the two sampled sub-graphs + the gathered (log1p) count rows of its nodes."""
        gn, ge = node_batch.graph, edge_batch.graph
        a_seed, a_sub, b_seed, b_sub = gn.n_dst, gn.n_nodes, ge.n_dst, ge.n_nodes
        dev = gn.rowptr.device
        m_rowptr = torch.empty(a_seed + b_seed + 1, dtype=torch.int32, device=dev)
        m_col = torch.empty(max(gn.e_dst + ge.e_dst, 1), dtype=torch.int32, device=dev)     # seed rows only
        m_nid = torch.empty(a_sub + b_sub, dtype=torch.int32, device=dev)
        call("nc_merge_subgraphs", gn.rowptr, gn.col, node_batch.n_id, a_seed, a_sub, ge.rowptr, ge.col,
             edge_batch.n_id, b_seed, b_sub, m_rowptr, m_col, m_nid)
        gm = SubgraphCSR(m_rowptr, m_col, a_sub + b_sub, a_seed + b_seed, max(gn.max_deg, ge.max_deg),
                         e_dst=gn.e_dst + ge.e_dst)
        x_enc = ops.csr_gather_rows(node_batch.store, m_nid, a_sub + b_sub, self.log_variational_,
                                    self.compute_dtype, pad8(self.encoder.n_input), background=background)
        return gm, x_enc

    stage_labels = True

    def stage_joint_input(self, node_batch, edge_batch, stream: "torch.cuda.Stream"):
        """Build the joint encoder input of a FUTURE iteration on ``stream`` while the compute stream runs the
        current one.  With the count store in pinned host memory the row gather is bound by the host link
        (~20 MB per iteration at C2); staged one iteration ahead it leaves the critical path.
        ``forward_joint`` picks the staged input up (and waits for its event) when it is given the same pair."""
        main = torch.cuda.current_stream()
        stream.wait_stream(main)                  # the batches' tensors were produced / published on ``main``
        with torch.cuda.stream(stream):
            gm, x_enc = self._joint_input(node_batch, edge_batch, background=True)
            labels = None
            if self.stage_labels and self.node_label_method_ != "one-hop-attention":   # parameter-free: a pure function of the batch
                nb = int(node_batch.batch_size)
                labels = (ops.csr_gather_rows(node_batch.store, node_batch.n_id, nb, False, torch.float32,
                                              background=True),
                          self.rna_node_label_aggregator.forward_store(node_batch.store, node_batch.n_id,
                                                                       node_batch.graph, nb))
            ev = torch.cuda.Event()
            ev.record(stream)
        cached = [t for v in node_batch.graph._norm.values() for t in v if t is not None]   # gcn_weights cache
        for t in (gm.rowptr, gm.col, x_enc) + (labels or ()) + tuple(cached):
            t.record_stream(main)
        edge_batch._joint_input = (node_batch, gm, x_enc, ev, labels)

    def forward_joint(self, node_batch, edge_batch, use_only_active_gps: bool = False):
        """``(forward(node_batch, "omics"), forward(edge_batch, "graph"))`` with the two encoder passes merged:
        ``nc_merge_subgraphs`` builds the block-diagonal problem [node seeds | edge seeds | others], the count
        rows are gathered once, every encoder GEMM / attention kernel (forward and backward) runs once on
        the union instead of once per batch."""
        staged = getattr(edge_batch, "_joint_input", None)
        if staged is not None and staged[0] is node_batch:
            _, gm, x_enc, ev, labels = staged
            if ev is not None:
                torch.cuda.current_stream().wait_event(ev)      # staged on a side stream (``stage_joint_input``)
            edge_batch._joint_input = None
        else:
            gm, x_enc = self._joint_input(node_batch, edge_batch)
            labels = None
        a_seed = node_batch.graph.n_dst
        ml_u = self.encoder.encode(x_enc, gm)
        node_out = self.forward(node_batch, "omics", use_only_active_gps, _latent=(ml_u, 0), _labels=labels)
        edge_out = self.forward(edge_batch, "graph", use_only_active_gps, _latent=(ml_u, a_seed))
        return node_out, edge_out

    def _update_atac_dynamic_masks(self):
        """Turn off peaks whose mapped genes were all regularised away (modules/vgpgae.py:813-857)."""
        with torch.no_grad():
            Fg = self.n_output_genes_
            w = torch.round(self.get_gp_weights(only_masked_features=False)[0], decimals=4)
            nz = torch.ne(w, 0)
            gp = self.gene_peaks_mask_
            gp = (gp.to_dense() if gp.is_sparse else gp).to(w.device, torch.float32)
            for ent, blk in (("target", nz[:Fg]), ("source", nz[Fg:])):
                m = torch.ne(blk.t().to(torch.float32) @ gp, 0)
                cur = getattr(self, f"{ent}_atac_dynamic_decoder_mask").to(w.device)
                setattr(self, f"{ent}_atac_dynamic_decoder_mask", cur & m)
            self._mask_version += 1

    # ------------------------------------------------------------------ loss
    def loss(self,
             edge_model_output: dict,
             node_model_output: dict,
             lambda_l1_masked: float,
             l1_targets_mask: Optional[torch.Tensor],
             l1_sources_mask: Optional[torch.Tensor],
             lambda_l1_addon: float,
             lambda_group_lasso: float,
             lambda_gene_expr_recon: float = 300.,
             lambda_chrom_access_recon: float = 100.,
             lambda_edge_recon: Optional[float] = 500000.,
             lambda_cat_covariates_contrastive: Optional[float] = 100000.,
             contrastive_logits_pos_ratio: float = 0.125,
             contrastive_logits_neg_ratio: float = 0.,
             edge_recon_active: bool = True,
             cat_covariates_contrastive_active: bool = True) -> dict:
        """Same composition as modules/vgpgae.py:1032-1143."""
        L = {}
        L["kl_reg_loss"] = compute_kl_reg_loss(node_model_output["mu"], node_model_output["logstd"])
        if self.include_edge_kl_loss_:
            L["kl_reg_loss"] = L["kl_reg_loss"] + compute_kl_reg_loss(edge_model_output["mu"],
                                                                     edge_model_output["logstd"])
        L["edge_recon_loss"] = lambda_edge_recon * compute_edge_recon_loss(
            edge_recon_logits=edge_model_output["edge_recon_logits"],
            edge_recon_labels=edge_model_output["edge_recon_labels"],
            edge_incl=edge_model_output["edge_incl"])
        if (edge_model_output["edge_same_cat_covariates_cat"] is not None) and (
                lambda_cat_covariates_contrastive > 0):
            L["cat_covariates_contrastive_loss"] = (
                lambda_cat_covariates_contrastive * compute_cat_covariates_contrastive_loss(
                    edge_recon_logits=edge_model_output["edge_recon_logits"],
                    edge_recon_labels=edge_model_output["edge_recon_labels"],
                    edge_same_cat_covariates_cat=edge_model_output["edge_same_cat_covariates_cat"],
                    contrastive_logits_pos_ratio=contrastive_logits_pos_ratio,
                    contrastive_logits_neg_ratio=contrastive_logits_neg_ratio))
        nbl = node_model_output.get("_nb_loss")
        if nbl is None:
            raise RuntimeError("node_model_output was not produced by VGPGAE.forward(decoder='omics')")
        L["gene_expr_recon_loss"] = lambda_gene_expr_recon * (nbl["target_rna"] + nbl["source_rna"])
        L["masked_gp_l1_reg_loss"] = lambda_l1_masked * compute_gp_l1_reg_loss(
            self, gp_type="prior", l1_targets_mask=l1_targets_mask, l1_sources_mask=l1_sources_mask) \
            if lambda_l1_masked != 0 else self._zero_scalar()
        L["group_lasso_reg_loss"] = lambda_group_lasso * compute_gp_group_lasso_reg_loss(self) \
            if lambda_group_lasso != 0 else self._zero_scalar()
        if self.n_addon_gp_ != 0:
            L["addon_gp_l1_reg_loss"] = lambda_l1_addon * compute_gp_l1_reg_loss(self, gp_type="addon") \
                if lambda_l1_addon != 0 else self._zero_scalar()
        if "atac" in self.modalities_:
            L["chrom_access_recon_loss"] = lambda_chrom_access_recon * (nbl["target_atac"] + nbl["source_atac"])
        # global / optimisation loss: the reference's chain of additions (modules/vgpgae.py:1100-1143),
        # evaluated as one stacked sum per loss (two small kernels instead of ~20)
        g_terms, o_terms = [L["kl_reg_loss"]], [L["kl_reg_loss"]]
        if self.include_edge_recon_loss_:
            g_terms.append(L["edge_recon_loss"])
            if edge_recon_active:
                o_terms.append(L["edge_recon_loss"])
        if self.include_cat_covariates_contrastive_loss_ and "cat_covariates_contrastive_loss" in L:
            g_terms.append(L["cat_covariates_contrastive_loss"])
            if cat_covariates_contrastive_active:
                o_terms.append(L["cat_covariates_contrastive_loss"])
        if self.include_gene_expr_recon_loss_:
            for k in ("gene_expr_recon_loss", "group_lasso_reg_loss", "masked_gp_l1_reg_loss"):
                g_terms.append(L[k])
                o_terms.append(L[k])
            if self.n_addon_gp_ != 0:
                g_terms.append(L["addon_gp_l1_reg_loss"])
                o_terms.append(L["addon_gp_l1_reg_loss"])
        if ("atac" in self.modalities_) and self.include_chrom_access_recon_loss_:
            g_terms.append(L["chrom_access_recon_loss"])
            o_terms.append(L["chrom_access_recon_loss"])
        o = torch.stack(o_terms).sum()
        same = len(g_terms) == len(o_terms)
        g = o if same else torch.stack(g_terms).sum()
        L["global_loss"] = g
        L["optim_loss"] = o
        return L

    def _zero_scalar(self) -> torch.Tensor:
        z = self._static_dev.get(("zero", self.device))
        if z is None:
            z = self._static_dev[("zero", self.device)] = torch.zeros((), device=self.device)
        return z

    # ------------------------------------------------------------------ gp weights / active mask
    @torch.no_grad()
    def get_gp_weights(self, only_masked_features: bool = False,
                       gp_type: Literal["all", "prior", "addon", "masked"] = "all") -> List[torch.Tensor]:
        """Decoder gene-program weights per modality, [2 * n_features, n_gps]
        (modules/vgpgae.py:1146-1202)."""
        out = []
        for mod in self.modalities_:
            t = getattr(self, f"target_{mod}_decoder").nb_means_normalized_decoder
            s = getattr(self, f"source_{mod}_decoder").nb_means_normalized_decoder
            w = torch.cat((t.masked_l.weight.data.clone(), s.masked_l.weight.data.clone()), dim=0)
            if gp_type != "masked" and self.n_addon_gp_ > 0:
                wa = torch.cat((t.addon_l.weight.data.clone(), s.addon_l.weight.data.clone()), dim=0)
                if gp_type == "all":
                    w = torch.cat([w, wa], dim=1)
                elif gp_type == "addon":
                    w = wa
            if only_masked_features:
                w = w[self.features_idx_dict_[f"masked_{mod}_idx"], :]
            out.append(w)
        return out

    @torch.no_grad()
    def _gp_block_stats(self, layer: str):
        """(sum_f |W|, #non-zero W) per gene program over the target ++ source rna decoder weights."""
        ws = [getattr(getattr(self, f"{ent}_rna_decoder").nb_means_normalized_decoder, layer).weight
              for ent in ("target", "source")]
        n = ws[0].size(1)
        abs_sum = torch.zeros(n, dtype=torch.float32, device=ws[0].device)
        nnz = torch.zeros(n, dtype=torch.float32, device=ws[0].device)
        for w in ws:
            wd = w.detach()
            call("nc_weight_col_stats", wd, wd.stride(0), wd.size(0), n, None, None, 0, abs_sum, None, nnz, None)
        return abs_sum, nnz

    @torch.no_grad()
    def get_active_gp_mask(self,
                           abs_gp_weights_agg_mode: Literal["sum", "nzmeans", "sum+nzmeans"] = "sum+nzmeans",
                           return_gp_weights: bool = False,
                           normalize_gp_weights_with_features_scale_factors: bool = False
                           ) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        """Active gene programs: aggregated |weight| * running mean |mu| above
        ``active_gp_thresh_ratio_`` times the maximum, by sum OR by mean over non-zero weights,
        thresholded separately for prior and add-on programs unless ``active_gp_type_ == 'mixed'``
        (modules/vgpgae.py:1205-1334)."""
        if abs_gp_weights_agg_mode not in ("sum", "nzmeans", "sum+nzmeans", "nzmedians", "sum+nzmedians"):
            raise ValueError(f"unknown aggregation mode {abs_gp_weights_agg_mode}")
        if normalize_gp_weights_with_features_scale_factors or "nzmedians" in abs_gp_weights_agg_mode:
            # analysis-time variants (medians / per-feature scale factors): evaluated on the materialised weight
            # matrix, statement by statement as modules/vgpgae.py:1254-1330 -- not on the training path
            return self._active_gp_mask_general(abs_gp_weights_agg_mode, return_gp_weights,
                                                normalize_gp_weights_with_features_scale_factors)
        dev = self.device
        Pp, Pa = self.n_prior_gp_, self.n_addon_gp_
        abs_p, nnz_p = self._gp_block_stats("masked_l")
        blocks = [(slice(None, Pp), abs_p, nnz_p)]
        if Pa > 0:
            abs_a, nnz_a = self._gp_block_stats("addon_l")
            if self.active_gp_type_ == "mixed":
                blocks = [(slice(None, None), torch.cat((abs_p, abs_a)), torch.cat((nnz_p, nnz_a)))]
            else:
                blocks.append((slice(Pp, None), abs_a, nnz_a))
        active = torch.zeros(Pp + Pa, dtype=torch.bool, device=dev)
        ratio = self.active_gp_thresh_ratio_
        for sl, abs_sum, nnz in blocks:
            ema = self.running_mean_abs_mu[sl]
            sums = ema.abs() * abs_sum
            if abs_gp_weights_agg_mode in ("sum", "sum+nzmeans"):
                active[sl] |= sums >= ratio * sums.max()
            if abs_gp_weights_agg_mode in ("nzmeans", "sum+nzmeans"):
                cnt = torch.where(ema != 0, nnz, torch.zeros_like(nnz))
                nzmeans = torch.nan_to_num(sums / cnt)
                active[sl] |= nzmeans >= ratio * nzmeans.max()
        if return_gp_weights:
            return active, self.get_gp_weights(gp_type="all")[0][:, active]
        return active

    @torch.no_grad()
    def _active_gp_mask_general(self, mode: str, return_gp_weights: bool, normalize: bool):
        dev = self.device
        Pp, Pa = self.n_prior_gp_, self.n_addon_gp_
        active = torch.zeros(Pp + Pa, dtype=torch.bool, device=dev)
        if self.active_gp_type_ == "mixed":
            types = [("all", slice(None, None))]
        elif Pa > 0:
            types = [("masked", slice(None, Pp)), ("addon", slice(Pp, None))]
        else:
            types = [("masked", slice(None, Pp))]
        ratio = self.active_gp_thresh_ratio_
        for gp_type, sl in types:
            w = self.get_gp_weights(only_masked_features=False, gp_type=gp_type)[0]
            if normalize:
                w = w / self.features_scale_factors_[:, None].to(dev)
            w = self.running_mean_abs_mu[sl] * w
            sums = w.abs().sum(0)
            if mode in ("sum", "sum+nzmeans", "sum+nzmedians"):
                active[sl] |= sums >= ratio * sums.max()
            if mode in ("nzmeans", "sum+nzmeans"):
                nzmeans = torch.nan_to_num(sums / torch.count_nonzero(w, dim=0))
                active[sl] |= nzmeans >= ratio * nzmeans.max()
            if mode in ("nzmedians", "sum+nzmedians"):
                with_nan = torch.where(w == 0, torch.full_like(w, float("nan")), w.abs())
                med = torch.nan_to_num(torch.nanmedian(with_nan, dim=0).values)
                active[sl] |= med >= 0.01 * med.max()               # the reference hard-codes 0.01 here (:1317)
        if return_gp_weights:
            return active, self.get_gp_weights(gp_type="all")[0][:, active]
        return active

    # ------------------------------------------------------------------ decoder outputs (post-hoc retrieval)
    @torch.no_grad()
    def get_omics_decoder_outputs(self, node_batch, only_active_gps: bool = True) -> dict:
        """Negative-binomial means of every decoder for the seeds of ``node_batch`` over ALL output features
        (modules/vgpgae.py:1439-1620): encoder -> z (= mu in eval mode) with inactive gene programs zeroed ->
        library size = sum of the node labels over all features -> exp(log lib) * softmax(decoder logits).
        No column selection here, unlike ``forward``.  Same kernels as the training path:
        ``nc_nb_loss_fused`` with every column selected returns the means."""
        graph, x_enc, target_rows, neighbor_rows, embed = self._prepare(node_batch, "omics")
        P = self.n_prior_gp_ + self.n_addon_gp_
        nb = int(node_batch.batch_size)
        ml = self.encoder.encode(x_enc, graph)[:nb]
        mu, logstd = ml[:, :P].contiguous(), ml[:, P:].contiguous()
        eps = self._draw_eps(mu)
        active_u8 = self.get_active_gp_mask().to(torch.uint8) if only_active_gps else None
        z = mu if (eps is None and active_u8 is None) else ReparamFn.apply(mu, logstd, eps, active_u8)
        x_t, x_n = target_rows(nb), neighbor_rows(nb)
        Fg = self.n_output_genes_
        cols = {"rna": slice(0, Fg), "atac": slice(Fg, Fg + self.n_output_peaks_)}
        inj = {"rna": "gene_expr_decoder", "atac": "chrom_access_decoder"}
        emb_b = embed[:nb] if embed is not None else None
        out = {"node_labels": {}}
        for mod in self.modalities_:
            for ent, xsrc in (("target", x_t), ("source", x_n)):
                key = f"{ent}_{mod}"
                dec = getattr(self, f"{key}_decoder")
                dyn = self._atac_dynamic_mask_on_device(key) if mod == "atac" else None
                e = emb_b if (emb_b is not None and inj[mod] in self.cat_covariates_embeds_injection_) else None
                logits = dec.logits(z, e, dyn)
                labels = xsrc[:, cols[mod]].contiguous()
                every = torch.ones(labels.size(1), dtype=torch.uint8, device=labels.device)
                _, means = compute_omics_recon_nb_loss_from_logits(logits, labels, getattr(self, f"{key}_theta"), every)
                out["node_labels"][key] = labels
                out[f"{key}_nb_means"] = means
        return out

    def log_module_hyperparams_to_mlflow(self, excluded_attr: list = ["features_idx_dict_", "gene_peaks_mask_"]):
        """mlflow is not part of this build; logging is a no-op (modules/vgpgae.py:1336-1366)."""
        return None

    # ------------------------------------------------------------------ latent inference
    def get_latent_representation(self, node_batch, only_active_gps: bool = True, return_mu_std: bool = False
                                  ) -> Union[torch.Tensor, Tuple[torch.Tensor, torch.Tensor]]:
        """modules/vgpgae.py:1368-1437."""
        graph, x_enc, _, _, _ = self._prepare(node_batch, "omics")
        P = self.n_prior_gp_ + self.n_addon_gp_
        ml = self.encoder.encode(x_enc, graph)[:int(node_batch.batch_size)]
        mu, logstd = ml[:, :P], ml[:, P:]
        if only_active_gps:
            a = self.get_active_gp_mask()
            mu, logstd = mu[:, a], logstd[:, a]
        if return_mu_std:
            return mu, torch.exp(logstd)
        eps = self._draw_eps(mu)
        if eps is None:
            return mu
        return ReparamFn.apply(mu.contiguous(), logstd.contiguous(), eps, None)
