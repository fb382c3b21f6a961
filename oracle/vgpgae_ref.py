"""
CPU restatement of the VGPGAE module: forward, loss, active-GP mask and latent inference
(TEST INFRASTRUCTURE -- see oracle/__init__.py).  Own words, plain torch, any float dtype.

Follows (relative to /root/reference/src/nichecompass):
  modules/vgpgae.py:541-930   forward           -> VGPGAERef.forward
  modules/vgpgae.py:932-1143  loss              -> VGPGAERef.loss
  modules/vgpgae.py:1146-1334 gp weights / active-GP mask
  modules/vgpgae.py:1368-1437 latent inference
  modules/losses.py           all six loss functions
  modules/vgaemodulemixin.py:34-41 reparameterisation (eps is injected, never drawn here)
  nn/encoders.py:144-208, nn/decoders.py:39-62,133-169, nn/layers.py:83-144,
  nn/layercomponents.py:53-81, nn/utils.py:8-36, nn/aggregators.py (via pyg_semantics)

Parameters live in a flat ``dict`` keyed exactly like the reference ``state_dict`` so that a
reference checkpoint, this oracle and the CUDA module can be driven by the same tensors.
"""
import math
from typing import Dict, List, Optional

import torch
import torch.nn.functional as F

from . import bf16_emulation as emu
from . import pyg_semantics as pyg


# ----------------------------------------------------------------------------- losses
def kl_reg_loss(mu, logstd):
    """modules/losses.py:255-284."""
    return -0.5 * torch.mean(torch.sum(1 + 2 * logstd - mu ** 2 - torch.exp(logstd) ** 2, 1))


def edge_recon_loss(logits, labels, edge_incl=None):
    """modules/losses.py:123-166: BCE-with-logits, pos_weight = #neg / #pos."""
    if edge_incl is not None:
        logits, labels = logits[edge_incl], labels[edge_incl]
    pos = (labels == 1.).sum()
    neg = (labels == 0.).sum()
    pos_weight = neg / pos
    return F.binary_cross_entropy_with_logits(logits, labels, pos_weight=pos_weight)


def nb_recon_loss(x, mu, theta, eps=1e-8):
    """modules/losses.py:169-212."""
    log_theta_mu_eps = torch.log(theta + mu + eps)
    ll = (theta * (torch.log(theta + eps) - log_theta_mu_eps)
          + x * (torch.log(mu + eps) - log_theta_mu_eps)
          + torch.lgamma(x + theta) - torch.lgamma(theta) - torch.lgamma(x + 1))
    return torch.mean(-ll.sum(-1))


def contrastive_loss(logits, labels, same_cat_list, pos_ratio, neg_ratio):
    """modules/losses.py:15-120 (labels are modified in place, as in the reference)."""
    if same_cat_list is None or (pos_ratio == 0 and neg_ratio == 0):
        return torch.tensor(0.)
    total = torch.zeros((), dtype=logits.dtype)
    for same in same_cat_list:
        diff_logits = logits[~same].detach()
        n_pos = math.ceil(pos_ratio * len(diff_logits))
        n_neg = math.ceil(neg_ratio * len(diff_logits))
        pos_thresh = (torch.topk(diff_logits, n_pos).values[-1] if n_pos > 0
                      else torch.tensor(float("inf")))
        neg_thresh = (torch.topk(diff_logits, n_neg, largest=False).values[-1] if n_neg > 0
                      else torch.tensor(float("-inf")))
        pos_ex = (~same) & (logits >= pos_thresh)
        neg_ex = (~same) & (logits <= neg_thresh)
        labels[pos_ex] = 1
        labels[neg_ex] = 0
        sel = pos_ex | neg_ex
        total = total + F.binary_cross_entropy_with_logits(logits[sel], labels[sel])
    return total


def cosine_similarity_rows(a, b, eps=1e-8):
    """nn/utils.py:8-36."""
    an = a.norm(dim=1, keepdim=True).clamp_min(eps)
    bn = b.norm(dim=1, keepdim=True).clamp_min(eps)
    return ((a / an) * (b / bn)).sum(1)


# ----------------------------------------------------------------------------- module
class VGPGAERef:
    """Functional VGPGAE.  ``cfg`` keys: n_input, n_output_genes, n_output_peaks (0),
    n_hidden, n_prior_gp, n_addon_gp, conv_layer ('gatv2conv'|'gcnconv'), heads, n_layers,
    node_label_method, log_variational, active_gp_thresh_ratio, active_gp_type,
    include_edge_kl_loss, include_edge_recon_loss, include_gene_expr_recon_loss,
    include_chrom_access_recon_loss, include_cat_covariates_contrastive_loss,
    cat_covariates_embeds_injection, cat_covariates_embeds_nums, cat_covariates_no_edges,
    masks: target_rna_mask/source_rna_mask [P_prior, F] (+ *_atac_mask [P_prior, Fp],
    gene_peaks_mask dense [F, Fp]), features_idx_dict."""

    def __init__(self, cfg: dict, params: Dict[str, torch.Tensor]):
        self.cfg = cfg
        self.p = params
        self.training = True
        dt = params["encoder.fc_l1.weight"].dtype
        P = cfg["n_prior_gp"] + cfg["n_addon_gp"]
        self.running_mean_abs_mu = params.get("running_mean_abs_mu",
                                              torch.zeros(P, dtype=dt)).clone()
        self.modalities = ["rna"] + (["atac"] if cfg.get("n_output_peaks", 0) > 0 else [])
        idx = cfg["features_idx_dict"]
        self.addon_mask = {}
        self.dynamic_mask = {}
        for ent in ("target", "source"):
            if cfg["n_addon_gp"] > 0:
                m = torch.zeros(cfg["n_addon_gp"], cfg["n_output_genes"], dtype=dt)
                m[:, idx[f"{ent}_unmasked_rna_idx"]] = 1.
                self.addon_mask[f"{ent}_rna"] = m
                if "atac" in self.modalities:
                    self.addon_mask[f"{ent}_atac"] = (
                        (m.to(torch.int64) @ cfg["gene_peaks_mask"].to(torch.int64)) > 0)
            self.dynamic_mask[f"{ent}_rna"] = torch.ones(P, cfg["n_output_genes"], dtype=torch.bool)
            if "atac" in self.modalities:
                self.dynamic_mask[f"{ent}_atac"] = torch.ones(P, cfg["n_output_peaks"], dtype=torch.bool)

    # ------------------------------------------------------------------ building blocks
    def _embed(self, cats):
        n = len(self.cfg.get("cat_covariates_embeds_nums", []))
        if n == 0:
            return None
        return torch.cat([self.p[f"cat_covariate{i}_embedder.weight"][cats[:, i]]
                          for i in range(n)], dim=1)

    def _conv(self, name, h, edge_index):
        c, p = self.cfg, self.p
        if c["conv_layer"] == "gatv2conv":
            if c.get("emulate_bf16", False):       # tests of the CUDA bf16 mode: oracle/bf16_emulation.py
                return pyg.gatv2_conv(h, edge_index,
                                      emu.round_fwd(p[f"encoder.{name}.lin_l.weight"]), p[f"encoder.{name}.lin_l.bias"],
                                      emu.round_fwd(p[f"encoder.{name}.lin_r.weight"]), p[f"encoder.{name}.lin_r.bias"],
                                      p[f"encoder.{name}.att"], p[f"encoder.{name}.bias"], c["heads"],
                                      round_fn=emu.round_both)
            return pyg.gatv2_conv(h, edge_index,
                                  p[f"encoder.{name}.lin_l.weight"], p[f"encoder.{name}.lin_l.bias"],
                                  p[f"encoder.{name}.lin_r.weight"], p[f"encoder.{name}.lin_r.bias"],
                                  p[f"encoder.{name}.att"], p[f"encoder.{name}.bias"], c["heads"])
        return pyg.gcn_conv(h, edge_index, p[f"encoder.{name}.lin.weight"], p[f"encoder.{name}.bias"])

    def encoder(self, x_enc, edge_index, embed=None):
        """nn/encoders.py:144-208 (dropout 0).  ``n_fc_layers`` 2: fc_l2 -> ReLU -> BatchNorm1d (:180-182), batch
        statistics (biased variance, eps 1e-5) and the momentum-0.1 update of the running buffers in training,
        the running statistics in eval (torch.nn.BatchNorm1d defaults, nn/encoders.py:96)."""
        c = self.cfg
        if embed is not None and c.get("cat_covariates_embed_mode", "input") == "input":
            x_enc = torch.cat((x_enc, embed), dim=1)
        if c.get("emulate_bf16", False):
            # bf16 operands (x_enc, W1 shadow), exact accumulation; d hidden rounded after the ReLU mask,
            # hidden itself stored in bf16 (oracle/bf16_emulation.py; single fc layer, GATv2 only)
            assert c.get("n_fc_layers", 1) == 1 and c.get("n_layers", 1) == 1 and c["conv_layer"] == "gatv2conv"
            pre = F.linear(emu.round_fwd(x_enc), emu.round_fwd(self.p["encoder.fc_l1.weight"]),
                           self.p["encoder.fc_l1.bias"])
            h = emu.round_fwd(torch.relu(emu.round_bwd(pre)))
        else:
            h = torch.relu(F.linear(x_enc, self.p["encoder.fc_l1.weight"], self.p["encoder.fc_l1.bias"]))
        if c.get("n_fc_layers", 1) == 2:
            h = torch.relu(F.linear(h, self.p["encoder.fc_l2.weight"], self.p["encoder.fc_l2.bias"]))
            bn = "encoder.fc_l2_bn."
            if self.training:
                self.p[bn + "num_batches_tracked"] += 1
            h = F.batch_norm(h, self.p[bn + "running_mean"], self.p[bn + "running_var"], self.p[bn + "weight"],
                             self.p[bn + "bias"], self.training, 0.1, 1e-5)
        if embed is not None and c.get("cat_covariates_embed_mode", "input") == "hidden":
            h = torch.cat((h, embed), dim=1)
        if c.get("n_layers", 1) == 2:
            h = torch.relu(self._conv("conv_l1", h, edge_index))
        mu = self._conv("conv_mu", h, edge_index)
        logstd = self._conv("conv_logstd", h, edge_index)
        if c["n_addon_gp"] != 0:
            mu = torch.cat((mu, self._conv("addon_conv_mu", h, edge_index)), dim=1)
            logstd = torch.cat((logstd, self._conv("addon_conv_logstd", h, edge_index)), dim=1)
        return mu, logstd

    def aggregate(self, x, edge_index, modality="rna", return_agg_weights=False):
        m = self.cfg["node_label_method"]
        if m == "one-hop-norm":
            return pyg.one_hop_gcn_norm_aggregate(x, edge_index, return_agg_weights)
        if m == "one-hop-sum":
            return pyg.one_hop_sum_aggregate(x, edge_index)
        pre = f"{modality}_node_label_aggregator"
        return pyg.one_hop_attention_aggregate(
            x, edge_index, self.p[f"{pre}.linear_l_l.weight"], self.p[f"{pre}.linear_r_l.weight"],
            self.p[f"{pre}.attn"], self.cfg.get("aggregator_heads", 4),
            return_agg_weights=return_agg_weights)

    def decoder(self, ent_mod, z, log_lib, embed=None, dynamic_mask=None):
        """nn/decoders.py:133-169 -> nn/layers.py:83-144 -> nn/layercomponents.py:53-81.
        Returns nb_means over ALL output features (softmax first, column select later)."""
        c, p = self.cfg, self.p
        pre = f"{ent_mod}_decoder.nb_means_normalized_decoder"
        Pp, Pa = c["n_prior_gp"], c["n_addon_gp"]
        zp, za = z[:, :Pp], z[:, Pp:Pp + Pa]

        def masked_linear(key, inp, dyn):
            w, m = p[f"{pre}.{key}.weight"], p[f"{pre}.{key}.mask"]
            if c.get("emulate_bf16", False):
                assert dyn is None and embed is None
                all_ones = bool((m == 1).all())
                return emu.MaskedLinearBF16.apply(inp, w * m, all_ones or not emu.mask_is_sparse(m))
            if dyn is not None:
                dyn_t = dyn.t().to(w.dtype)
                with torch.no_grad():
                    w.mul_(dyn_t)                  # weight.data *= dynamic_mask (side effect)
                return F.linear(inp, w * m * dyn_t)
            return F.linear(inp, w * m)

        out = masked_linear("masked_l", zp, None if dynamic_mask is None else dynamic_mask[:Pp])
        if Pa != 0:
            out = out + masked_linear("addon_l", za,
                                      None if dynamic_mask is None else dynamic_mask[Pp:])
        if embed is not None:
            e = F.linear(embed, p[f"{pre}.cat_covariates_embed_l.weight"])
            if Pa != 0:
                out = out + e
            else:
                mod = ent_mod.split("_")[1]
                ent = ent_mod.split("_")[0]
                idx = c["features_idx_dict"][f"{ent}_masked_{mod}_idx"]
                out = out.clone()
                out[:, idx] = out[:, idx] + e[:, idx]
        return torch.exp(log_lib) * torch.softmax(out, dim=-1)

    # ------------------------------------------------------------------ gp weights / active mask
    def get_gp_weights(self, gp_type="all", modality="rna"):
        """modules/vgpgae.py:1146-1202 ('masked' is what get_active_gp_mask passes for the
        prior block)."""
        p = self.p
        tw = p[f"target_{modality}_decoder.nb_means_normalized_decoder.masked_l.weight"].detach()
        sw = p[f"source_{modality}_decoder.nb_means_normalized_decoder.masked_l.weight"].detach()
        w = torch.cat((tw, sw), dim=0)
        if self.cfg["n_addon_gp"] > 0 and gp_type != "masked":
            ta = p[f"target_{modality}_decoder.nb_means_normalized_decoder.addon_l.weight"].detach()
            sa = p[f"source_{modality}_decoder.nb_means_normalized_decoder.addon_l.weight"].detach()
            wa = torch.cat((ta, sa), dim=0)
            if gp_type == "all":
                w = torch.cat([w, wa], dim=1)
            elif gp_type == "addon":
                w = wa
        return w

    def get_active_gp_mask(self):
        """modules/vgpgae.py:1205-1334, mode 'sum+nzmeans'."""
        c = self.cfg
        Pp, Pa = c["n_prior_gp"], c["n_addon_gp"]
        active = torch.zeros(Pp + Pa, dtype=torch.bool)
        if c.get("active_gp_type", "separate") == "mixed":
            types = [("all", slice(None, None))]
        elif Pa > 0:
            types = [("masked", slice(None, Pp)), ("addon", slice(Pp, None))]
        else:
            types = [("masked", slice(None, Pp))]
        ratio = c["active_gp_thresh_ratio"]
        for gp_type, sl in types:
            w = self.running_mean_abs_mu[sl] * self.get_gp_weights(gp_type)
            sums = w.abs().sum(0)
            active[sl] |= sums >= ratio * sums.max()
            nzmeans = torch.nan_to_num(sums / torch.count_nonzero(w, dim=0))
            active[sl] |= nzmeans >= ratio * nzmeans.max()
        return active

    # ------------------------------------------------------------------ forward
    def forward(self, batch, decoder: str, use_only_active_gps=False, eps=None,
                update_atac_dynamic_decoder_mask=False):
        """modules/vgpgae.py:541-930.  ``eps``: the N(0,1) draw of reparameterize
        ([len(batch_idx), P]); required in training mode."""
        c = self.cfg
        x, edge_index = batch.x, batch.edge_index
        if decoder == "omics":
            batch_idx = torch.arange(batch.batch_size)
        else:
            batch_idx = torch.cat((batch.edge_label_index[0], batch.edge_label_index[1]), 0)
        x_enc = torch.log(1 + x) if c.get("log_variational", True) else x
        embed = self._embed(getattr(batch, "cat_covariates_cats", None))
        inj = c.get("cat_covariates_embeds_injection", [])
        mu_all, logstd_all = self.encoder(x_enc, edge_index, embed if "encoder" in inj else None)
        mu, logstd = mu_all[batch_idx], logstd_all[batch_idx]
        out = {"mu": mu, "logstd": logstd}
        z = mu + eps * torch.exp(logstd) if self.training else mu
        active = None
        if use_only_active_gps:
            active = self.get_active_gp_mask()
            z = z.clone()
            z[:, ~active] = 0
            if not self.training:
                # in eval mode the reference's z IS self.mu (vgaemodulemixin.py:39-41), so the
                # in-place zeroing at vgpgae.py:637 also lands in output["mu"]
                mu = z
                out["mu"] = z
        if decoder == "graph":
            out["edge_recon_labels"] = batch.edge_label
            n_cov = len(c.get("cat_covariates_embeds_nums", []))
            if n_cov > 0:
                cats = batch.cat_covariates_cats
                same = [cats[batch.edge_label_index[0], i] == cats[batch.edge_label_index[1], i]
                        for i in range(n_cov)]
                incl = [s if no_edge else torch.ones_like(s)
                        for s, no_edge in zip(same, c["cat_covariates_no_edges"])]
                out["edge_same_cat_covariates_cat"] = same
                out["edge_incl"] = torch.all(torch.stack(incl), dim=0)
            else:
                out["edge_same_cat_covariates_cat"] = None
                out["edge_incl"] = None
            half = z.shape[0] // 2
            out["edge_recon_logits"] = cosine_similarity_rows(z[:half], z[half:])
            return out

        # ---- omics
        with torch.no_grad():
            if self.training:
                self.running_mean_abs_mu = (0.1 * mu.detach().abs().sum(0) / mu.size(0)
                                            + 0.9 * self.running_mean_abs_mu)
            if use_only_active_gps:
                self.running_mean_abs_mu[~active] = 0
                for k in self.dynamic_mask:
                    self.dynamic_mask[k][~active, :] = 0
        Fg = c["n_output_genes"]
        x_rna, x_atac = x[:, :Fg], x[:, Fg:]
        emb_b = embed[batch_idx] if embed is not None else None
        out["node_labels"] = {}
        self.theta_idx = {}
        for mod, xm, inj_key in (("rna", x_rna, "gene_expr_decoder"),
                                 ("atac", x_atac, "chrom_access_decoder")):
            if mod not in self.modalities:
                continue
            x_nbr, _ = self.aggregate(xm, edge_index, mod)
            if mod == "atac" and update_atac_dynamic_decoder_mask:
                self._update_atac_dynamic_mask()
            for ent, lab_src in (("target", xm), ("source", x_nbr)):
                key = f"{ent}_{mod}"
                static = c[f"{key}_mask"].to(torch.float64)
                if c["n_addon_gp"] > 0:
                    static = torch.cat((static, self.addon_mask[key].to(torch.float64)), dim=0)
                n_gps_per_feat = (static * self.dynamic_mask[key]).sum(0)
                ridx = torch.nonzero(n_gps_per_feat).flatten()
                self.theta_idx[key] = ridx
                labels = lab_src[batch_idx][:, ridx]
                out["node_labels"][key] = labels
                log_lib = torch.log(labels.sum(1, keepdim=True))
                means = self.decoder(key, z, log_lib,
                                     emb_b if (emb_b is not None and inj_key in inj) else None,
                                     self.dynamic_mask[key] if mod == "atac" else None)
                out[f"{key}_nb_means"] = means[:, ridx]
        return out

    def get_omics_decoder_outputs(self, batch, only_active_gps=True, eps=None):
        """modules/vgpgae.py:1439-1620: NB means over ALL output features (no column selection), library size =
        row sum of the node labels over all features; z = mu in eval mode."""
        c = self.cfg
        x, edge_index = batch.x, batch.edge_index
        nb = batch.batch_size
        x_enc = torch.log(1 + x) if c.get("log_variational", True) else x
        embed = self._embed(getattr(batch, "cat_covariates_cats", None))
        inj = c.get("cat_covariates_embeds_injection", [])
        mu_all, logstd_all = self.encoder(x_enc, edge_index, embed if "encoder" in inj else None)
        mu, logstd = mu_all[:nb], logstd_all[:nb]
        z = mu + eps * torch.exp(logstd) if self.training else mu.clone()
        if only_active_gps:
            z[:, ~self.get_active_gp_mask()] = 0
        Fg = c["n_output_genes"]
        out = {"node_labels": {}}
        for mod, xm, inj_key in (("rna", x[:, :Fg], "gene_expr_decoder"), ("atac", x[:, Fg:], "chrom_access_decoder")):
            if mod not in self.modalities:
                continue
            x_nbr, _ = self.aggregate(xm, edge_index, mod)
            for ent, lab in (("target", xm[:nb]), ("source", x_nbr[:nb])):
                key = f"{ent}_{mod}"
                out["node_labels"][key] = lab
                e = embed[:nb] if (embed is not None and inj_key in inj) else None
                out[f"{key}_nb_means"] = self.decoder(key, z, torch.log(lab.sum(1, keepdim=True)), e,
                                                      self.dynamic_mask[key] if mod == "atac" else None)
        return out

    def _update_atac_dynamic_mask(self):
        """modules/vgpgae.py:813-857."""
        c = self.cfg
        Fg = c["n_output_genes"]
        w = torch.round(self.get_gp_weights("all"), decimals=4)
        nz = w != 0
        gp = c["gene_peaks_mask"].to(torch.float32)
        for ent, blk in (("target", nz[:Fg]), ("source", nz[Fg:])):
            m = (blk.t().to(torch.float32) @ gp) != 0
            self.dynamic_mask[f"{ent}_atac"] = self.dynamic_mask[f"{ent}_atac"] & m

    # ------------------------------------------------------------------ loss
    def loss(self, edge_out, node_out, lambda_l1_masked=0., l1_targets_mask=None,
             l1_sources_mask=None, lambda_l1_addon=0., lambda_group_lasso=0.,
             lambda_gene_expr_recon=300., lambda_chrom_access_recon=100.,
             lambda_edge_recon=500000., lambda_cat_covariates_contrastive=0.,
             contrastive_logits_pos_ratio=0., contrastive_logits_neg_ratio=0.,
             edge_recon_active=True, cat_covariates_contrastive_active=True):
        """modules/vgpgae.py:932-1143."""
        c, p = self.cfg, self.p
        L = {}
        L["kl_reg_loss"] = kl_reg_loss(node_out["mu"], node_out["logstd"])
        if c.get("include_edge_kl_loss", True):
            L["kl_reg_loss"] = L["kl_reg_loss"] + kl_reg_loss(edge_out["mu"], edge_out["logstd"])
        L["edge_recon_loss"] = lambda_edge_recon * edge_recon_loss(
            edge_out["edge_recon_logits"], edge_out["edge_recon_labels"], edge_out["edge_incl"])
        if edge_out["edge_same_cat_covariates_cat"] is not None and lambda_cat_covariates_contrastive > 0:
            L["cat_covariates_contrastive_loss"] = lambda_cat_covariates_contrastive * contrastive_loss(
                edge_out["edge_recon_logits"], edge_out["edge_recon_labels"],
                edge_out["edge_same_cat_covariates_cat"],
                contrastive_logits_pos_ratio, contrastive_logits_neg_ratio)

        def nb(key):
            theta = torch.exp(p[f"{key}_theta"][self.theta_idx[key]])
            return nb_recon_loss(node_out["node_labels"][key], node_out[f"{key}_nb_means"], theta)

        L["gene_expr_recon_loss"] = lambda_gene_expr_recon * nb("target_rna")
        L["gene_expr_recon_loss"] = L["gene_expr_recon_loss"] + lambda_gene_expr_recon * nb("source_rna")

        def dec_w(ent, layer):
            return p[f"{ent}_rna_decoder.nb_means_normalized_decoder.{layer}.weight"]

        # L1 on prior-GP weights (modules/losses.py:287-339); masks are [F, P_prior] bool or None
        l1 = 0.
        for ent, m in (("target", l1_targets_mask), ("source", l1_sources_mask)):
            w = dec_w(ent, "masked_l")
            l1 = l1 + (w[m].abs().sum() if m is not None else w.abs().sum())
        L["masked_gp_l1_reg_loss"] = lambda_l1_masked * l1
        # group lasso (modules/losses.py:215-252)
        gl = 0.
        for ent in ("target", "source"):
            w = dec_w(ent, "masked_l")
            m = p[f"{ent}_rna_decoder.nb_means_normalized_decoder.masked_l.mask"]
            gl = gl + (torch.sqrt(torch.count_nonzero(m, dim=0))      # int64 -> fp32 sqrt, as the reference
                       * torch.linalg.vector_norm(w, ord=2, dim=0)).sum()
        L["group_lasso_reg_loss"] = lambda_group_lasso * gl
        if c["n_addon_gp"] != 0:
            L["addon_gp_l1_reg_loss"] = lambda_l1_addon * (
                dec_w("target", "addon_l").abs().sum() + dec_w("source", "addon_l").abs().sum())
        if "atac" in self.modalities:
            L["chrom_access_recon_loss"] = lambda_chrom_access_recon * nb("target_atac")
            L["chrom_access_recon_loss"] = (L["chrom_access_recon_loss"]
                                            + lambda_chrom_access_recon * nb("source_atac"))
        g = L["kl_reg_loss"]
        o = L["kl_reg_loss"]
        if c.get("include_edge_recon_loss", True):
            g = g + L["edge_recon_loss"]
            if edge_recon_active:
                o = o + L["edge_recon_loss"]
        if c.get("include_cat_covariates_contrastive_loss", False) and "cat_covariates_contrastive_loss" in L:
            g = g + L["cat_covariates_contrastive_loss"]
            if cat_covariates_contrastive_active:
                o = o + L["cat_covariates_contrastive_loss"]
        if c.get("include_gene_expr_recon_loss", True):
            for k in ("gene_expr_recon_loss", "group_lasso_reg_loss", "masked_gp_l1_reg_loss"):
                g = g + L[k]
                o = o + L[k]
            if c["n_addon_gp"] != 0:
                g = g + L["addon_gp_l1_reg_loss"]
                o = o + L["addon_gp_l1_reg_loss"]
        if "atac" in self.modalities and c.get("include_chrom_access_recon_loss", True):
            g = g + L["chrom_access_recon_loss"]
            o = o + L["chrom_access_recon_loss"]
        L["global_loss"], L["optim_loss"] = g, o
        return L

    # ------------------------------------------------------------------ latent inference
    def get_latent_representation(self, batch, only_active_gps=True, return_mu_std=False, eps=None):
        """modules/vgpgae.py:1368-1437."""
        c = self.cfg
        x_enc = torch.log(1 + batch.x) if c.get("log_variational", True) else batch.x
        embed = self._embed(getattr(batch, "cat_covariates_cats", None))
        inj = c.get("cat_covariates_embeds_injection", [])
        mu, logstd = self.encoder(x_enc, batch.edge_index, embed if "encoder" in inj else None)
        mu, logstd = mu[:batch.batch_size], logstd[:batch.batch_size]
        if only_active_gps:
            a = self.get_active_gp_mask()
            mu, logstd = mu[:, a], logstd[:, a]
        if return_mu_std:
            return mu, torch.exp(logstd)
        return mu + eps * torch.exp(logstd) if self.training else mu


# ----------------------------------------------------------------------------- param init
def init_params(cfg: dict, seed: int = 0, dtype=torch.float32) -> Dict[str, torch.Tensor]:
    """Random parameters with the reference's shapes / state_dict keys and
    initialisers of the same family (kaiming-uniform Linear, glorot PyG Linear / att, zero
    biases, N(0,1) theta and embeddings, decoder weights zeroed by their mask
    -- nn/layercomponents.py:51)."""
    g = torch.Generator().manual_seed(seed)
    F_in, Hd = cfg["n_input"], cfg["n_hidden"]
    Pp, Pa, H = cfg["n_prior_gp"], cfg["n_addon_gp"], cfg.get("heads", 4)
    inj = cfg.get("cat_covariates_embeds_injection", [])
    n_emb = sum(cfg.get("cat_covariates_embeds_nums", []))
    p: Dict[str, torch.Tensor] = {}

    def uni(shape, a):
        return (torch.rand(shape, generator=g, dtype=torch.float64) * 2 - 1).mul(a).to(dtype)

    def glorot(shape):
        return uni(shape, (6.0 / (shape[-2] + shape[-1])) ** 0.5)

    enc_in = F_in + (n_emb if "encoder" in inj else 0)
    if cfg.get("n_fc_layers", 1) == 2:                       # nn/encoders.py:93-96
        mid = int(enc_in / 2)
        p["encoder.fc_l1.weight"] = uni((mid, enc_in), 1 / math.sqrt(enc_in))
        p["encoder.fc_l1.bias"] = uni((mid,), 1 / math.sqrt(enc_in))
        p["encoder.fc_l2.weight"] = uni((Hd, mid), 1 / math.sqrt(mid))
        p["encoder.fc_l2.bias"] = uni((Hd,), 1 / math.sqrt(mid))
        p["encoder.fc_l2_bn.weight"] = 1 + uni((Hd,), 0.2)         # off the 1 / 0 defaults to exercise both
        p["encoder.fc_l2_bn.bias"] = uni((Hd,), 0.1)
        p["encoder.fc_l2_bn.running_mean"] = torch.zeros(Hd, dtype=dtype)
        p["encoder.fc_l2_bn.running_var"] = torch.ones(Hd, dtype=dtype)
        p["encoder.fc_l2_bn.num_batches_tracked"] = torch.zeros((), dtype=torch.int64)
    else:
        p["encoder.fc_l1.weight"] = uni((Hd, enc_in), 1 / math.sqrt(enc_in))
        p["encoder.fc_l1.bias"] = uni((Hd,), 1 / math.sqrt(enc_in))
    convs = [("conv_mu", Pp), ("conv_logstd", Pp)]
    if cfg.get("n_layers", 1) == 2:
        convs.insert(0, ("conv_l1", Hd))
    if Pa:
        convs += [("addon_conv_mu", Pa), ("addon_conv_logstd", Pa)]
    for name, C in convs:
        if cfg["conv_layer"] == "gatv2conv":
            p[f"encoder.{name}.att"] = glorot((1, H, C))
            for side in ("lin_l", "lin_r"):
                p[f"encoder.{name}.{side}.weight"] = glorot((H * C, Hd))
                p[f"encoder.{name}.{side}.bias"] = uni((H * C,), 0.1)   # non-zero to exercise bias paths
            p[f"encoder.{name}.bias"] = uni((C,), 0.1)
        else:
            p[f"encoder.{name}.lin.weight"] = glorot((C, Hd))
            p[f"encoder.{name}.bias"] = uni((C,), 0.1)
    idx = cfg["features_idx_dict"]
    mods = [("rna", cfg["n_output_genes"], "gene_expr_decoder")]
    if cfg.get("n_output_peaks", 0) > 0:
        mods.append(("atac", cfg["n_output_peaks"], "chrom_access_decoder"))
    for mod, Fo, inj_key in mods:
        for ent in ("target", "source"):
            pre = f"{ent}_{mod}_decoder.nb_means_normalized_decoder"
            m = cfg[f"{ent}_{mod}_mask"].t().to(dtype)                 # [Fo, Pp]
            p[f"{pre}.masked_l.mask"] = m
            p[f"{pre}.masked_l.weight"] = uni((Fo, Pp), 1 / math.sqrt(max(Pp, 1))) * m
            if Pa:
                am = torch.zeros(Pa, Fo, dtype=dtype)
                if mod == "rna":
                    am[:, idx[f"{ent}_unmasked_rna_idx"]] = 1.
                else:
                    rm = torch.zeros(Pa, cfg["n_output_genes"], dtype=torch.int64)
                    rm[:, idx[f"{ent}_unmasked_rna_idx"]] = 1
                    am = ((rm @ cfg["gene_peaks_mask"].to(torch.int64)) > 0).to(dtype)
                p[f"{pre}.addon_l.mask"] = am.t().contiguous()
                p[f"{pre}.addon_l.weight"] = uni((Fo, Pa), 1 / math.sqrt(Pa)) * am.t()
            if n_emb and inj_key in inj:
                p[f"{pre}.cat_covariates_embed_l.weight"] = uni((Fo, n_emb), 1 / math.sqrt(n_emb))
            p[f"{ent}_{mod}_theta"] = torch.randn(Fo, generator=g, dtype=torch.float64).to(dtype)
    for i, (ncat, nemb) in enumerate(zip(cfg.get("cat_covariates_num_cats", []),
                                         cfg.get("cat_covariates_embeds_nums", []))):
        p[f"cat_covariate{i}_embedder.weight"] = torch.randn(ncat, nemb, generator=g,
                                                             dtype=torch.float64).to(dtype)
    if cfg["node_label_method"] == "one-hop-attention":
        Fg, Ha = cfg["n_input"], cfg.get("aggregator_heads", 4)
        p["rna_node_label_aggregator.linear_l_l.weight"] = glorot((Ha * Fg, Fg))
        p["rna_node_label_aggregator.linear_r_l.weight"] = glorot((Ha * Fg, Fg))
        p["rna_node_label_aggregator.attn"] = glorot((1, Ha, Fg))
    p["running_mean_abs_mu"] = torch.zeros(Pp + Pa, dtype=dtype)
    return p


NON_TRAINABLE_SUFFIXES = (".mask", "running_mean_abs_mu", ".running_mean", ".running_var", ".num_batches_tracked")


def trainable_keys(params: Dict[str, torch.Tensor]) -> List[str]:
    return [k for k in params if not k.endswith(NON_TRAINABLE_SUFFIXES)]
