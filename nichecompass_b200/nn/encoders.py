"""
Encoder of the VGPGAE (reference: nn/encoders.py:12-208).  Same constructor, parameters and
state_dict keys; the arithmetic is one fused autograd Function over hand-written kernels
(tcgen05 / SIMT GEMMs + the fused GATv2 edge kernels or the CSR SpMM for GCN).
"""
from typing import Literal, Optional, Tuple

import torch
import torch.nn as nn

from ..functional import HIDDEN_INPUT, EncoderGATv2Fn, EncoderGCNFn, LinearActFn, SubgraphCSR, new_cache_key
from ..ops import pad8
from .conv import GATv2Conv, GCNConv


class Encoder(nn.Module):
    """fc layer -> ReLU -> message passing heads producing ``mu`` and ``logstd`` of the latent
    normal (prior gene programs ++ add-on gene programs).

    Parameters are those of the reference ``Encoder``; ``use_bn`` is accepted and unused
    exactly as there (nn/encoders.py:52-53 documents it, the body never reads it).
    Kernel coverage: ``n_layers`` 1 or 2 (``conv_l1`` over the whole sampled sub-graph, ReLU, then the
    latent convs on the seed rows).  ``n_fc_layers == 1`` with ``dropout_rate == 0`` (the defaults) is ONE
    fused function; ``n_fc_layers == 2`` (fc_l2 + BatchNorm1d, nn/encoders.py:93-96,180-182) or dropout
    composes the same kernels layer by layer (``_encode_composed``), the convs reading the hidden tensor as is.
    """

    def __init__(self,
                 n_input: int,
                 n_cat_covariates_embed_input: int,
                 n_hidden: int,
                 n_latent: int,
                 n_addon_latent: int = 100,
                 n_fc_layers: int = 1,
                 conv_layer: Literal["gcnconv", "gatv2conv"] = "gatv2conv",
                 n_layers: int = 1,
                 cat_covariates_embed_mode: Literal["input", "hidden"] = "input",
                 n_attention_heads: int = 4,
                 dropout_rate: float = 0.,
                 activation=torch.relu,
                 use_bn: bool = True):
        super().__init__()
        print("ENCODER -> "
              f"n_input: {n_input}, n_cat_covariates_embed_input: {n_cat_covariates_embed_input}, "
              f"n_hidden: {n_hidden}, n_latent: {n_latent}, n_addon_latent: {n_addon_latent}, "
              f"n_fc_layers: {n_fc_layers}, n_layers: {n_layers}, conv_layer: {conv_layer}, "
              f"n_attention_heads: {n_attention_heads if conv_layer == 'gatv2conv' else '0'}, "
              f"dropout_rate: {dropout_rate}, ")
        if n_fc_layers not in (1, 2) or n_layers not in (1, 2):
            raise NotImplementedError("nichecompass_b200 kernels cover n_fc_layers in (1, 2) and n_layers in (1, 2)")
        if cat_covariates_embed_mode != "input" and n_cat_covariates_embed_input != 0:
            raise NotImplementedError("cat_covariates_embed_mode='hidden' is not implemented")
        self.n_latent = n_latent
        self.n_addon_latent = n_addon_latent
        self.n_layers = n_layers
        self.n_fc_layers = n_fc_layers
        self.conv_layer = conv_layer
        self.n_attention_heads = n_attention_heads
        self.cat_covariates_embed_mode = cat_covariates_embed_mode
        self.n_input = n_input + n_cat_covariates_embed_input
        if n_fc_layers == 2:                               # nn/encoders.py:93-96
            self.fc_l1 = nn.Linear(self.n_input, int(self.n_input / 2))
            self.fc_l2 = nn.Linear(int(self.n_input / 2), n_hidden)
            self.fc_l2_bn = nn.BatchNorm1d(n_hidden)
            self._key_fc = (new_cache_key(self), new_cache_key(self))
        else:
            self.fc_l1 = nn.Linear(self.n_input, n_hidden)
        self.dropout_rate = dropout_rate
        self.dropout = nn.Dropout(dropout_rate)
        conv = (lambda c: GATv2Conv(n_hidden, c, heads=n_attention_heads)) if conv_layer == "gatv2conv" \
            else (lambda c: GCNConv(n_hidden, c))
        if conv_layer not in ("gatv2conv", "gcnconv"):
            raise ValueError(f"unknown conv_layer {conv_layer}")
        if n_layers == 2:
            self.conv_l1 = conv(n_hidden)                  # nn/encoders.py:106-108 / 123-127
            self._key_l1 = new_cache_key(self)
        self.conv_mu = conv(n_latent)
        self.conv_logstd = conv(n_latent)
        if n_addon_latent != 0:
            self.addon_conv_mu = conv(n_addon_latent)
            self.addon_conv_logstd = conv(n_addon_latent)
        self.compute_dtype = torch.float32
        self._key = new_cache_key(self)

    # ------------------------------------------------------------------ fused path
    @property
    def n_gps(self) -> int:
        return self.n_latent + self.n_addon_latent

    def _convs(self):
        """Convs in output order with their output column offsets in the (mu | logstd) matrix."""
        P = self.n_gps
        convs = [(self.conv_mu, 0)]
        if self.n_addon_latent != 0:
            convs.append((self.addon_conv_mu, self.n_latent))
        convs.append((self.conv_logstd, P))
        if self.n_addon_latent != 0:
            convs.append((self.addon_conv_logstd, P + self.n_latent))
        return convs

    def arena_groups(self):
        """Parameters the fused kernels consume stacked: laid out back to back by ``FlatAdam`` so
        that the stacked matrices, their bf16 shadows and their gradients are views."""
        convs = [c for c, _ in self._convs()]
        if not self.fused:
            # composed path: fc weights are ordinary parameters of the arena, the conv stacks as below
            if self.conv_layer == "gatv2conv":
                return [[c.lin_l.weight for c in convs], [c.lin_r.weight for c in convs],
                        [c.lin_l.bias for c in convs], [c.lin_r.bias for c in convs],
                        [c.att for c in convs], [c.bias for c in convs]]
            return [[c.lin.weight for c in convs], [c.bias for c in convs]]
        if self.conv_layer == "gatv2conv":
            groups = [[c.lin_l.weight for c in convs], [c.lin_r.weight for c in convs],
                      [c.lin_l.bias for c in convs], [c.lin_r.bias for c in convs],
                      [c.att for c in convs], [c.bias for c in convs], [self.fc_l1.weight], [self.fc_l1.bias]]
            if self.n_layers == 2:
                groups += [[t] for t in self.conv_l1.param_tuple()]
            return groups
        groups = [[c.lin.weight for c in convs], [c.bias for c in convs], [self.fc_l1.weight], [self.fc_l1.bias]]
        if self.n_layers == 2:
            groups += [[self.conv_l1.lin.weight], [self.conv_l1.bias]]
        return groups

    @property
    def fused(self) -> bool:
        """fc_l1 + ReLU + convs in one autograd function (the default configuration)."""
        return self.n_fc_layers == 1 and self.dropout_rate == 0.

    def _encode_composed(self, x_enc: torch.Tensor, graph: SubgraphCSR) -> torch.Tensor:
        """nn/encoders.py:179-207 layer by layer: fc_l1 -> ReLU -> dropout [-> fc_l2 -> ReLU -> dropout ->
        BatchNorm1d] [-> conv_l1 -> ReLU -> dropout] -> latent convs.  GEMMs and message passing are the
        library's kernels; dropout and BatchNorm (batch statistics over the sampled sub-graph's nodes in
        training, running statistics in eval) are torch element-wise / reduction ops on [n_sub, Hd] fp32."""
        convs = self._convs()
        offs = [o for _, o in convs]
        act = self.compute_dtype
        gat = self.conv_layer == "gatv2conv"
        k1 = self._key_fc[0] if self.n_fc_layers == 2 else self._key
        h = self.dropout(LinearActFn.apply(x_enc, self.fc_l1.weight, self.fc_l1.bias, act, True, (k1, "fc1")))
        if self.n_fc_layers == 2:
            h = self.dropout(LinearActFn.apply(h, self.fc_l2.weight, self.fc_l2.bias, act, True,
                                               (self._key_fc[1], "fc2")))
            h = self.fc_l2_bn(h)
        Hd = h.size(1)
        if self.n_layers == 2:
            g_all = SubgraphCSR(graph.rowptr, graph.col, graph.n_nodes, graph.n_nodes, graph.max_deg)
            if gat:
                h = EncoderGATv2Fn.apply(h, g_all, Hd, self.n_attention_heads, [0], Hd, self._key_l1, act,
                                         HIDDEN_INPUT, None, *self.conv_l1.param_tuple())
            else:
                h = EncoderGCNFn.apply(h, g_all, Hd, [0], Hd, self._key_l1, act, HIDDEN_INPUT, None,
                                       self.conv_l1.lin.weight, self.conv_l1.bias)
            h = self.dropout(torch.relu(h))
        if gat:
            tensors = [t for c, _ in convs for t in c.param_tuple()]
            return EncoderGATv2Fn.apply(h, graph, Hd, self.n_attention_heads, offs, 2 * self.n_gps, self._key, act,
                                        HIDDEN_INPUT, None, *tensors)
        tensors = [t for c, _ in convs for t in (c.lin.weight, c.bias)]
        return EncoderGCNFn.apply(h, graph, Hd, offs, 2 * self.n_gps, self._key, act, HIDDEN_INPUT, None, *tensors)

    def encode(self, x_enc: torch.Tensor, graph: SubgraphCSR) -> torch.Tensor:
        """``x_enc``: [n_sub, pad8(n_input)] in ``compute_dtype`` (log1p counts ++ optional
        embedding, zero padded).  Returns (mu | logstd) [graph.n_dst, 2 * n_gps] fp32."""
        if not self.fused:
            return self._encode_composed(x_enc, graph)
        convs = self._convs()
        offs = [o for _, o in convs]
        act = self.compute_dtype
        gat = self.conv_layer == "gatv2conv"
        W1, b1, n_in, key = self.fc_l1.weight, self.fc_l1.bias, self.n_input, self._key
        if self.n_layers == 2:
            # hidden = relu(conv_l1(relu(fc_l1(x)))) for EVERY node of the sampled sub-graph (nn/encoders.py:191-194);
            # the latent convs then start from that fp32 tensor (``W1 = None``: relu + cast inside the function)
            Hd = self.fc_l1.weight.size(0)
            g_all = SubgraphCSR(graph.rowptr, graph.col, graph.n_nodes, graph.n_nodes, graph.max_deg)
            if gat:
                x_enc = EncoderGATv2Fn.apply(x_enc, g_all, n_in, self.n_attention_heads, [0], Hd, self._key_l1, act,
                                             W1, b1, *self.conv_l1.param_tuple())
            else:
                x_enc = EncoderGCNFn.apply(x_enc, g_all, n_in, [0], Hd, self._key_l1, act, W1, b1,
                                           self.conv_l1.lin.weight, self.conv_l1.bias)
            W1, b1, n_in = None, None, Hd
        if gat:
            tensors = [t for c, _ in convs for t in c.param_tuple()]
            return EncoderGATv2Fn.apply(x_enc, graph, n_in, self.n_attention_heads, offs, 2 * self.n_gps, key, act,
                                        W1, b1, *tensors)
        tensors = [t for c, _ in convs for t in (c.lin.weight, c.bias)]
        return EncoderGCNFn.apply(x_enc, graph, n_in, offs, 2 * self.n_gps, key, act, W1, b1, *tensors)

    def pack_input(self, x: torch.Tensor, cat_covariates_embed: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Dense fp32 [n, F] (already log-transformed) (+ embedding) -> padded compute-dtype tile."""
        n = x.size(0)
        buf = torch.zeros(n, pad8(self.n_input), dtype=self.compute_dtype, device=x.device)
        buf[:, :x.size(1)] = x
        if cat_covariates_embed is not None and self.cat_covariates_embed_mode == "input":
            if cat_covariates_embed.requires_grad:
                full = torch.cat((x, cat_covariates_embed), dim=1).to(self.compute_dtype)
                return torch.nn.functional.pad(full, (0, buf.size(1) - full.size(1)))
            buf[:, x.size(1):x.size(1) + cat_covariates_embed.size(1)] = cat_covariates_embed
        return buf

    # ------------------------------------------------------------------ reference signature
    def forward(self,
                x: torch.Tensor,
                edge_index: torch.Tensor,
                cat_covariates_embed: Optional[torch.Tensor] = None
                ) -> Tuple[torch.Tensor, torch.Tensor]:
        """Reference call signature (nn/encoders.py:144-208): returns ``mu`` and ``logstd`` for
        ALL nodes of the batch."""
        graph = SubgraphCSR.from_edge_index(edge_index, x.size(0))
        ml = self.encode(self.pack_input(x, cat_covariates_embed), graph)
        P = self.n_gps
        return ml[:, :P], ml[:, P:]
