"""Mirror of ``nichecompass.nn`` (reference: nn/__init__.py:11-19); ``FCOmicsFeatureDecoder`` is an
ablation-only module of the reference and is not part of the accelerated path."""
from .aggregators import (OneHopAttentionNodeLabelAggregator,
                          OneHopGCNNormNodeLabelAggregator,
                          OneHopSumNodeLabelAggregator)
from .conv import GATv2Conv, GCNConv
from .decoders import CosineSimGraphDecoder, MaskedOmicsFeatureDecoder
from .encoders import Encoder
from .layercomponents import MaskedLinear
from .layers import AddOnMaskedLayer

__all__ = ["OneHopAttentionNodeLabelAggregator",
           "OneHopGCNNormNodeLabelAggregator",
           "OneHopSumNodeLabelAggregator",
           "CosineSimGraphDecoder",
           "MaskedOmicsFeatureDecoder",
           "Encoder",
           "MaskedLinear",
           "AddOnMaskedLayer",
           "GATv2Conv",
           "GCNConv"]
