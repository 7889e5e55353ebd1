"""Mirror of ``mDT/src/modules/__init__.py:1-7`` (same exported names)."""
from .graphormer_graph_encoder_layer import GraphEncoderStack, GraphormerGraphEncoderLayer
from .graphormer_layers import FusedAttnBias, GraphAttnBias, GraphNodeFeature, init_params
from .multi_graphormer_fusion_layer import GraphFusionLayer, GraphFusionStack
from .multigraphormer_graph_encoder import MultiGraphormerGraphEncoder, init_graphormer_params
from .multihead_attention import MultiheadAttention

__all__ = ["MultiheadAttention", "GraphNodeFeature", "GraphAttnBias", "GraphormerGraphEncoderLayer", "GraphEncoderStack",
           "GraphFusionLayer", "GraphFusionStack", "MultiGraphormerGraphEncoder", "init_graphormer_params", "init_params",
           "FusedAttnBias"]
