"""``GraphormerModel`` / ``GraphormerEncoder`` registered as ``multi_graphormer`` with archs
``multi_graphormer`` and ``multi_graphormer_base`` (mDT/src/models/multi_modal_discussion_transformer.py:22-344).
Same ``args`` fields, parameter names (dead ones included, for checkpoint compatibility) and outputs."""

from __future__ import annotations

import logging

import torch
import torch.nn as nn

from .._registry import FairseqEncoder, FairseqEncoderModel, register_model, register_model_architecture
from ..modules import MultiGraphormerGraphEncoder, init_graphormer_params
from ..modules._fairseq_compat import LayerNorm, get_activation_fn

logger = logging.getLogger(__name__)

__all__ = ["GraphormerModel", "GraphormerEncoder", "base_architecture", "graphormer_base_architecture"]

# dataset-side sizes the task config supplies (tasks/task.py:29-113)
TASK_DEFAULTS = dict(num_atoms=512 * 9, num_in_degree=512, num_out_degree=512, num_edges=512 * 3, num_spatial=512,
                     num_edge_dis=128, edge_type="multi_hop", multi_hop_max_dist=5, num_classes=2, max_nodes=10000)


@register_model("multi_graphormer")
class GraphormerModel(FairseqEncoderModel):
    def __init__(self, args, encoder):
        super().__init__(encoder)
        self.args = args
        if getattr(args, "apply_graphormer_init", False):
            self.apply(init_graphormer_params)
        self.encoder_embed_dim = args.encoder_embed_dim

    @staticmethod
    def add_args(parser):
        """Same flags as the reference (multi_modal_discussion_transformer.py:32-158)."""
        for flag in ("--dropout", "--attention-dropout", "--act-dropout"):
            parser.add_argument(flag, type=float, metavar="D")
        for flag in ("--encoder-ffn-embed-dim", "--encoder-layers", "--encoder-attention-heads", "--num_fusion_layers",
                     "--num_graph_stack", "--num_fusion_stack", "--num_bottleneck_tokens", "--encoder-embed-dim", "--split",
                     "--max-positions"):
            parser.add_argument(flag, type=int, metavar="N")
        for flag in ("--share-encoder-input-output-embed", "--encoder-learned-pos", "--no-token-positional-embeddings",
                     "--apply-graphormer-init", "--encoder-normalize-before", "--pre-layernorm", "--freeze_initial_encoders"):
            parser.add_argument(flag, action="store_true")
        parser.add_argument("--activation-fn", choices=["relu", "gelu", "tanh", "linear"])

    def max_nodes(self):
        return self.encoder.max_nodes

    @classmethod
    def build_model(cls, args, task=None):
        base_architecture(args)
        if not hasattr(args, "max_nodes"):
            args.max_nodes = getattr(args, "tokens_per_sample", TASK_DEFAULTS["max_nodes"])
        return cls(args, GraphormerEncoder(args))

    def forward(self, batched_data, **kwargs):
        return self.encoder(batched_data, **kwargs)


class GraphormerEncoder(FairseqEncoder):
    def __init__(self, args):
        super().__init__(dictionary=None)
        for k, v in TASK_DEFAULTS.items():
            if not hasattr(args, k):
                setattr(args, k, v)
        self.max_nodes = args.max_nodes
        self.graph_encoder = MultiGraphormerGraphEncoder(
            num_atoms=args.num_atoms, num_in_degree=args.num_in_degree, num_out_degree=args.num_out_degree,
            num_edges=args.num_edges, num_spatial=args.num_spatial, num_edge_dis=args.num_edge_dis,
            edge_type=args.edge_type, multi_hop_max_dist=args.multi_hop_max_dist,
            num_bottle_neck=args.num_bottleneck_tokens, num_fusion_layers=args.num_fusion_layers,
            num_fusion_stack=args.num_fusion_stack, num_graph_stack=args.num_graph_stack,
            num_encoder_layers=args.encoder_layers, embedding_dim=args.encoder_embed_dim,
            ffn_embedding_dim=args.encoder_ffn_embed_dim, num_attention_heads=args.encoder_attention_heads,
            dropout=args.dropout, attention_dropout=args.attention_dropout, activation_dropout=args.act_dropout,
            encoder_normalize_before=args.encoder_normalize_before, pre_layernorm=getattr(args, "pre_layernorm", False),
            apply_graphormer_init=args.apply_graphormer_init, activation_fn=args.activation_fn,
            freeze_initial_encoders=getattr(args, "freeze_initial_encoders", False),
            bert_config=getattr(args, "bert_config", None), vit_config=getattr(args, "vit_config", None))
        ge = self.graph_encoder
        self.node_encoder_stack = nn.ModuleList([ge.text_pooler, ge.text_dropout, ge.node_classifier])      # :215-221
        ge.text_first_token_only = type(ge.text_pooler).__name__ == "BertPooler"   # the head reads hidden_states[:, 0] only
        self.share_input_output_embed = getattr(args, "share_encoder_input_output_embed", False)
        self.load_softmax = not getattr(args, "remove_head", False)
        # unused heads kept for state-dict compatibility (:231-247)
        self.masked_lm_pooler = nn.Linear(args.encoder_embed_dim, args.encoder_embed_dim)
        self.lm_head_transform_weight = nn.Linear(args.encoder_embed_dim, args.encoder_embed_dim)
        self.activation_fn = get_activation_fn(args.activation_fn)
        self.layer_norm = LayerNorm(args.encoder_embed_dim)
        self.embed_out = None
        self.lm_output_learned_bias = None
        if self.load_softmax:
            self.lm_output_learned_bias = nn.Parameter(torch.zeros(1))
            if self.share_input_output_embed:
                raise NotImplementedError
            self.embed_out = nn.Linear(args.encoder_embed_dim, args.num_classes, bias=False)

    def forward(self, batched_data, **unused):
        """-> ``(node logits [n_real, 2], global embedding [B, 768])`` (:256-276)."""
        text, graph, global_embedding = self.graph_encoder(batched_data)
        for layer in self.node_encoder_stack:   # pooler (token 0) -> dropout -> classifier, on both streams
            text = layer(text)
            graph = layer(graph)
        return (text + graph) / 2, global_embedding

    def upgrade_state_dict_named(self, state_dict, name):
        if not self.load_softmax:
            for k in list(state_dict.keys()):
                if "embed_out.weight" in k or "lm_output_learned_bias" in k:
                    del state_dict[k]
        return state_dict


def _default(args, name, value):
    setattr(args, name, getattr(args, name, value))


@register_model_architecture("multi_graphormer", "multi_graphormer")
def base_architecture(args):
    """Defaults of multi_modal_discussion_transformer.py:290-317."""
    for name, value in (("dropout", 0.1), ("attention_dropout", 0.1), ("act_dropout", 0.0), ("encoder_ffn_embed_dim", 4096),
                        ("encoder_layers", 6), ("encoder_attention_heads", 8), ("split", 0), ("encoder_embed_dim", 1024),
                        ("share_encoder_input_output_embed", False), ("no_token_positional_embeddings", False),
                        ("num_bottleneck_tokens", 4), ("num_fusion_layers", 4), ("num_graph_stack", 1),
                        ("num_fusion_stack", 1), ("apply_graphormer_init", False), ("activation_fn", "gelu"),
                        ("encoder_normalize_before", True)):
        _default(args, name, value)


@register_model_architecture("multi_graphormer", "multi_graphormer_base")
def graphormer_base_architecture(args):
    """Defaults of multi_modal_discussion_transformer.py:320-344."""
    for name, value in (("encoder_embed_dim", 80), ("split", 0), ("encoder_layers", 12), ("encoder_attention_heads", 8),
                        ("encoder_ffn_embed_dim", 80), ("activation_fn", "gelu"), ("freeze_initial_encoders", False),
                        ("encoder_normalize_before", True), ("apply_graphormer_init", False),
                        ("share_encoder_input_output_embed", False), ("no_token_positional_embeddings", False),
                        ("pre_layernorm", False)):
        _default(args, name, value)
    base_architecture(args)
