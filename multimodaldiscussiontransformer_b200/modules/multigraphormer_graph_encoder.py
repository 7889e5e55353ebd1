"""``MultiGraphormerGraphEncoder`` (mDT/src/modules/multigraphormer_graph_encoder.py:42-464): graph
transformer stacks over the discussion tree interleaved with BERT / ViT fusion stacks through the
bottleneck tokens.  Same constructor signature, parameter names and outputs as the reference; the
graph half runs on libmdt_b200 (node features, attention bias, biased attention, row moves) and
no step of the forward synchronises with the host when the batch comes from this package's
collator (``node_index`` / ``bias_code`` / ``x_nodes`` / ``x_image_rows`` keys)."""

from __future__ import annotations

import logging
from typing import Optional, Tuple

import torch
import torch.nn as nn

from .. import ops
from ._fairseq_compat import FairseqDropout, LayerNorm
from .graphormer_graph_encoder_layer import GraphEncoderStack
from .graphormer_layers import GraphAttnBias, GraphNodeFeature
from .multi_graphormer_fusion_layer import (GraphFusionStack, TextPack, image_rows, packed_bert_layer, use_fused_intermediates, use_fused_vit_layers, use_patch_gemm,
                                            use_fused_layer_norms, use_mdt_attention)
from .multihead_attention import MultiheadAttention

logger = logging.getLogger(__name__)


def init_graphormer_params(module):
    """multigraphormer_graph_encoder.py:18-39: N(0, 0.02) weights drawn on the CPU, zero biases / padding rows."""

    def normal_(data):
        data.copy_(data.cpu().normal_(mean=0.0, std=0.02).to(data.device))

    if isinstance(module, nn.Linear):
        normal_(module.weight.data)
        if module.bias is not None:
            module.bias.data.zero_()
    if isinstance(module, nn.Embedding):
        normal_(module.weight.data)
        if module.padding_idx is not None:
            module.weight.data[module.padding_idx].zero_()
    if isinstance(module, MultiheadAttention):
        normal_(module.q_proj.weight.data)
        normal_(module.k_proj.weight.data)
        normal_(module.v_proj.weight.data)


def _chunks(xs, k):
    return [xs[i * k:(i + 1) * k] for i in range((len(xs) + k - 1) // k)]


class MultiGraphormerGraphEncoder(nn.Module):
    #: use the in-kernel bias synthesis (code + lut) when the batch carries ``bias_code``
    fused_bias = True
    #: run the BERT side of the fusion layers on packed (unpadded) tokens when the batch carries ``x_text_pack`` and
    #: the model computes in bf16/fp16; hidden states at padding positions come back as zeros
    pack_text = False
    #: try the local Hugging Face cache for the pretrained trunks before falling back to config init
    pretrained = False

    def __init__(self, num_atoms: int, num_in_degree: int, num_out_degree: int, num_edges: int, num_spatial: int,
                 num_edge_dis: int, num_bottle_neck: int, num_fusion_layers: int, edge_type: str,
                 multi_hop_max_dist: int, num_fusion_stack: int = 1, num_graph_stack: int = 1,
                 num_encoder_layers: int = 12, embedding_dim: int = 768, ffn_embedding_dim: int = 768,
                 num_attention_heads: int = 32, dropout: float = 0.1, attention_dropout: float = 0.1,
                 activation_dropout: float = 0.1, layerdrop: float = 0.0, encoder_normalize_before: bool = False,
                 pre_layernorm: bool = False, apply_graphormer_init: bool = False, activation_fn: str = "gelu",
                 embed_scale: float = None, freeze_embeddings: bool = False, n_trans_layers_to_freeze: int = 0,
                 export: bool = False, traceable: bool = False, q_noise: float = 0.0, qn_block_size: int = 8,
                 freeze_initial_encoders: bool = False, bert_config=None, vit_config=None) -> None:
        super().__init__()
        self.dropout_module = FairseqDropout(dropout, module_name=self.__class__.__name__)
        self.layerdrop = layerdrop
        self.embedding_dim = embedding_dim
        self.apply_graphormer_init = apply_graphormer_init
        self.traceable = traceable
        num_encoder_layers = num_fusion_layers                                     # :87 (overrides the argument)
        self.graph_node_feature = GraphNodeFeature(
            num_heads=num_attention_heads, num_atoms=num_atoms, num_in_degree=num_in_degree,
            num_out_degree=num_out_degree, hidden_dim=embedding_dim, n_layers=num_encoder_layers)
        self.graph_attn_bias = GraphAttnBias(
            num_heads=num_attention_heads, num_atoms=num_atoms, num_edges=num_edges, num_spatial=num_spatial,
            num_edge_dis=num_edge_dis, edge_type=edge_type, multi_hop_max_dist=multi_hop_max_dist,
            hidden_dim=embedding_dim, n_layers=num_encoder_layers)
        self.embed_scale = embed_scale
        if q_noise > 0:
            raise NotImplementedError("quant-noise (q_noise > 0) is outside the B200 hot path")
        self.quant_noise = None
        self.emb_layer_norm = LayerNorm(self.embedding_dim, export=export) if encoder_normalize_before else None
        if pre_layernorm:
            self.final_layer_norm = LayerNorm(self.embedding_dim, export=export)
        if self.layerdrop > 0.0:
            raise NotImplementedError("layerdrop > 0 is outside the B200 hot path (the published run uses 0)")
        self.layers = nn.ModuleList([])
        (self.vit_model, self.vit_pooler, vit_other_layers, self.text_model, self.text_pooler, text_other_layers,
         self.node_classifier, self.text_dropout) = self.build_vit_bert_encoders(
            num_fusion_layers + 1, attention_dropout, activation_dropout, bert_config, vit_config)
        use_fused_layer_norms(self.vit_model)  # the trunk ViT blocks' LayerNorms on csrc/layernorm.cu as well
        use_mdt_attention(self.vit_model)      # ... and their self-attention on the tcgen05 kernels
        #: packed text only: return [n_real, 1, 768] (token 0 of every comment, all a pooling head reads) instead of the
        #: re-padded [n_real, 100, 768]; GraphormerEncoder switches it on, a caller that wants every token leaves it off
        self.text_first_token_only = False
        use_fused_intermediates(self.vit_model)  # ... and dense -> GELU on csrc/gelu.cu
        use_fused_vit_layers(self.vit_model)     # ... and each block as one node (residual adds / dropouts fused)
        use_patch_gemm(self.vit_model)           # ... and the patch embedding as the GEMM it is
        use_fused_layer_norms(self.text_model)   # the BERT embeddings' LayerNorm (the blocks' LayerNorms are fused with their tails)
        use_fused_intermediates(self.text_model)
        self.fusion_layers = nn.ModuleList([
            GraphFusionStack(t, v, num_bottle_neck, use_projection=True)
            for t, v in zip(_chunks(list(text_other_layers), num_fusion_stack), _chunks(list(vit_other_layers), num_fusion_stack))])
        # one more graph stack than fusion stacks is built; stack len(fusion_layers)-1 is never run (:173-191, :413-446)
        self.layers.extend([
            self.build_graphormer_graph_encoder_layer(
                num_layers=num_graph_stack, embedding_dim=self.embedding_dim, ffn_embedding_dim=ffn_embedding_dim,
                num_attention_heads=num_attention_heads, dropout=self.dropout_module.p,
                attention_dropout=attention_dropout, activation_dropout=activation_dropout, activation_fn=activation_fn,
                export=export, q_noise=q_noise, qn_block_size=qn_block_size, pre_layernorm=pre_layernorm)
            for _ in range(len(self.fusion_layers) + 1)])
        logger.info("graph stacks: %d (x%d layers), fusion stacks: %d (x%d layers)", len(self.layers), num_graph_stack,
                    len(self.fusion_layers), num_fusion_stack)
        self.num_bottle_neck = num_bottle_neck
        self.bottle_neck = nn.Embedding(num_bottle_neck, 768)

        def set_requires_grad(m, flag):
            if m is not None:
                for p in m.parameters():
                    p.requires_grad = flag

        if freeze_embeddings:
            raise NotImplementedError("Freezing embeddings is not implemented yet.")
        if freeze_initial_encoders:                                               # :223-228
            set_requires_grad(self.text_model, False)
            set_requires_grad(self.vit_model, False)
            set_requires_grad(self.node_classifier, True)
            set_requires_grad(self.text_pooler, True)
            set_requires_grad(self.vit_pooler, True)
        for layer in range(n_trans_layers_to_freeze):
            set_requires_grad(self.layers[layer], False)

    # ------------------------------------------------------------------------------------------
    def build_vit_bert_encoders(self, num_fusion_layers, attention_dropout, activation_dropout, bert_config=None,
                                vit_config=None):
        """BERT-base / ViT-base trunks with their last ``num_fusion_layers`` layers peeled off (:233-278).

        The reference downloads ``bert-base-uncased`` and ``google/vit-base-patch16-224``; without
        network access the same architectures are built from their default configs (random init) and
        a checkpoint supplies the weights (parameter names are identical)."""
        from transformers import AutoModel, AutoModelForSequenceClassification, BertConfig, BertForSequenceClassification
        from transformers import ViTConfig, ViTModel

        kw = dict(hidden_dropout_prob=activation_dropout, attention_probs_dropout_prob=attention_dropout)
        vit_model = bert = None
        if self.pretrained and bert_config is None and vit_config is None:
            try:
                vit_model = AutoModel.from_pretrained("google/vit-base-patch16-224", local_files_only=True, **kw)
                bert = AutoModelForSequenceClassification.from_pretrained("bert-base-uncased", local_files_only=True, **kw)
            except Exception as e:  # no cache, no network
                logger.warning("pretrained trunks unavailable (%s): building from config", e)
                vit_model = bert = None
        if bert is None:
            bc = bert_config if bert_config is not None else BertConfig(num_labels=2, **kw)
            vc = vit_config if vit_config is not None else ViTConfig(**kw)
            vit_model = ViTModel(vc)
            bert = BertForSequenceClassification(bc)
        bert_model = bert.bert
        vit_pooler, bert_pooler = vit_model.pooler, bert_model.pooler
        if num_fusion_layers == 0:
            vit_other_layers, bert_other_layers = [], []
        else:
            vit_other_layers = vit_model.encoder.layer[-num_fusion_layers:]
            vit_model.encoder.layer = vit_model.encoder.layer[:-num_fusion_layers]
            bert_other_layers = bert_model.encoder.layer[-num_fusion_layers:]
            bert_model.encoder.layer = bert_model.encoder.layer[:-num_fusion_layers]
        return (vit_model, vit_pooler, vit_other_layers, bert_model, bert_pooler, bert_other_layers, bert.classifier,
                bert.dropout)

    def build_graphormer_graph_encoder_layer(self, embedding_dim, ffn_embedding_dim, num_attention_heads, dropout,
                                             attention_dropout, activation_dropout, activation_fn, export, q_noise,
                                             qn_block_size, pre_layernorm, num_layers=1):
        return GraphEncoderStack(
            num_layers=num_layers, embedding_dim=embedding_dim, ffn_embedding_dim=ffn_embedding_dim,
            num_attention_heads=num_attention_heads, dropout=dropout, attention_dropout=attention_dropout,
            activation_dropout=activation_dropout, activation_fn=activation_fn, export=export, q_noise=q_noise,
            qn_block_size=qn_block_size, pre_layernorm=pre_layernorm)

    # ------------------------------------------------------------------------------------------
    @staticmethod
    def _node_rows(batched_data) -> torch.Tensor:
        """int32 rows of the real comments in the flattened ``[B*N]`` layout."""
        if batched_data.get("node_index") is not None:
            return batched_data["node_index"]
        # reference-format batch: derive the list from the bool mask (one host sync, as every
        # boolean-mask index of the reference, multigraphormer_graph_encoder.py:322-324)
        return torch.nonzero(batched_data["x_token_mask"].reshape(-1)).squeeze(1).to(torch.int32)

    def forward(self, batched_data, last_state_only: bool = False, token_embeddings: Optional[torch.Tensor] = None,
                attn_mask: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """-> ``(bert_output [n_real,100,768], bottle_neck [n_real,nb,768], global_embedding [B,768])`` (:310-464)."""
        B, N = batched_data["x"].shape[:2]
        T = N + 1
        node_rows = self._node_rows(batched_data)
        if batched_data.get("x_nodes") is not None:          # node-major text shipped by this package's collator
            ids, type_ids, att = batched_data["x_nodes"][0], batched_data["x_nodes"][1], batched_data["x_nodes"][2]
        else:
            rows = node_rows.long()
            L = batched_data["x"].shape[2]
            ids = batched_data["x"].reshape(B * N, L).index_select(0, rows)
            type_ids = batched_data["x_token_type_ids"].reshape(B * N, L).index_select(0, rows)
            att = batched_data["x_attention_mask"].reshape(B * N, L).index_select(0, rows)
        n_real = ids.shape[0]
        dtype = self.bottle_neck.weight.dtype
        pack = None
        if self.pack_text and batched_data.get("x_text_pack") is not None and dtype in (torch.bfloat16, torch.float16):
            # packed (unpadded) text tokens from the embeddings on: the trunk layers in front of the fusion stack and the
            # BERT side of every fusion layer never see a padding token
            lens, total, longest = batched_data["x_text_pack"]
            pack = TextPack(lens, total, longest, self.num_bottle_neck, ids.shape[1])
            text_tokens = pack.gather_text(self.text_model.embeddings(input_ids=ids, token_type_ids=type_ids))
            for layer in self.text_model.encoder.layer:
                text_tokens = packed_bert_layer(layer, text_tokens, pack.text_only)
            bert_output = None
        else:
            bert_output = self.text_model(token_type_ids=type_ids, attention_mask=att, input_ids=ids).last_hidden_state
        vit_output = None
        if batched_data["x_images"] is not None:
            vit_output = self.vit_model(batched_data["x_images"].to(dtype)).last_hidden_state
        img_rows = batched_data.get("x_image_rows")
        if img_rows is None and vit_output is not None:
            img_rows = image_rows(batched_data["x_image_indexes"])

        bottle_neck = self.bottle_neck.weight.to(dtype).unsqueeze(0).repeat(n_real, 1, 1)                    # :339
        am = torch.cat([att.new_ones(n_real, self.num_bottle_neck), att], dim=1)
        ext_mask = ((1.0 - am[:, None, None, :].to(torch.float32)) * torch.finfo(torch.half).min).to(dtype)  # :341-354
        if pack is not None:
            x_text = pack.pack_tokens(text_tokens, bottle_neck)
            x_text, vit_output, bottle_neck = self.fusion_layers[0].forward_packed(x_text, vit_output, bottle_neck, pack, img_rows)
        else:
            bert_output, vit_output, bottle_neck = self.fusion_layers[0](bert_output, vit_output, bottle_neck, ext_mask, img_rows)

        # node-major -> padded [B,N,C]                                                                   :363-371
        graph_data = ops.scatter_rows(bottle_neck.new_zeros(B * N, 768), bottle_neck[:, 0, :], node_rows).view(B, N, 768)
        padding_mask = torch.cat([graph_data.new_zeros(B, 1, dtype=torch.bool), graph_data[:, :, 0].eq(0)], dim=1)  # :376-382
        x = self.graph_node_feature(graph_data, batched_data["in_degree"], batched_data["out_degree"])       # :386-388
        if self.fused_bias and attn_mask is None and batched_data.get("bias_code") is not None:
            attn_bias = self.graph_attn_bias.forward_fused(batched_data)
        else:
            attn_bias = self.graph_attn_bias(batched_data)                                                   # :392
        if self.embed_scale is not None:
            x = x * self.embed_scale
        if self.emb_layer_norm is not None:
            x = self.emb_layer_norm(x)
        x = self.dropout_module(x)
        x = x.transpose(0, 1).contiguous()                                                                   # [T,B,C] :407

        # rows of the real comments in the [T*B] layout of x: node (b, i) sits at t = i + 1
        nr = node_rows.long()
        tb_rows = (((nr % N) + 1) * B + nr // N).to(torch.int32)
        inner_states = [] if last_state_only else [x]
        for g_layer, f_layer in zip(self.layers, self.fusion_layers[1:]):                                    # :413-439
            x, _ = g_layer(x, self_attn_padding_mask=padding_mask, self_attn_mask=attn_mask, self_attn_bias=attn_bias)
            xf = x.reshape(T * B, self.embedding_dim)
            bottle_neck = torch.cat([ops.gather_rows(xf, tb_rows).unsqueeze(1), bottle_neck[:, 1:]], dim=1)   # :425
            if pack is not None:
                x_text, vit_output, bottle_neck = f_layer.forward_packed(x_text, vit_output, bottle_neck, pack, img_rows)
            else:
                bert_output, vit_output, bottle_neck = f_layer(bert_output, vit_output, bottle_neck, ext_mask, img_rows)
            x = ops.scatter_rows(xf, bottle_neck[:, 0, :], tb_rows).view(T, B, self.embedding_dim)           # :435
            if not last_state_only:
                inner_states.append(x)
        x, _ = self.layers[-1](x, self_attn_padding_mask=padding_mask, self_attn_mask=attn_mask, self_attn_bias=attn_bias)
        if pack is not None:
            if self.text_first_token_only:
                # the head pools token 0 of every comment (BertPooler): one row per comment instead of re-padding
                # [n_real, 100, 768] (630 MB) forward and an index_add over 222 K rows backward
                first = pack.cu_seqlens[:-1] + pack.nb
                bert_output = ops.gather_rows(x_text, first.contiguous()).unsqueeze(1)
            else:
                bert_output = pack.unpack_text(x_text)
        return bert_output, bottle_neck, x[0, :, :]
