"""``GraphFusionLayer`` / ``GraphFusionStack`` (mDT/src/modules/multi_graphormer_fusion_layer.py:8-192):
bottleneck tokens are prepended to the text tokens of a BERT layer and to the patches of a ViT
layer; comments that carry an image average the two bottleneck outputs.  The BERT / ViT layer
arithmetic is Hugging Face's (third-party for the reference too); the glue here is sync-free
when ``x_image_indexes`` is given as an int row list."""

from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn


def _first(out):
    return out[0] if isinstance(out, (tuple, list)) else out


def image_rows(x_image_indexes: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """bool ``[n_real]`` (reference format; costs one device->host sync) or int row list -> int64 rows."""
    if x_image_indexes is None:
        return None
    if x_image_indexes.dtype == torch.bool:
        return torch.nonzero(x_image_indexes).squeeze(1)
    return x_image_indexes.long()


class GraphFusionLayer(nn.Module):
    def __init__(self, bert_layer, vit_layer, num_bottle_neck_tokens: int, use_projection: bool = False) -> None:
        super().__init__()
        self.bert_encoder = bert_layer
        self.vit_encoder = vit_layer
        self.gradient_checkpointing = False
        self.num_bottle_neck_tokens = num_bottle_neck_tokens
        if use_projection:  # constructed, never applied (multi_graphormer_fusion_layer.py:22-27)
            self.bert_projection = nn.Linear(768, 768)
            self.vit_projection = nn.Linear(768, 768)
        else:
            self.bert_projection = nn.Identity()
            self.vit_projection = nn.Identity()

    def bert_forward(self, hidden_states, attention_mask=None, head_mask=None, encoder_hidden_states=None,
                     encoder_attention_mask=None):
        return _first(self.bert_encoder(hidden_states, attention_mask))

    def vit_forward(self, hidden_states):
        return _first(self.vit_encoder(hidden_states))

    def forward(self, bert_hidden_states, vit_hidden_states, bottle_neck, bert_attention_mask=None, x_image_indexes=None):
        nb = self.num_bottle_neck_tokens
        bert_out = self.bert_forward(torch.cat([bottle_neck, bert_hidden_states], dim=1), bert_attention_mask)
        bert_hidden_output, bottle_neck_output = bert_out[:, nb:], bert_out[:, :nb]
        if vit_hidden_states is not None:
            rows = image_rows(x_image_indexes)
            # the ViT layer sees the *input* bottleneck, not BERT's output (:57)
            vit_out = self.vit_forward(torch.cat([bottle_neck.index_select(0, rows), vit_hidden_states], dim=1))
            vit_hidden_output = vit_out[:, nb:]
            mixed = (vit_out[:, :nb] + bottle_neck_output.index_select(0, rows)) / 2          # :63-66
            bottle_neck_output = bottle_neck_output.index_copy(0, rows, mixed)
        else:
            vit_hidden_output = None
        return bert_hidden_output, vit_hidden_output, bottle_neck_output


class GraphFusionStack(nn.Module):
    def __init__(self, bert_layers, vit_layers, num_bottle_neck_tokens, use_projection=False) -> None:
        super().__init__()
        self.fusion_layers = nn.ModuleList([
            GraphFusionLayer(b, v, num_bottle_neck_tokens, use_projection) for b, v in zip(bert_layers, vit_layers)])

    def forward(self, bert_hidden_states, vit_hidden_states, bottle_neck, bert_attention_mask=None, x_image_indexes=None):
        rows = image_rows(x_image_indexes)  # resolved once per stack
        for f_layer in self.fusion_layers:
            bert_hidden_states, vit_hidden_states, bottle_neck = f_layer(
                bert_hidden_states, vit_hidden_states, bottle_neck, bert_attention_mask, rows)
        return bert_hidden_states, vit_hidden_states, bottle_neck
