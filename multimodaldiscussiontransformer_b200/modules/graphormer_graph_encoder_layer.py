"""``GraphormerGraphEncoderLayer`` / ``GraphEncoderStack`` with the reference's signatures and
parameter names (mDT/src/modules/graphormer_graph_encoder_layer.py:13-195)."""

from __future__ import annotations

from typing import Callable, Optional

import torch
import torch.nn as nn

from .. import ops
from ._fairseq_compat import FairseqDropout, LayerNorm, get_activation_fn, quant_noise
from .multihead_attention import MultiheadAttention


class GraphormerGraphEncoderLayer(nn.Module):
    def __init__(self, embedding_dim: int = 768, ffn_embedding_dim: int = 3072, num_attention_heads: int = 8,
                 dropout: float = 0.1, attention_dropout: float = 0.1, activation_dropout: float = 0.1,
                 activation_fn: str = "relu", export: bool = False, q_noise: float = 0.0, qn_block_size: int = 8,
                 init_fn: Callable = None, pre_layernorm: bool = False) -> None:
        super().__init__()
        if init_fn is not None:
            init_fn()
        self.embedding_dim = embedding_dim
        self.num_attention_heads = num_attention_heads
        self.attention_dropout = attention_dropout
        self.q_noise = q_noise
        self.qn_block_size = qn_block_size
        self.pre_layernorm = pre_layernorm
        self.dropout_module = FairseqDropout(dropout, module_name=self.__class__.__name__)
        self.activation_dropout_module = FairseqDropout(activation_dropout, module_name=self.__class__.__name__)
        self.activation_name = activation_fn
        self.activation_fn = get_activation_fn(activation_fn)
        self.self_attn = self.build_self_attention(self.embedding_dim, num_attention_heads, dropout=attention_dropout,
                                                   self_attention=True, q_noise=q_noise, qn_block_size=qn_block_size)
        self.self_attn_layer_norm = LayerNorm(self.embedding_dim, export=export)
        self.fc1 = self.build_fc1(self.embedding_dim, ffn_embedding_dim, q_noise=q_noise, qn_block_size=qn_block_size)
        self.fc2 = self.build_fc2(ffn_embedding_dim, self.embedding_dim, q_noise=q_noise, qn_block_size=qn_block_size)
        self.final_layer_norm = LayerNorm(self.embedding_dim, export=export)

    def build_fc1(self, input_dim, output_dim, q_noise, qn_block_size):
        return quant_noise(nn.Linear(input_dim, output_dim), q_noise, qn_block_size)

    def build_fc2(self, input_dim, output_dim, q_noise, qn_block_size):
        return quant_noise(nn.Linear(input_dim, output_dim), q_noise, qn_block_size)

    def build_self_attention(self, embed_dim, num_attention_heads, dropout, self_attention, q_noise, qn_block_size):
        return MultiheadAttention(embed_dim, num_attention_heads, dropout=dropout, self_attention=True, q_noise=q_noise,
                                  qn_block_size=qn_block_size)

    def forward(self, x: torch.Tensor, self_attn_bias=None, self_attn_mask: Optional[torch.Tensor] = None,
                self_attn_padding_mask: Optional[torch.Tensor] = None):
        """``x [T,B,C]`` -> ``(x, None)``; post-LN by default, pre-LN with ``pre_layernorm`` (:103-142)."""
        residual = x
        if self.pre_layernorm:
            x = self.self_attn_layer_norm(x)
        x, attn = self.self_attn(query=x, key=x, value=x, attn_bias=self_attn_bias,
                                 key_padding_mask=self_attn_padding_mask, need_weights=False, attn_mask=self_attn_mask)
        x = self._add_norm(x, residual, self.self_attn_layer_norm)
        residual = x
        if self.pre_layernorm:
            x = self.final_layer_norm(x)
        if (self.activation_name == "gelu" and x.is_cuda and x.dtype == torch.bfloat16 and self.fc1.weight.dtype == torch.bfloat16
                and self.fc1.out_features % 8 == 0):
            # fc1 -> gelu (fp32 arithmetic, as fairseq's) on csrc/gelu.cu: one pass instead of cast / gelu / cast, and the
            # backward produces the activation gradient and fc1's bias gradient together
            x = self.activation_dropout_module(ops.linear_gelu(x, self.fc1.weight, self.fc1.bias))
        else:
            x = self.activation_dropout_module(self.activation_fn(self.fc1(x)))
        x = self._add_norm(self.fc2(x), residual, self.final_layer_norm)
        return x, attn

    def _add_norm(self, x, residual, norm):
        """``residual + dropout(x)`` followed by the post-LN (:127-130, :139-141); one fused pass on the bf16 CUDA path."""
        if (not self.pre_layernorm and x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] in (256, 512, 768, 1024)
                and norm.weight is not None and norm.weight.dtype == torch.bfloat16):
            return ops.dropout_add_layer_norm(x, residual, norm.weight, norm.bias, self.dropout_module.p, self.training, norm.eps)
        x = residual + self.dropout_module(x)
        return x if self.pre_layernorm else norm(x)


class GraphEncoderStack(nn.Module):
    """``num_layers`` encoder layers sharing one bias / padding mask (:145-195)."""

    def __init__(self, num_layers, embedding_dim: int = 768, ffn_embedding_dim: int = 3072, num_attention_heads: int = 8,
                 dropout: float = 0.1, attention_dropout: float = 0.1, activation_dropout: float = 0.1,
                 activation_fn: str = "relu", export: bool = False, q_noise: float = 0.0, qn_block_size: int = 8,
                 init_fn: Callable = None, pre_layernorm: bool = False):
        super().__init__()
        self.layers = nn.ModuleList([
            GraphormerGraphEncoderLayer(embedding_dim, ffn_embedding_dim, num_attention_heads, dropout, attention_dropout,
                                        activation_dropout, activation_fn, export, q_noise, qn_block_size, init_fn,
                                        pre_layernorm)
            for _ in range(num_layers)
        ])

    def forward(self, x: torch.Tensor, self_attn_bias=None, self_attn_mask: Optional[torch.Tensor] = None,
                self_attn_padding_mask: Optional[torch.Tensor] = None):
        if len(self.layers) == 0:
            # the reference leaves `attn` unbound here and raises (graphormer_graph_encoder_layer.py:163-166,193-195)
            raise UnboundLocalError("GraphEncoderStack with num_layers == 0 cannot run (as in the reference)")
        attn = None
        for layer in self.layers:
            x, attn = layer(x, self_attn_bias, self_attn_mask, self_attn_padding_mask)
        return x, attn
