"""The handful of fairseq helpers the reference modules import (fairseq is a third-party
dependency of the reference, version unpinned, not installable here -- SURVEY.md 8c).

Restated by their published behaviour, cited at the reference's call sites:
  * ``utils.get_activation_fn``            graphormer_graph_encoder_layer.py:50
  * ``FairseqDropout``                     multihead_attention.py:36, graphormer_graph_encoder_layer.py:42-47
  * ``LayerNorm``                          graphormer_graph_encoder_layer.py:61,77
  * ``quant_noise``                        multihead_attention.py:54-66 (identity at p == 0)
"""

from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F


def get_activation_fn(name: str):
    if name == "relu":
        return F.relu
    if name == "gelu":
        return lambda x: F.gelu(x.float()).type_as(x)  # fairseq computes gelu in fp32
    if name == "tanh":
        return torch.tanh
    if name == "linear":
        return lambda x: x
    raise RuntimeError(f"--activation-fn {name} not supported")


class FairseqDropout(nn.Module):
    def __init__(self, p: float, module_name=None):
        super().__init__()
        self.p = p
        self.module_name = module_name

    def forward(self, x, inplace: bool = False):
        if self.p > 0 and self.training:
            return F.dropout(x, p=self.p, training=True, inplace=inplace)
        return x


def LayerNorm(normalized_shape, eps=1e-5, elementwise_affine=True, export=False):
    return nn.LayerNorm(normalized_shape, eps, elementwise_affine)


def quant_noise(module, p, block_size):
    if p > 0:
        raise NotImplementedError("quant-noise (q_noise > 0) is outside the B200 hot path")
    return module
