"""``GraphNodeFeature`` / ``GraphAttnBias`` with the reference's constructor signatures and
parameter names (mDT/src/modules/graphormer_layers.py:16-110), running on libmdt_b200."""

from __future__ import annotations

import math
from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from .. import ops


def init_params(module, n_layers):
    """graphormer_layers.py:7-13 -- note that padding rows of embeddings are re-drawn too."""
    if isinstance(module, nn.Linear):
        module.weight.data.normal_(mean=0.0, std=0.02 / math.sqrt(n_layers))
        if module.bias is not None:
            module.bias.data.zero_()
    if isinstance(module, nn.Embedding):
        module.weight.data.normal_(mean=0.0, std=0.02)


class FusedAttnBias(NamedTuple):
    """Attention bias that is never materialised: ``bias[b,h,i,j] = lut[h, code[b,i,j]]`` (255 = -inf).

    Accepted wherever the modules of this package take ``attn_bias`` / ``self_attn_bias``.
    """

    code: torch.Tensor  # uint8 [B, T, pitch]
    lut: torch.Tensor   # float32 [H, 256], differentiable w.r.t. the GraphAttnBias parameters

    def materialise(self, dtype=torch.float32) -> torch.Tensor:
        B, T, _ = self.code.shape
        c = self.code[:, :, :T].long()
        bias = self.lut[:, c.reshape(-1)].reshape(self.lut.shape[0], B, T, T).permute(1, 0, 2, 3)
        return bias.masked_fill((c == 255).unsqueeze(1), float("-inf")).to(dtype)


class GraphNodeFeature(nn.Module):
    """Degree embeddings + graph token (graphormer_layers.py:16-50): one fused gather-add kernel."""

    def __init__(self, num_heads, num_atoms, num_in_degree, num_out_degree, hidden_dim, n_layers):
        super().__init__()
        self.num_heads = num_heads
        self.num_atoms = num_atoms
        self.atom_encoder = nn.Embedding(num_atoms + 1, hidden_dim, padding_idx=0)  # dead in mDT, kept for checkpoints
        self.in_degree_encoder = nn.Embedding(num_in_degree, hidden_dim, padding_idx=0)
        self.out_degree_encoder = nn.Embedding(num_out_degree, hidden_dim, padding_idx=0)
        self.graph_token = nn.Embedding(1, hidden_dim)
        self.apply(lambda module: init_params(module, n_layers=n_layers))

    def forward(self, x, in_degree, out_degree):
        return ops.node_feature(x, in_degree, out_degree, self.in_degree_encoder.weight, self.out_degree_encoder.weight,
                                self.graph_token.weight)


class GraphAttnBias(nn.Module):
    """Spatial-position bias per head (graphormer_layers.py:53-110); there is no edge term in mDT."""

    def __init__(self, num_heads, num_atoms, num_edges, num_spatial, num_edge_dis, hidden_dim, edge_type,
                 multi_hop_max_dist, n_layers):
        super().__init__()
        self.num_heads = num_heads
        self.multi_hop_max_dist = multi_hop_max_dist
        self.edge_encoder = nn.Embedding(num_edges + 1, num_heads, padding_idx=0)      # never read (:86-110)
        self.edge_type = edge_type
        if self.edge_type == "multi_hop":
            self.edge_dis_encoder = nn.Embedding(num_edge_dis * num_heads * num_heads, 1)  # never read
        self.spatial_pos_encoder = nn.Embedding(num_spatial, num_heads, padding_idx=0)
        self.graph_token_virtual_distance = nn.Embedding(1, num_heads)
        self.apply(lambda module: init_params(module, n_layers=n_layers))

    def forward(self, batched_data) -> torch.Tensor:
        """``[B,H,T,T]`` tensor, as the reference returns it (one coalesced kernel instead of ~8 passes)."""
        return ops.attn_bias(batched_data["attn_bias"], batched_data["spatial_pos"], self.spatial_pos_encoder.weight,
                             self.graph_token_virtual_distance.weight)

    def forward_fused(self, batched_data) -> FusedAttnBias:
        """The same bias as ``forward`` as (byte code, per-head table): consumed by the attention kernel directly.

        Needs ``batched_data["bias_code"]`` from this package's collator.  Row 0 of the table (padding_idx)
        takes part in the forward value but receives no gradient, like ``nn.Embedding``.
        """
        code = batched_data["bias_code"]
        table = self.spatial_pos_encoder.weight.float()
        H = self.num_heads
        n = min(table.shape[0], 254)
        rows = torch.cat([table[:1].detach(), table[1:n]], dim=0).t()                       # [H, n]
        parts = [rows]
        if n < 254:
            parts.append(rows.new_zeros(H, 254 - n))
        parts.append(self.graph_token_virtual_distance.weight.float().view(H, 1))            # code 254
        parts.append(rows.new_zeros(H, 1))                                                   # code 255 (-inf in-kernel)
        return FusedAttnBias(code, torch.cat(parts, dim=1))
