"""``MultiheadAttention`` with the reference's signature and parameter names
(mDT/src/modules/multihead_attention.py:17-248); the attention core runs on the tcgen05 kernel."""

from __future__ import annotations

import math
from typing import Optional, Tuple, Union

import torch
import torch.nn.functional as F
from torch import Tensor, nn

from .. import ops
from ._fairseq_compat import FairseqDropout, quant_noise
from .graphormer_layers import FusedAttnBias

KERNEL_HEAD_DIM = 64


class MultiheadAttention(nn.Module):
    """Self-attention with an additive per-head bias.  ``key`` / ``value`` are ignored, as in the
    reference (multihead_attention.py:134-136 projects ``query`` three times)."""

    def __init__(self, embed_dim, num_heads, kdim=None, vdim=None, dropout=0.0, bias=True, self_attention=False,
                 q_noise=0.0, qn_block_size=8):
        super().__init__()
        self.embed_dim = embed_dim
        self.kdim = kdim if kdim is not None else embed_dim
        self.vdim = vdim if vdim is not None else embed_dim
        self.qkv_same_dim = self.kdim == embed_dim and self.vdim == embed_dim
        self.num_heads = num_heads
        self.dropout_module = FairseqDropout(dropout, module_name=self.__class__.__name__)
        self.head_dim = embed_dim // num_heads
        assert self.head_dim * num_heads == self.embed_dim, "embed_dim must be divisible by num_heads"
        assert self.head_dim <= KERNEL_HEAD_DIM, "the B200 attention kernel covers head_dim <= 64"
        self.scaling = self.head_dim ** -0.5
        self.self_attention = self_attention
        assert self.self_attention, "Only support self attention"
        assert not self.self_attention or self.qkv_same_dim, (
            "Self-attention requires query, key and value to be of the same size")
        self.k_proj = quant_noise(nn.Linear(self.kdim, embed_dim, bias=bias), q_noise, qn_block_size)
        self.v_proj = quant_noise(nn.Linear(self.vdim, embed_dim, bias=bias), q_noise, qn_block_size)
        self.q_proj = quant_noise(nn.Linear(embed_dim, embed_dim, bias=bias), q_noise, qn_block_size)
        self.out_proj = quant_noise(nn.Linear(embed_dim, embed_dim, bias=bias), q_noise, qn_block_size)
        self.reset_parameters()
        self.onnx_trace = False

    def prepare_for_onnx_export_(self):
        raise NotImplementedError

    def reset_parameters(self):
        gain = 1 / math.sqrt(2) if self.qkv_same_dim else 1.0  # multihead_attention.py:75-89
        nn.init.xavier_uniform_(self.k_proj.weight, gain=gain)
        nn.init.xavier_uniform_(self.v_proj.weight, gain=gain)
        nn.init.xavier_uniform_(self.q_proj.weight, gain=gain)
        nn.init.xavier_uniform_(self.out_proj.weight)
        if self.out_proj.bias is not None:
            nn.init.constant_(self.out_proj.bias, 0.0)

    # ------------------------------------------------------------------------------------
    def _project_qkv(self, x: Tensor):
        """One [3C, C] GEMM over the fused parameter storage of q_proj / k_proj / v_proj (``ops.linear_qkv``: no ``cat`` of
        the weights per forward, the weight gradient comes back as three slices); q is then scaled by d^-0.5
        (multihead_attention.py:137)."""
        T, B, C = x.shape
        if x.is_cuda:
            qkv = ops.linear_qkv(x, self.q_proj, self.k_proj, self.v_proj)
        else:  # host-side checks of the module logic (no kernels involved)
            qkv = torch.cat([self.q_proj(x), self.k_proj(x), self.v_proj(x)], dim=-1)
        qkv = qkv.view(T, B, 3, self.num_heads, self.head_dim)
        if qkv.dtype != torch.bfloat16:
            qkv = qkv.to(torch.bfloat16)
        if self.head_dim != KERNEL_HEAD_DIM:  # zero-padded head dim: same scores, same outputs
            qkv = F.pad(qkv, (0, KERNEL_HEAD_DIM - self.head_dim))
        q = (qkv[:, :, 0] * self.scaling).permute(1, 0, 2, 3)
        return q, qkv[:, :, 1].permute(1, 0, 2, 3), qkv[:, :, 2].permute(1, 0, 2, 3)  # [B, T, H, 64] views

    def forward(self, query, key: Optional[Tensor], value: Optional[Tensor],
                attn_bias: Union[Tensor, FusedAttnBias, None], key_padding_mask: Optional[Tensor] = None,
                need_weights: bool = True, attn_mask: Optional[Tensor] = None, before_softmax: bool = False,
                need_head_weights: bool = False) -> Tuple[Tensor, Optional[Tensor]]:
        """Input ``[T, B, C]``; returns ``(attn [T, B, C], None)`` (the hot path calls with need_weights=False)."""
        if need_head_weights:
            need_weights = True
        tgt_len, bsz, embed_dim = query.size()
        assert embed_dim == self.embed_dim, f"query dim {embed_dim} != {self.embed_dim}"
        if key is not None:
            src_len, key_bsz, _ = key.size()
            assert key_bsz == bsz and value is not None and src_len == tgt_len
        if need_weights or before_softmax:
            raise NotImplementedError(
                "the B200 attention kernel never materialises the [B*H,T,T] weights; the mDT hot path calls "
                "MultiheadAttention with need_weights=False (graphormer_graph_encoder_layer.py:124)")
        if key_padding_mask is not None and key_padding_mask.dim() == 0:
            key_padding_mask = None
        if key_padding_mask is not None:
            assert key_padding_mask.size(0) == bsz and key_padding_mask.size(1) == tgt_len

        q, k, v = self._project_qkv(query)
        bias = code = lut = None
        if isinstance(attn_bias, FusedAttnBias):
            if attn_mask is None:
                code, lut = attn_bias.code, attn_bias.lut
            else:
                bias = attn_bias.materialise()
        elif attn_bias is not None:
            bias = attn_bias.view(bsz, self.num_heads, tgt_len, tgt_len)
            if bias.dtype not in (torch.float32, torch.bfloat16):
                bias = bias.float()
        if attn_mask is not None:  # additive [T,T] mask (multihead_attention.py:176-178)
            m = attn_mask.to(torch.float32).view(1, 1, tgt_len, tgt_len)
            bias = m.expand(bsz, self.num_heads, tgt_len, tgt_len) if bias is None else bias.float() + m
        # attention dropout (multihead_attention.py:196) happens inside the kernel: counter-hash mask, regenerated in the backward
        p_drop = float(self.dropout_module.p) if self.training else 0.0
        o = ops.biased_attention(q, k, v, bias=bias, code=code, lut=lut, key_padding_mask=key_padding_mask, dropout_p=p_drop)
        # o: [B,T,H,64] stored t-major, i.e. [T,B,H,64] in memory
        o = o.permute(1, 0, 2, 3)
        if self.head_dim != KERNEL_HEAD_DIM:
            o = o[..., : self.head_dim]
        attn = o.reshape(tgt_len, bsz, embed_dim).to(query.dtype)
        return self.out_proj(attn), None

    def apply_sparse_mask(self, attn_weights, tgt_len: int, src_len: int, bsz: int):
        return attn_weights

    def upgrade_state_dict_named(self, state_dict, name):
        """Split legacy ``in_proj_weight`` / ``in_proj_bias`` into q/k/v (multihead_attention.py:219-248)."""
        prefix = name + "." if name != "" else ""
        for k in [k for k in state_dict.keys() if k.endswith(prefix + "in_proj_weight")]:
            w = state_dict.pop(k)
            dim = w.shape[0] // 3
            for i, nm in enumerate(("q_proj", "k_proj", "v_proj")):
                state_dict[prefix + nm + ".weight"] = w[i * dim:(i + 1) * dim]
            kb = prefix + "in_proj_bias"
            if kb in state_dict:
                bvec = state_dict.pop(kb)
                for i, nm in enumerate(("q_proj", "k_proj", "v_proj")):
                    state_dict[prefix + nm + ".bias"] = bvec[i * dim:(i + 1) * dim]
