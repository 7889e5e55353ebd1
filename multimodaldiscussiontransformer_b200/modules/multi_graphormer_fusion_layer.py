"""``GraphFusionLayer`` / ``GraphFusionStack`` (mDT/src/modules/multi_graphormer_fusion_layer.py:8-192):
bottleneck tokens are prepended to the text tokens of a BERT layer and to the patches of a ViT
layer; comments that carry an image average the two bottleneck outputs.  The BERT / ViT layer
arithmetic is Hugging Face's (third-party for the reference too); the glue here is sync-free
when ``x_image_indexes`` is given as an int row list."""

from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn as nn

from .. import ops


def _first(out):
    return out[0] if isinstance(out, (tuple, list)) else out


def _rows_i32(rows: torch.Tensor) -> torch.Tensor:
    """int32 copy of a row list, cached on the tensor (one conversion per batch, not per layer)."""
    r = getattr(rows, "_mdt_i32", None)
    if r is None:
        r = rows.to(torch.int32).contiguous()
        try:
            rows._mdt_i32 = r
        except Exception:
            pass
    return r


def image_rows(x_image_indexes: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """bool ``[n_real]`` (reference format; costs one device->host sync) or int row list -> int64 rows."""
    if x_image_indexes is None:
        return None
    if x_image_indexes.dtype == torch.bool:
        return torch.nonzero(x_image_indexes).squeeze(1)
    return x_image_indexes.long()


class TextPack:
    """Packed (unpadded) layout of the BERT side of the fusion layers.

    Every comment i is one sequence ``[nb bottleneck slots | len_i real text tokens]``; the sequences are laid end
    to end in one ``[total, C]`` buffer, so the layer GEMMs, LayerNorms, GELUs and the attention only touch real
    tokens (padding is ~46 % of a HatefulDiscussions-shaped batch).  Built on the device from the per-comment
    lengths without a host sync: the totals come from the host side of the collator.
    Padding tokens never influence a real token in the reference either (additive key mask,
    multigraphormer_graph_encoder.py:341-354), so logits, loss and gradients are unchanged; the hidden states
    returned *at padding positions* are zeros instead of the reference's don't-care values.
    """

    def __init__(self, text_lens: torch.Tensor, text_tokens: int, text_max: int, nb: int, L: int):
        dev = text_lens.device
        n = text_lens.numel()
        lens = text_lens.long()
        self.nb, self.L, self.n = nb, L, n
        self.total = text_tokens + nb * n
        self.max_seqlen = text_max + nb
        cu = torch.zeros(n + 1, dtype=torch.int64, device=dev)
        cu[1:] = torch.cumsum(lens + nb, 0)
        self.cu_seqlens = cu.to(torch.int32)
        self._plan = None
        seq = torch.repeat_interleave(torch.arange(n, device=dev), lens, output_size=text_tokens)
        cu_text = torch.cumsum(lens, 0) - lens
        pos = torch.arange(text_tokens, device=dev) - cu_text[seq]
        self.text_rows = cu[seq] + nb + pos                       # packed row of every real text token
        self.padded_rows = seq * L + pos                          # its row in the padded [n * L] layout
        self.bn_rows = (cu[:-1, None] + torch.arange(nb, device=dev)[None, :]).reshape(-1)  # [n * nb]
        # the same sequences without bottleneck slots (the trunk layers in front of the fusion stack)
        cu0 = torch.zeros(n + 1, dtype=torch.int64, device=dev)
        cu0[1:] = torch.cumsum(lens, 0)
        self.text_only = _SeqLens(cu0.to(torch.int32), text_max, text_tokens)

    @property
    def plan(self):
        """Tile plan of the packed attention kernel (csrc/attn_varlen.cu): built once, shared by all layers and passes."""
        if self._plan is None:
            self._plan = ops.varlen_plan(self.cu_seqlens, self.total)
        return self._plan

    def gather_text(self, text_padded: torch.Tensor) -> torch.Tensor:
        """padded ``[n, L, C]`` -> the real tokens ``[text_tokens, C]`` in sequence order (layout of ``text_only``)."""
        C = text_padded.shape[-1]
        return text_padded.reshape(-1, C).index_select(0, self.padded_rows)

    def pack_tokens(self, text_tokens: torch.Tensor, bottle_neck: torch.Tensor) -> torch.Tensor:
        """real tokens in sequence order + bottleneck -> the fusion layout."""
        C = text_tokens.shape[-1]
        x = text_tokens.new_empty(self.total, C)
        return x.index_copy(0, self.text_rows, text_tokens).index_copy(0, self.bn_rows, bottle_neck.reshape(-1, C))

    def unpack_text(self, x: torch.Tensor) -> torch.Tensor:
        C = x.shape[-1]
        out = x.new_zeros(self.n * self.L, C)
        return out.index_copy(0, self.padded_rows, x.index_select(0, self.text_rows)).view(self.n, self.L, C)


class _SeqLens:
    def __init__(self, cu_seqlens, max_seqlen, total):
        self.cu_seqlens, self.max_seqlen, self.total = cu_seqlens, max_seqlen, total
        self._plan = None

    @property
    def plan(self):
        if self._plan is None:
            self._plan = ops.varlen_plan(self.cu_seqlens, self.total)
        return self._plan


#: "mdt" = this library's tcgen05 kernels (csrc/attn_varlen.cu); "flash" = flash-attn's varlen entry, kept only as the
#: A/B comparison of benchmarks/text_attn_probe.py (MDT_TEXT_ATTN=flash) -- never a fallback: "mdt" raises when it cannot run.
TEXT_ATTN_IMPL = os.environ.get("MDT_TEXT_ATTN", "mdt")
#: "0" keeps the block as separate autograd nodes (A/B runs); the arithmetic and the kernels are the same
FUSED_TEXT_LAYER = os.environ.get("MDT_FUSED_TEXT_LAYER", "1") != "0"
FUSED_VIT_LAYER = os.environ.get("MDT_FUSED_VIT_LAYER", "1") != "0"
_FLASH = None


def _varlen_attention(qkv, pack, dropout_p: float):
    """Self-attention over the packed sequences: ``[total, 3, H, 64]`` -> ``[total, H, 64]``; one ``[total, 3, H, 64]``
    gradient comes back, so autograd never re-assembles three slices."""
    if TEXT_ATTN_IMPL == "mdt":
        if pack.max_seqlen > 128:
            raise RuntimeError(f"packed text attention handles sequences of <= 128 rows, got {pack.max_seqlen}")
        return ops.varlen_attention(qkv, pack.plan, dropout_p=dropout_p)
    global _FLASH
    if _FLASH is None:
        from flash_attn import flash_attn_varlen_qkvpacked_func
        _FLASH = flash_attn_varlen_qkvpacked_func
    return _FLASH(qkv, pack.cu_seqlens, pack.max_seqlen, dropout_p=dropout_p)


def packed_bert_layer(layer, x: torch.Tensor, pack, take_rows=None):
    """Hugging Face ``BertLayer`` arithmetic (post-LN block, same parameters, same dropout rates) on packed tokens.

    The three projections run as ONE GEMM over the fused storage of the ``query / key / value`` parameters
    (``ops.linear_qkv``: the parameters keep their names and shapes, only their ``.data`` become adjacent slices of one
    buffer): ``x`` is read once, no weights are concatenated per call, and the backward has one input-gradient GEMM and one
    bias-gradient reduction instead of three of each plus two gradient-accumulation adds over ``[total, C]``."""
    att, sa = layer.attention, layer.attention.self
    H = sa.num_attention_heads
    d = x.shape[-1] // H
    if (FUSED_TEXT_LAYER and TEXT_ATTN_IMPL == "mdt" and x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and d == 64
            and x.shape[-1] in (256, 512, 768, 1024) and getattr(layer.intermediate, "_mdt_exact_gelu", False)
            and sa.query.bias is not None):
        # the whole block as one autograd node (ops.packed_post_ln_layer): no gradient-accumulation adds and no separate
        # bias-gradient reductions in the backward
        if pack.max_seqlen > 128:
            raise RuntimeError(f"packed text attention handles sequences of <= 128 rows, got {pack.max_seqlen}")
        return ops.packed_post_ln_layer(x, pack.plan, H, sa.dropout.p, att.output.dropout.p, sa.query, sa.key, sa.value,
                                        att.output.dense, att.output.LayerNorm, layer.intermediate.dense, layer.output.dense,
                                        layer.output.LayerNorm, layer.training, take_rows=take_rows)
    qkv = ops.linear_qkv(x, sa.query, sa.key, sa.value).view(-1, 3, H, d)
    ctx = _varlen_attention(qkv, pack, sa.dropout.p if layer.training else 0.0).reshape(x.shape)
    if x.dtype == torch.bfloat16 and x.shape[-1] in (256, 512, 768, 1024):
        # dense -> dropout -> + residual -> LayerNorm tails in one pass each way (libmdt_b200, csrc/layernorm.cu)
        ln1, ln2 = att.output.LayerNorm, layer.output.LayerNorm
        h = ops.dropout_add_layer_norm(att.output.dense(ctx), x, ln1.weight, ln1.bias, att.output.dropout.p, layer.training, ln1.eps)
        inter = layer.intermediate(h)
        return ops.dropout_add_layer_norm(layer.output.dense(inter), h, ln2.weight, ln2.bias, layer.output.dropout.p,
                                          layer.training, ln2.eps)
    h = att.output.LayerNorm(att.output.dropout(att.output.dense(ctx)) + x)
    inter = layer.intermediate(h)
    return layer.output.LayerNorm(layer.output.dropout(layer.output.dense(inter)) + h)


class _PrependTokens(torch.autograd.Function):
    """``cat([bottle_neck, hidden], dim=1)`` as two strided copies into one buffer.  ``hidden`` is the ``[:, nb:]`` view of
    the previous layer's output, and ``torch.cat`` serves a non-contiguous input with its scalar (non-vectorised) copy
    kernel (0.5 TB/s on the ViT tokens); the gradients are the two views of the incoming gradient, as in ``cat``."""

    @staticmethod
    def forward(ctx, bn, hidden):
        nb = bn.shape[1]
        out = hidden.new_empty(hidden.shape[0], nb + hidden.shape[1], hidden.shape[2])
        out[:, :nb].copy_(bn)
        out[:, nb:].copy_(hidden)
        ctx.nb = nb
        return out

    @staticmethod
    def backward(ctx, g):
        return g[:, :ctx.nb], g[:, ctx.nb:]


def prepend_tokens(bn: torch.Tensor, hidden: torch.Tensor) -> torch.Tensor:
    if hidden.is_cuda and not hidden.is_contiguous():
        return _PrependTokens.apply(bn, hidden)
    return torch.cat([bn, hidden], dim=1)


class FusedLayerNorm(nn.LayerNorm):
    """``nn.LayerNorm`` (same parameters, same state-dict keys) whose bf16 CUDA forward / backward is the one-pass
    kernel of csrc/layernorm.cu; anything else takes PyTorch's implementation."""

    def forward(self, x):
        if (x.is_cuda and x.dtype == torch.bfloat16 and len(self.normalized_shape) == 1 and self.weight is not None
                and self.bias is not None and self.weight.dtype == torch.bfloat16
                and self.normalized_shape[0] in (256, 512, 768, 1024)):
            return ops.dropout_add_layer_norm(x, None, self.weight, self.bias, 0.0, False, self.eps)
        return super().forward(x)


#: "mdt" = the ViT blocks' self-attention on this library's tcgen05 kernels (bf16 CUDA tensors); "sdpa" = Hugging Face's
#: default (cuDNN / flash SDPA), kept as the A/B comparison (MDT_VIT_ATTN=sdpa).
VIT_ATTN_IMPL = os.environ.get("MDT_VIT_ATTN", "mdt")


def _mdt_vit_attention(module, query, key, value, attention_mask, scaling=None, dropout: float = 0.0, **kwargs):
    """Hugging Face attention-interface function (``ALL_ATTENTION_FUNCTIONS``): ``query / key / value`` are ``[B, H, T, d]``.

    bf16 CUDA tensors with d = 64 and no mask (every ViT block of mDT: 197 patches [+ 4 bottleneck tokens],
    multi_graphormer_fusion_layer.py:73-100) run on ``mdt_biased_attn_fwd/bwd`` (no bias, in-kernel dropout); anything else
    (the fp32 parity configurations) takes Hugging Face's own SDPA function, as the reference does."""
    if (query.is_cuda and query.dtype == torch.bfloat16 and query.shape[-1] == 64 and attention_mask is None
            and VIT_ATTN_IMPL == "mdt"):
        scale = query.shape[-1] ** -0.5 if scaling is None else scaling
        out = ops.biased_attention((query * scale).transpose(1, 2), key.transpose(1, 2), value.transpose(1, 2),
                                   dropout_p=float(dropout))
        return out, None
    from transformers.modeling_utils import ALL_ATTENTION_FUNCTIONS

    return ALL_ATTENTION_FUNCTIONS["sdpa"](module, query, key, value, attention_mask, scaling=scaling, dropout=dropout, **kwargs)


def use_mdt_attention(module: nn.Module) -> nn.Module:
    """Route the self-attention of every Hugging Face block below ``module`` through ``_mdt_vit_attention``."""
    from transformers.modeling_utils import ALL_ATTENTION_FUNCTIONS

    if "mdt_b200" not in ALL_ATTENTION_FUNCTIONS.valid_keys():
        ALL_ATTENTION_FUNCTIONS.register("mdt_b200", _mdt_vit_attention)
    for m in module.modules():
        cfg = getattr(m, "config", None)
        if cfg is not None and hasattr(cfg, "_attn_implementation") and cfg._attn_implementation in ("sdpa", "eager", None):
            cfg._attn_implementation = "mdt_b200"
    return module


class FusedIntermediate(nn.Module):
    """Hugging Face ``BertIntermediate`` / ``ViTIntermediate`` (``dense`` -> exact GELU; same parameters and state-dict keys)
    whose bf16 CUDA forward is ``ops.linear_gelu``: the activation is one pass of csrc/gelu.cu and its backward also
    yields the bias gradient of ``dense``; anything else takes Hugging Face's own forward."""

    def forward(self, hidden_states):
        d = self.dense
        if (hidden_states.is_cuda and hidden_states.dtype == torch.bfloat16 and d.weight.dtype == torch.bfloat16
                and d.out_features % 8 == 0 and getattr(self, "_mdt_exact_gelu", False)):
            return ops.linear_gelu(hidden_states, d.weight, d.bias)
        return self.intermediate_act_fn(d(hidden_states))


def use_fused_intermediates(module: nn.Module) -> nn.Module:
    """Re-class the ``dense -> gelu`` blocks below ``module`` whose activation is the exact (erf) GELU."""
    from transformers.activations import GELUActivation

    for m in module.modules():
        if type(m).__name__ in ("BertIntermediate", "ViTIntermediate") and hasattr(m, "dense"):
            act = getattr(m, "intermediate_act_fn", None)
            exact = (isinstance(act, GELUActivation) and not getattr(act, "use_gelu_python", False)) or act is torch.nn.functional.gelu
            m._mdt_exact_gelu = bool(exact)
            m.__class__ = type("Fused" + type(m).__name__, (FusedIntermediate, type(m)), {})
    return module


class FusedViTLayer(nn.Module):
    """Hugging Face ``ViTLayer`` (same parameters and state-dict keys) whose bf16 CUDA forward is the single autograd node
    ``ops.vit_pre_ln_layer``; anything else (fp32 parity runs, ``MDT_FUSED_VIT_LAYER=0``) takes Hugging Face's forward."""

    def forward(self, hidden_states, *args, take_tokens=None, **kwargs):
        sa = self.attention.attention
        if (FUSED_VIT_LAYER and VIT_ATTN_IMPL == "mdt" and hidden_states.is_cuda and hidden_states.dtype == torch.bfloat16
                and hidden_states.dim() == 3 and hidden_states.shape[-1] in (256, 512, 768, 1024)
                and hidden_states.shape[-1] == 64 * sa.num_attention_heads and sa.query.bias is not None
                and getattr(self.intermediate, "_mdt_exact_gelu", False) and not args and not kwargs):
            return ops.vit_pre_ln_layer(hidden_states, sa.num_attention_heads, sa.dropout_prob, self.output.dropout.p,
                                        self.layernorm_before, sa.query, sa.key, sa.value, self.attention.output.dense,
                                        self.layernorm_after, self.intermediate.dense, self.output.dense, self.training,
                                        take_tokens=take_tokens)
        return super().forward(hidden_states, *args, **kwargs)


def use_fused_vit_layers(module: nn.Module) -> nn.Module:
    """Re-class every Hugging Face ``ViTLayer`` at or below ``module``."""
    for m in module.modules():
        if type(m).__name__ == "ViTLayer":
            m.__class__ = type("FusedViTLayer_", (FusedViTLayer, type(m)), {})
    return module


class PatchEmbeddingsAsGemm(nn.Module):
    """Hugging Face ``ViTPatchEmbeddings`` (a ``Conv2d`` whose stride equals its kernel; same parameters and state-dict keys)
    computed as what it is -- one GEMM over the flattened, non-overlapping patches -- on bf16 CUDA tensors: cuDNN runs the
    convolution form through two layout transposes and an implicit-GEMM kernel (2.6 ms for 410 images; the patch gather plus
    a cuBLAS GEMM take ~0.2 ms).  Anything else takes Hugging Face's forward."""

    def forward(self, pixel_values, interpolate_pos_encoding: bool = False):
        pr = self.projection
        B, Cin, Hh, Ww = pixel_values.shape
        ph, pw = self.patch_size
        if (pixel_values.is_cuda and pixel_values.dtype == torch.bfloat16 and pr.weight.dtype == torch.bfloat16
                and not interpolate_pos_encoding and Cin == self.num_channels and tuple(pr.stride) == (ph, pw)
                and tuple(pr.kernel_size) == (ph, pw) and tuple(pr.padding) == (0, 0) and Hh % ph == 0 and Ww % pw == 0):
            gh, gw = Hh // ph, Ww // pw
            patches = pixel_values.reshape(B, Cin, gh, ph, gw, pw).permute(0, 2, 4, 1, 3, 5).reshape(B * gh * gw, Cin * ph * pw)
            out = torch.nn.functional.linear(patches, pr.weight.reshape(pr.out_channels, -1), pr.bias)
            return out.view(B, gh * gw, pr.out_channels)
        return super().forward(pixel_values, interpolate_pos_encoding)


def use_patch_gemm(module: nn.Module) -> nn.Module:
    """Re-class every Hugging Face ``ViTPatchEmbeddings`` at or below ``module``."""
    for m in module.modules():
        if type(m).__name__ == "ViTPatchEmbeddings":
            m.__class__ = type("PatchEmbeddingsAsGemm_", (PatchEmbeddingsAsGemm, type(m)), {})
    return module


def use_fused_layer_norms(module: nn.Module) -> nn.Module:
    """Re-class every plain ``nn.LayerNorm`` below ``module`` (the ViT blocks are pre-LN: no residual to fuse)."""
    for m in module.modules():
        if type(m) is nn.LayerNorm:
            m.__class__ = FusedLayerNorm
    return module


class GraphFusionLayer(nn.Module):
    def __init__(self, bert_layer, vit_layer, num_bottle_neck_tokens: int, use_projection: bool = False) -> None:
        super().__init__()
        self.bert_encoder = use_fused_intermediates(bert_layer)
        self.vit_encoder = use_fused_vit_layers(use_fused_intermediates(use_mdt_attention(use_fused_layer_norms(vit_layer))))
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


    def forward_packed(self, x_packed, vit_hidden_states, bottle_neck, pack: TextPack, x_image_indexes=None):
        """Same block on the packed layout: ``x_packed`` already holds the text tokens; the bottleneck slots are
        (re)written from ``bottle_neck`` and read back after the BERT layer."""
        nb = self.num_bottle_neck_tokens
        C = x_packed.shape[-1]
        out = packed_bert_layer(self.bert_encoder, ops.replace_rows(x_packed, pack.bn_rows, bottle_neck.reshape(-1, C)), pack,
                                take_rows=pack.bn_rows)
        if isinstance(out, tuple):      # the single-node block hands the bottleneck rows out itself
            x_packed, bn_rows_out = out
        else:
            x_packed, bn_rows_out = out, out.index_select(0, pack.bn_rows)
        bottle_neck_output = bn_rows_out.view(-1, nb, C)
        if vit_hidden_states is not None:
            rows = image_rows(x_image_indexes)
            if isinstance(self.vit_encoder, FusedViTLayer) and x_packed.is_cuda:
                # The ViT tokens live in ONE [n_img, nb + 197, C] buffer across the fusion layers: its bottleneck slots are
                # rewritten in place, the block hands its first nb output tokens out itself, row moves go through
                # mdt_gather_rows / mdt_scatter_rows (ATen's index_add for 380 rows of 3072 elements takes 92 us).
                rows32 = _rows_i32(rows)
                n = bottle_neck.shape[0]
                bn_img = ops.gather_rows(bottle_neck.reshape(n, nb * C), rows32).view(-1, nb, C)
                if getattr(vit_hidden_states, "_mdt_has_bn", False):
                    x_vit = ops.replace_tokens(vit_hidden_states, bn_img)
                else:
                    x_vit = prepend_tokens(bn_img, vit_hidden_states)
                out = self.vit_encoder(x_vit, take_tokens=nb)
                if isinstance(out, tuple):
                    vit_hidden_output, vit_bn = out
                    vit_hidden_output._mdt_has_bn = True
                else:
                    vit_bn, vit_hidden_output = ops.split_tokens(_first(out), nb)
                bno = bottle_neck_output.reshape(n, nb * C)
                mixed = (vit_bn.reshape(-1, nb * C) + ops.gather_rows(bno, rows32)) / 2
                bottle_neck_output = ops.scatter_rows(bno, mixed, rows32).view(n, nb, C)
            else:
                vit_out = self.vit_forward(prepend_tokens(bottle_neck.index_select(0, rows), vit_hidden_states))
                vit_bn, vit_hidden_output = vit_out[:, :nb], vit_out[:, nb:]
                mixed = (vit_bn + bottle_neck_output.index_select(0, rows)) / 2
                bottle_neck_output = bottle_neck_output.index_copy(0, rows, mixed)
        else:
            vit_hidden_output = None
        return x_packed, vit_hidden_output, bottle_neck_output


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

    def forward_packed(self, x_packed, vit_hidden_states, bottle_neck, pack: TextPack, x_image_indexes=None):
        rows = image_rows(x_image_indexes)
        for f_layer in self.fusion_layers:
            x_packed, vit_hidden_states, bottle_neck = f_layer.forward_packed(x_packed, vit_hidden_states, bottle_neck, pack, rows)
        return x_packed, vit_hidden_states, bottle_neck
