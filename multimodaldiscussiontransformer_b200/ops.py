"""``torch.autograd.Function`` wrappers over the C ABI of ``libmdt_b200.so``.

PyTorch supplies device memory, streams and the autograd graph; all arithmetic of the hot
path happens in the sm_100a kernels behind ``include/mdt_b200.h``.  Nothing here falls back
to PyTorch operators: a missing library or a CPU tensor raises.
"""

from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import _lib


# When set to a list, the attention wrappers append (kind, start_event, end_event, B, H, T) around
# their C-ABI call on the current stream: bench.py derives the kernels' live in-step durations from it.
# The LayerNorm wrappers append ("ln_fwd" | "ln_bwd", start, end, rows, C, flags) the same way
# (ln_fwd flags: 2 = has residual, 1 = writes z; ln_bwd flags = number of gradient tensors written).
ATTN_EVENTS = None

#: MDT_OWN_WGRAD=0 sends every weight-gradient GEMM to cuBLAS (A/B runs)
OWN_WGRAD = __import__("os").environ.get("MDT_OWN_WGRAD", "1") != "0"


class _timed:
    def __init__(self, kind, B, H, T):
        self.rec = ATTN_EVENTS is not None
        self.meta = (kind, B, H, T)

    def __enter__(self):
        if self.rec:
            self.s = torch.cuda.Event(enable_timing=True)
            self.e = torch.cuda.Event(enable_timing=True)
            self.s.record()

    def __exit__(self, *exc):
        if self.rec:
            self.e.record()
            ATTN_EVENTS.append((self.meta[0], self.s, self.e) + self.meta[1:])


def next_seed() -> int:
    """Seed of an in-kernel counter-hash dropout mask, drawn from torch's CPU generator: ``torch.manual_seed`` reproduces
    every mask, and no device work or synchronisation is involved (a CPU tensor's ``.item()``)."""
    return int(torch.randint(0, 2 ** 62, (1,)).item())


def _require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("libmdt_b200 operators need CUDA tensors (there is no CPU path)")


# ------------------------------------------------------------------------------------------
# A5  GraphAttnBias                                   graphormer_layers.py:86-110
# ------------------------------------------------------------------------------------------
class _AttnBias(torch.autograd.Function):
    @staticmethod
    def forward(ctx, attn_bias, spatial_pos, table, virt):
        _require_cuda(attn_bias, spatial_pos, table, virt)
        lib = _lib.load()
        B, T, _ = attn_bias.shape
        N = T - 1
        H = table.shape[1]
        attn_bias = attn_bias.contiguous()
        sp = spatial_pos.to(torch.int32).contiguous()
        tab = table.detach().float().contiguous()
        vt = virt.detach().float().contiguous().view(-1)
        out = torch.empty((B, H, T, T), dtype=attn_bias.dtype, device=attn_bias.device)
        with torch.cuda.device(attn_bias.device):
            _lib.check(lib.mdt_attn_bias_fwd(_lib.ptr(attn_bias), _lib.ptr(sp), _lib.ptr(tab), _lib.ptr(vt), B, H, N,
                                             tab.shape[0], _lib.dtype_code(attn_bias.dtype), _lib.ptr(out),
                                             _lib.stream()), "mdt_attn_bias_fwd")
        ctx.save_for_backward(sp)
        ctx.meta = (B, H, N, tab.shape[0], table.dtype, virt.dtype, virt.shape)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (sp,) = ctx.saved_tensors
        B, H, N, S, tdt, vdt, vshape = ctx.meta
        lib = _lib.load()
        go = grad_out.contiguous()
        g_table = torch.zeros((S, H), dtype=torch.float32, device=go.device)
        g_virt = torch.zeros((H,), dtype=torch.float32, device=go.device)
        with torch.cuda.device(go.device):
            _lib.check(lib.mdt_attn_bias_bwd(_lib.ptr(go), _lib.ptr(sp), B, H, N, S, _lib.dtype_code(go.dtype),
                                             _lib.ptr(g_table), _lib.ptr(g_virt), _lib.stream()), "mdt_attn_bias_bwd")
        return None, None, g_table.to(tdt), g_virt.to(vdt).view(vshape)


def attn_bias(attn_bias_in: torch.Tensor, spatial_pos: torch.Tensor, table: torch.Tensor, virt: torch.Tensor):
    """``[B,T,T]`` + ``[B,N,N]`` -> materialised ``[B,H,T,T]`` (GraphAttnBias.forward)."""
    return _AttnBias.apply(attn_bias_in, spatial_pos, table, virt)


# ------------------------------------------------------------------------------------------
# A4  GraphNodeFeature                                graphormer_layers.py:39-50
# ------------------------------------------------------------------------------------------
class _NodeFeature(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, in_degree, out_degree, w_in, w_out, token):
        _require_cuda(x, in_degree, out_degree, w_in, w_out, token)
        lib = _lib.load()
        B, N, C = x.shape
        dt = x.dtype
        x = x.contiguous()
        ind = in_degree.to(torch.int64).contiguous()
        outd = out_degree.to(torch.int64).contiguous()
        wi, wo, tk = (w.detach().to(dt).contiguous() for w in (w_in, w_out, token))
        out = torch.empty((B, N + 1, C), dtype=dt, device=x.device)
        with torch.cuda.device(x.device):
            _lib.check(lib.mdt_node_feature_fwd(_lib.ptr(x), _lib.ptr(ind), _lib.ptr(outd), _lib.ptr(wi), _lib.ptr(wo),
                                                _lib.ptr(tk), B, N, C, wi.shape[0], wo.shape[0], _lib.dtype_code(dt),
                                                _lib.ptr(out), _lib.stream()), "mdt_node_feature_fwd")
        ctx.save_for_backward(ind, outd)
        ctx.meta = (B, N, C, w_in.shape, w_out.shape, token.shape, w_in.dtype, w_out.dtype, token.dtype)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        ind, outd = ctx.saved_tensors
        B, N, C, s_in, s_out, s_tok, d_in, d_out, d_tok = ctx.meta
        lib = _lib.load()
        go = grad_out.contiguous()
        gx = torch.empty((B, N, C), dtype=go.dtype, device=go.device) if ctx.needs_input_grad[0] else None
        g_in = torch.zeros(s_in, dtype=torch.float32, device=go.device)
        g_out = torch.zeros(s_out, dtype=torch.float32, device=go.device)
        g_tok = torch.zeros((C,), dtype=torch.float32, device=go.device)
        with torch.cuda.device(go.device):
            _lib.check(lib.mdt_node_feature_bwd(_lib.ptr(go), _lib.ptr(ind), _lib.ptr(outd), B, N, C, s_in[0], s_out[0],
                                                _lib.dtype_code(go.dtype), _lib.ptr(gx), _lib.ptr(g_in), _lib.ptr(g_out),
                                                _lib.ptr(g_tok), _lib.stream()), "mdt_node_feature_bwd")
        return gx, None, None, g_in.to(d_in), g_out.to(d_out), g_tok.to(d_tok).view(s_tok)


def node_feature(x, in_degree, out_degree, w_in, w_out, token):
    """``[B,N,C]`` -> ``[B,N+1,C]`` (GraphNodeFeature.forward)."""
    return _NodeFeature.apply(x, in_degree, out_degree, w_in, w_out, token)


# ------------------------------------------------------------------------------------------
# K12  node-major <-> padded row moves      multigraphormer_graph_encoder.py:363-371,425,435
# ------------------------------------------------------------------------------------------
def _move(scatter: bool, src: torch.Tensor, index: torch.Tensor, dst: torch.Tensor) -> None:
    lib = _lib.load()
    _require_cuda(src, index, dst)
    assert src.dim() == 2 and dst.dim() == 2 and src.stride(1) == 1 and dst.stride(1) == 1
    assert index.dtype == torch.int32 and index.is_contiguous()
    fn = lib.mdt_scatter_rows if scatter else lib.mdt_gather_rows
    with torch.cuda.device(src.device):
        _lib.check(fn(_lib.ptr(src), src.stride(0), _lib.ptr(index), index.numel(), src.shape[1],
                      _lib.dtype_code(src.dtype), _lib.ptr(dst), dst.stride(0), _lib.stream()),
                   "mdt_scatter_rows" if scatter else "mdt_gather_rows")


def _i32(rows: torch.Tensor) -> torch.Tensor:
    """int32 copy of a row list for mdt_gather_rows / mdt_scatter_rows, cached on the tensor (one conversion per batch)."""
    if rows.dtype == torch.int32 and rows.is_contiguous():
        return rows
    r = getattr(rows, "_mdt_i32", None)
    if r is None:
        r = rows.to(torch.int32).contiguous()
        try:
            rows._mdt_i32 = r
        except Exception:
            pass
    return r


class _GatherRows(torch.autograd.Function):
    """``out[r] = src[index[r]]``; ``index`` holds distinct rows."""

    @staticmethod
    def forward(ctx, src, index):
        out = torch.empty((index.numel(), src.shape[1]), dtype=src.dtype, device=src.device)
        _move(False, src, index, out)
        ctx.save_for_backward(index)
        ctx.rows = src.shape[0]
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (index,) = ctx.saved_tensors
        g = torch.zeros((ctx.rows, grad_out.shape[1]), dtype=grad_out.dtype, device=grad_out.device)
        _move(True, grad_out.contiguous(), index, g)
        return g, None


class _ScatterRows(torch.autograd.Function):
    """``out = base.clone(); out[index[r]] = src[r]`` (rows of ``base`` that are overwritten get no gradient)."""

    @staticmethod
    def forward(ctx, base, src, index):
        out = base.clone()
        _move(True, src.contiguous(), index, out)
        ctx.save_for_backward(index)
        return out

    @staticmethod
    def backward(ctx, grad_out):
        (index,) = ctx.saved_tensors
        go = grad_out.contiguous()
        g_src = torch.empty((index.numel(), go.shape[1]), dtype=go.dtype, device=go.device)
        _move(False, go, index, g_src)
        g_base = None
        if ctx.needs_input_grad[0]:
            g_base = go.clone()
            zeros = torch.zeros_like(g_src)
            _move(True, zeros, index, g_base)
        return g_base, g_src, None


def gather_rows(src: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    return _GatherRows.apply(src, index)


def scatter_rows(base: torch.Tensor, src: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    return _ScatterRows.apply(base, src, index)


# ------------------------------------------------------------------------------------------
# A6  biased attention core (tcgen05 / TMEM / TMA)        multihead_attention.py:168-199
# ------------------------------------------------------------------------------------------
def _attn_dims(q, k, v, o) -> "_lib.AttnDims":
    d = _lib.AttnDims()
    d.B, d.T, d.H = q.shape[0], q.shape[1], q.shape[2]
    for name, t in (("q", q), ("k", k), ("v", v), ("o", o)):
        setattr(d, f"{name}_stride_b", t.stride(0))
        setattr(d, f"{name}_stride_t", t.stride(1))
        setattr(d, f"{name}_stride_h", t.stride(2))
    return d


def _check_bthd(name: str, t: torch.Tensor) -> torch.Tensor:
    if t.dtype != torch.bfloat16:
        raise TypeError(f"{name} must be bfloat16 (the tcgen05 kernel computes in bf16 with fp32 accumulation)")
    if t.dim() != 4 or t.shape[-1] != 64:
        raise ValueError(f"{name} must be [B, T, H, 64], got {tuple(t.shape)}")
    ok = t.stride(3) == 1 and all(s % 8 == 0 for s in t.stride()[:3]) and t.data_ptr() % 16 == 0
    return t if ok else t.contiguous()


def transposed_code(code: torch.Tensor) -> torch.Tensor:
    """``code_t[b, k, q] = code[b, q, k]`` (columns ``q >= T``: 255) for the backward's vector code loads.  Built once
    per ``code`` tensor -- every layer and head of a batch shares the same codes -- and cached on the tensor object."""
    ct = getattr(code, "_mdt_code_t", None)
    if ct is None:
        B, T, pitch = code.shape
        ct = torch.full_like(code, _lib.CODE_MASKED)
        ct[:, :, :T] = code[:, :, :T].transpose(1, 2)
        try:
            code._mdt_code_t = ct
        except Exception:  # a tensor subclass that refuses attributes: recompute next time
            pass
    return ct


def _empty_like_layout(q: torch.Tensor) -> torch.Tensor:
    """Output buffer [B,T,H,64] whose memory order follows q's (t-major for the encoder's [T,B,C] activations)."""
    B, T, H, D = q.shape
    if q.stride(1) > q.stride(0):
        return torch.empty((T, B, H, D), dtype=q.dtype, device=q.device).permute(1, 0, 2, 3)
    return torch.empty((B, T, H, D), dtype=q.dtype, device=q.device)


class _BiasedAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, bias, code, lut, key_pad, dropout_p, seed):
        _require_cuda(q, k, v, bias, code, lut, key_pad)
        lib = _lib.load()
        q, k, v = _check_bthd("q", q), _check_bthd("k", k), _check_bthd("v", v)
        B, T, H, _ = q.shape
        if bias is not None:
            if bias.shape != (B, H, T, T):
                raise ValueError(f"bias must be [B,H,T,T]={B, H, T, T}, got {tuple(bias.shape)}")
            bias = bias.contiguous()
        pitch = 0
        if code is not None:
            if code.dtype != torch.uint8 or code.dim() != 3 or code.shape[0] != B or code.shape[1] != T:
                raise ValueError("code must be uint8 [B, T, pitch]")
            code = code.contiguous()
            pitch = code.shape[2]
            lut = lut.detach().float().contiguous()
            if lut.shape != (H, 256):
                raise ValueError("lut must be [H, 256]")
        if key_pad is not None:
            key_pad = key_pad.to(torch.uint8).contiguous()
        out = _empty_like_layout(q)
        lse = torch.empty((B, H, T), dtype=torch.float32, device=q.device)
        dims = _attn_dims(q, k, v, out)
        with torch.cuda.device(q.device), _timed("fwd", B, H, T):
            _lib.check(lib.mdt_biased_attn_fwd(_lib.ptr(q), _lib.ptr(k), _lib.ptr(v), ctypes.byref(dims), _lib.ptr(bias),
                                               _lib.dtype_code(bias.dtype) if bias is not None else 0, _lib.ptr(code),
                                               pitch, _lib.ptr(lut) if code is not None else None, _lib.ptr(key_pad),
                                               _lib.ptr(out), _lib.ptr(lse), float(dropout_p), int(seed), _lib.stream()),
                       "mdt_biased_attn_fwd")
        ctx.save_for_backward(q, k, v, out, lse, bias, code, lut if code is not None else None, key_pad)
        ctx.pitch = pitch
        ctx.drop = (float(dropout_p), int(seed))
        return out

    @staticmethod
    def backward(ctx, grad_out):
        q, k, v, out, lse, bias, code, lut, key_pad = ctx.saved_tensors
        lib = _lib.load()
        B, T, H, _ = q.shape
        go = _check_bthd("grad_out", grad_out)
        views = getattr(ctx, "grad_views", None)   # a caller that wants dq / dk / dv inside one [B, T, 3, H, 64] buffer
        dq, dk, dv = views if views is not None else (_empty_like_layout(q), _empty_like_layout(q), _empty_like_layout(q))
        dq_accum = torch.zeros((B, H, T, 64), dtype=torch.float32, device=q.device)
        delta = torch.empty((B, H, T), dtype=torch.float32, device=q.device)
        g_bias = torch.empty_like(bias) if (bias is not None and ctx.needs_input_grad[3]) else None
        g_lut = torch.zeros((H, 256), dtype=torch.float32, device=q.device) if (
            code is not None and ctx.needs_input_grad[5]) else None
        dims = _attn_dims(q, k, v, out)
        # grad_out is addressed through the "o" strides of gdims; dq/dk/dv through q/k/v strides of gdims
        gdims = _attn_dims(dq, dk, dv, go)
        with torch.cuda.device(q.device), _timed("bwd", B, H, T):
            _lib.check(lib.mdt_biased_attn_bwd(
                _lib.ptr(q), _lib.ptr(k), _lib.ptr(v), _lib.ptr(out), _lib.ptr(go), _lib.ptr(lse), ctypes.byref(dims),
                _lib.ptr(bias), _lib.dtype_code(bias.dtype) if bias is not None else 0, _lib.ptr(code),
                _lib.ptr(transposed_code(code)) if code is not None else None, ctx.pitch, _lib.ptr(lut), _lib.ptr(key_pad), _lib.ptr(dq), _lib.ptr(dk), _lib.ptr(dv), ctypes.byref(gdims),
                _lib.ptr(dq_accum), _lib.ptr(delta), _lib.ptr(g_bias), _lib.ptr(g_lut), ctx.drop[0], ctx.drop[1],
                _lib.stream()), "mdt_biased_attn_bwd")
        return dq, dk, dv, g_bias, None, g_lut, None, None, None


def biased_attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, bias: Optional[torch.Tensor] = None,
                     code: Optional[torch.Tensor] = None, lut: Optional[torch.Tensor] = None,
                     key_padding_mask: Optional[torch.Tensor] = None, dropout_p: float = 0.0,
                     seed: Optional[int] = None) -> torch.Tensor:
    """``dropout(softmax_fp32(q k^T + bias [+ key padding -> -inf])) v`` on the tcgen05 kernel.

    ``q, k, v``: bf16 ``[B, T, H, 64]`` (any strides that are multiples of 8; ``q`` already scaled by
    ``d^-0.5`` as at multihead_attention.py:137).  ``bias``: materialised ``[B,H,T,T]`` (f32 | bf16) **or**
    ``code`` uint8 ``[B,T,pitch]`` with ``lut`` f32 ``[H,256]`` (bias = ``lut[h, code]``, code 255 = -inf).
    ``dropout_p`` drops attention probabilities (multihead_attention.py:196) with a counter-hash mask seeded by ``seed``
    (default: drawn from torch's CPU generator, so ``torch.manual_seed`` reproduces it); the backward regenerates it.
    Returns ``[B, T, H, 64]`` bf16 laid out like ``q``.
    """
    if bias is not None and code is not None:
        raise ValueError("pass either a materialised bias or code+lut")
    if code is not None and lut is None:
        raise ValueError("code needs lut")
    if not 0.0 <= dropout_p < 1.0:
        raise ValueError("dropout_p must be in [0, 1)")
    if seed is None:
        seed = next_seed() if dropout_p > 0.0 else 0
    return _BiasedAttention.apply(q, k, v, bias, code, lut, key_padding_mask, float(dropout_p), int(seed))


# ------------------------------------------------------------------------------------------
# N1  self-attention over packed (unpadded) sequences     multi_graphormer_fusion_layer.py:105-149
# ------------------------------------------------------------------------------------------
class VarlenPlan:
    """Tile plan of the packed-sequence attention (csrc/attn_varlen.cu): which consecutive sequences share a 128-row
    tile and every packed row's key window.  Built on the device from ``cu_seqlens`` without a sync; ``total`` (number of
    packed rows) is host-known like ``max_seqlen`` in any varlen attention API."""

    def __init__(self, cu_seqlens: torch.Tensor, total: int, group: bool = True):
        _require_cuda(cu_seqlens)
        lib = _lib.load()
        assert cu_seqlens.dtype == torch.int32 and cu_seqlens.is_contiguous()
        self.n_seq, self.total = cu_seqlens.numel() - 1, int(total)
        self.data = torch.empty(int(lib.mdt_varlen_attn_plan_words(self.n_seq, self.total)), dtype=torch.int32,
                                device=cu_seqlens.device)
        with torch.cuda.device(cu_seqlens.device):
            _lib.check(lib.mdt_varlen_attn_plan(_lib.ptr(cu_seqlens), self.n_seq, self.total, 1 if group else 0,
                                                _lib.ptr(self.data), _lib.stream()), "mdt_varlen_attn_plan")

    # host-side views (synchronise; tests and diagnostics only)
    def n_items(self) -> int:
        return int(self.data[0])

    def error(self) -> bool:
        return bool(int(self.data[1]))

    def item_rows(self):
        """first packed row of every item, plus ``total``"""
        n = self.n_items()
        return self.data[4:4 + 2 * (n + 1):2].tolist()

    def item_first_seq(self):
        n = self.n_items()
        return self.data[5:5 + 2 * (n + 1):2].tolist()


def varlen_plan(cu_seqlens: torch.Tensor, total: int, group: bool = True) -> VarlenPlan:
    return VarlenPlan(cu_seqlens, total, group)


class _VarlenAttention(torch.autograd.Function):
    @staticmethod
    def forward(ctx, qkv, plan, dropout_p, seed, scale):
        _require_cuda(qkv, plan.data)
        lib = _lib.load()
        if qkv.dtype != torch.bfloat16 or qkv.dim() != 4 or qkv.shape[1] != 3 or qkv.shape[3] != 64:
            raise ValueError(f"qkv must be bf16 [total, 3, H, 64], got {qkv.dtype} {tuple(qkv.shape)}")
        total, _, H, _ = qkv.shape
        if total != plan.total:
            raise ValueError(f"the plan was built for {plan.total} packed rows, qkv has {total}")
        if not (qkv.stride(3) == 1 and qkv.stride(2) == 64 and qkv.stride(1) == H * 64 and qkv.stride(0) % 8 == 0
                and qkv.data_ptr() % 16 == 0):
            qkv = qkv.contiguous()   # rows may be slices of a wider buffer; anything else is compacted
        out = torch.empty((total, H, 64), dtype=qkv.dtype, device=qkv.device)
        lse = torch.empty((H, total), dtype=torch.float32, device=qkv.device)
        rs = qkv.stride(0)
        with torch.cuda.device(qkv.device), _timed("vl_fwd", total, H, plan.n_seq):
            _lib.check(lib.mdt_varlen_attn_fwd(_lib.ptr(qkv[:, 0]), _lib.ptr(qkv[:, 1]), _lib.ptr(qkv[:, 2]), rs, rs, rs, total, H,
                                               plan.n_seq, _lib.ptr(plan.data), _lib.ptr(out), H * 64, _lib.ptr(lse),
                                               float(scale), float(dropout_p), int(seed), _lib.stream()), "mdt_varlen_attn_fwd")
        ctx.save_for_backward(qkv, out, lse, plan.data)
        ctx.plan = plan
        ctx.drop = (float(dropout_p), int(seed), float(scale))
        return out

    @staticmethod
    def backward(ctx, grad_out):
        qkv, out, lse, plan_data = ctx.saved_tensors
        plan = ctx.plan
        lib = _lib.load()
        total, _, H, _ = qkv.shape
        go = grad_out.contiguous()
        dqkv = torch.empty((total, 3, H, 64), dtype=qkv.dtype, device=qkv.device)  # every row is written (sequences tile [0, total))
        rs, gs = qkv.stride(0), 3 * H * 64
        with torch.cuda.device(qkv.device), _timed("vl_bwd", total, H, plan.n_seq):
            _lib.check(lib.mdt_varlen_attn_bwd(_lib.ptr(qkv[:, 0]), _lib.ptr(qkv[:, 1]), _lib.ptr(qkv[:, 2]), rs, rs, rs,
                                               _lib.ptr(out), H * 64, _lib.ptr(go), H * 64, _lib.ptr(lse), total, H,
                                               plan.n_seq, _lib.ptr(plan_data), _lib.ptr(dqkv[:, 0]),
                                               _lib.ptr(dqkv[:, 1]), _lib.ptr(dqkv[:, 2]), gs, gs, gs, ctx.drop[2], ctx.drop[0],
                                               ctx.drop[1], _lib.stream()), "mdt_varlen_attn_bwd")
        return dqkv, None, None, None, None


def varlen_attention(qkv: torch.Tensor, plan: VarlenPlan, dropout_p: float = 0.0, seed: Optional[int] = None,
                     softmax_scale: Optional[float] = None) -> torch.Tensor:
    """Self-attention inside every packed sequence: ``qkv`` bf16 ``[total, 3, H, 64]`` -> ``[total, H, 64]``.

    ``plan = varlen_plan(cu_seqlens, total)`` (once per batch): ``cu_seqlens`` int32 ``[n_seq + 1]`` tiles ``[0, total)`` with
    sequences of at most 128 rows.  The scores are scaled by ``softmax_scale`` (default ``64 ** -0.5``) inside the kernels
    (folded into the exponent), so ``q`` is passed unscaled."""
    if not 0.0 <= dropout_p < 1.0:
        raise ValueError("dropout_p must be in [0, 1)")
    if seed is None:
        seed = next_seed() if dropout_p > 0.0 else 0
    scale = 64 ** -0.5 if softmax_scale is None else float(softmax_scale)
    return _VarlenAttention.apply(qkv, plan, float(dropout_p), int(seed), scale)


# ------------------------------------------------------------------------------------------
# N2  dense -> dropout -> + residual -> LayerNorm tail of the post-LN blocks, one pass each way
#     (HF BertSelfOutput / BertOutput in the fusion layers; graphormer_graph_encoder_layer.py:127-130,139-141)
# ------------------------------------------------------------------------------------------
class _DropoutAddLayerNorm(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, p, eps, seed):
        _require_cuda(x, residual, weight, bias)
        lib = _lib.load()
        C = x.shape[-1]
        x2 = x.reshape(-1, C).contiguous()
        r2 = residual.reshape(-1, C).contiguous() if residual is not None else None
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        need_z = r2 is not None or p > 0.0
        z = torch.empty_like(x2) if need_z else None
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        w, b = weight.detach().contiguous(), bias.detach().contiguous()
        with torch.cuda.device(x.device), _timed("ln_fwd", rows, C, (2 if r2 is not None else 0) + (1 if need_z else 0)):
            _lib.check(lib.mdt_dropout_add_ln_fwd(_lib.ptr(x2), _lib.ptr(r2), _lib.ptr(w), _lib.ptr(b), _lib.ptr(y), _lib.ptr(z),
                                                  _lib.ptr(mean), _lib.ptr(rstd), rows, C, float(p), int(seed), float(eps),
                                                  _lib.stream()), "mdt_dropout_add_ln_fwd")
        ctx.save_for_backward(z if need_z else x2, mean, rstd, w)
        ctx.meta = (float(p), int(seed), x.shape, residual is not None, weight.dtype)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, gy):
        z, mean, rstd, w = ctx.saved_tensors
        p, seed, shape, has_res, wdtype = ctx.meta
        lib = _lib.load()
        C = shape[-1]
        gy2 = gy.reshape(-1, C).contiguous()
        rows = gy2.shape[0]
        need_dx = ctx.needs_input_grad[0]
        need_dres = has_res and ctx.needs_input_grad[1]
        # without dropout d x == d residual: one tensor serves both
        share = p == 0.0 and need_dx and need_dres
        dres = torch.empty_like(gy2) if (need_dres or (share and need_dx)) else None
        dx = torch.empty_like(gy2) if (need_dx and not share) else None
        dgamma = torch.zeros(C, dtype=torch.float32, device=gy.device)
        dbeta = torch.zeros(C, dtype=torch.float32, device=gy.device)
        if dx is None and dres is None:
            dres = torch.empty_like(gy2)
        with torch.cuda.device(gy.device), _timed("ln_bwd", rows, C, (1 if dx is not None else 0) + (1 if dres is not None else 0)):
            _lib.check(lib.mdt_dropout_add_ln_bwd(_lib.ptr(gy2), _lib.ptr(z), _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(w),
                                                  _lib.ptr(dx), _lib.ptr(dres), _lib.ptr(dgamma), _lib.ptr(dbeta), rows, C,
                                                  float(p), int(seed), _lib.stream()), "mdt_dropout_add_ln_bwd")
        g_x = (dres if share else dx).view(shape) if need_dx else None
        g_res = dres.view(shape) if need_dres else None
        return g_x, g_res, dgamma.to(wdtype), dbeta.to(wdtype), None, None, None


def dropout_add_layer_norm(x, residual, weight, bias, p: float = 0.0, training: bool = False, eps: float = 1e-5):
    """``LayerNorm(dropout(x, p) + residual)`` in one pass (bf16 tensors, C in {256, 512, 768, 1024}); ``residual`` may be
    None.  The dropout mask is a counter hash seeded from torch's CPU generator (reproducible under manual_seed)."""
    if x.dtype != torch.bfloat16:
        raise RuntimeError("dropout_add_layer_norm: bf16 only (libmdt_b200 has no other LayerNorm path)")
    p = float(p) if training else 0.0
    seed = next_seed() if p > 0.0 else 0
    return _DropoutAddLayerNorm.apply(x, residual, weight, bias, p, eps, seed)


# ------------------------------------------------------------------------------------------
# N2  GELU (exact erf) and dense -> GELU with the bias gradient folded into the activation's backward
#     graphormer_graph_encoder_layer.py:132-136; HF BertIntermediate / ViTIntermediate
# ------------------------------------------------------------------------------------------
def _gelu_ok(x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype == torch.bfloat16 and x.shape[-1] % 8 == 0


class _Gelu(torch.autograd.Function):
    @staticmethod
    def forward(ctx, u):
        lib = _lib.load()
        u2 = u.reshape(-1, u.shape[-1]).contiguous()
        g = torch.empty_like(u2)
        with torch.cuda.device(u.device):
            _lib.check(lib.mdt_gelu_fwd(_lib.ptr(u2), _lib.ptr(g), u2.shape[0], u2.shape[1], _lib.stream()), "mdt_gelu_fwd")
        ctx.save_for_backward(u2)
        return g.view(u.shape)

    @staticmethod
    def backward(ctx, dg):
        (u2,) = ctx.saved_tensors
        lib = _lib.load()
        dg2 = dg.reshape(u2.shape).contiguous()
        du = torch.empty_like(u2)
        with torch.cuda.device(dg.device):
            _lib.check(lib.mdt_gelu_bwd(_lib.ptr(dg2), _lib.ptr(u2), _lib.ptr(du), None, u2.shape[0], u2.shape[1], _lib.stream()),
                       "mdt_gelu_bwd")
        return du.view(dg.shape)


def gelu(u: torch.Tensor) -> torch.Tensor:
    """Exact (erf) GELU of a bf16 CUDA tensor, arithmetic in fp32 (``F.gelu(u.float()).type_as(u)``)."""
    if not _gelu_ok(u):
        raise RuntimeError("ops.gelu: bf16 CUDA tensors with a last dimension that is a multiple of 8 (no other path)")
    return _Gelu.apply(u)


class _LinearGelu(torch.autograd.Function):
    """``gelu(x @ W^T + b)``: the GEMMs stay cuBLAS (bias in its epilogue); the activation and, in the backward, the
    activation gradient together with the bias gradient (column sums) are one pass each of csrc/gelu.cu."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        lib = _lib.load()
        x2 = x.reshape(-1, x.shape[-1])
        with torch.cuda.device(x.device):
            g, u = _dense_gelu(x2, weight, bias)
        ctx.save_for_backward(x2, weight, u)
        ctx.has_bias = bias is not None
        ctx.bias_dtype = bias.dtype if bias is not None else None
        return g.view(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dg):
        x2, weight, u = ctx.saved_tensors
        lib = _lib.load()
        dg2 = dg.reshape(u.shape).contiguous()
        du = torch.empty_like(u)
        need_b = ctx.has_bias and ctx.needs_input_grad[2]
        dbias = torch.zeros(u.shape[1], dtype=torch.float32, device=u.device) if need_b else None
        with torch.cuda.device(dg.device):
            _lib.check(lib.mdt_gelu_bwd(_lib.ptr(dg2), _lib.ptr(u), _lib.ptr(du), _lib.ptr(dbias), u.shape[0], u.shape[1],
                                        _lib.stream()), "mdt_gelu_bwd")
        dx = (du @ weight).view(*dg.shape[:-1], weight.shape[1]) if ctx.needs_input_grad[0] else None
        dw = du.t() @ x2 if ctx.needs_input_grad[1] else None
        return dx, dw, dbias.to(ctx.bias_dtype) if need_b else None


def linear_gelu(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """``gelu(F.linear(x, weight, bias))`` for bf16 CUDA tensors (exact GELU in fp32)."""
    if not (_gelu_ok(x) and weight.dtype == torch.bfloat16 and weight.shape[0] % 8 == 0):
        raise RuntimeError("ops.linear_gelu: bf16 CUDA tensors, output width a multiple of 8 (no other path)")
    return _LinearGelu.apply(x, weight, bias)


# ------------------------------------------------------------------------------------------
# N2  q / k / v projections as ONE GEMM over a persistent fused parameter view
#     multihead_attention.py:134-136; HF BertSelfAttention.query / key / value
# ------------------------------------------------------------------------------------------
def fused_qkv_storage(q_lin, k_lin, v_lin):
    """Make the weights (and biases) of three ``nn.Linear`` adjacent slices of one buffer and return the fused views
    ``([3C, C], [3C] | None)``.  The parameters keep their names, shapes and identity (state-dict keys unchanged: only
    ``.data`` is re-pointed, once); the fused views feed a single GEMM without a ``torch.cat`` per forward, and the weight
    gradient of that GEMM is handed back as three slices without a copy.  Re-checked on every call (three pointer
    compares): ``module.to(...)`` re-allocates parameters, after which the storage is fused again."""
    ws = (q_lin.weight, k_lin.weight, v_lin.weight)
    n, esz = ws[0].numel(), ws[0].element_size()
    base = ws[0].data_ptr()
    cache = getattr(q_lin, "_mdt_fused", None)
    ok = (cache is not None and cache[0].data_ptr() == base and ws[1].data_ptr() == base + n * esz
          and ws[2].data_ptr() == base + 2 * n * esz and cache[0].dtype == ws[0].dtype)
    if not ok:
        with torch.no_grad():
            wf = torch.cat([w.data for w in ws], dim=0)
            C = ws[0].shape[0]
            for i, w in enumerate(ws):
                w.data = wf[i * C:(i + 1) * C]
            bf = None
            if q_lin.bias is not None:
                bs = (q_lin.bias, k_lin.bias, v_lin.bias)
                bf = torch.cat([b.data for b in bs], dim=0)
                for i, b in enumerate(bs):
                    b.data = bf[i * C:(i + 1) * C]
        cache = (wf, bf)
        q_lin._mdt_fused = cache
    return cache


class _Linear3(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, wq, wk, wv, bq, bk, bv, wf, bf):
        x2 = x.reshape(-1, x.shape[-1])
        y = torch.nn.functional.linear(x2, wf, bf)
        ctx.save_for_backward(x2, wf)
        ctx.has_bias = bf is not None
        return y.view(*x.shape[:-1], wf.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, wf = ctx.saved_tensors
        dy2 = dy.reshape(-1, wf.shape[0])
        C = wf.shape[0] // 3
        dx = (dy2 @ wf).view(*dy.shape[:-1], wf.shape[1]) if ctx.needs_input_grad[0] else None
        dw = dy2.t() @ x2 if any(ctx.needs_input_grad[1:4]) else None
        db = dy2.sum(0) if ctx.has_bias and any(ctx.needs_input_grad[4:7]) else None
        gw = [dw[i * C:(i + 1) * C] if dw is not None else None for i in range(3)]
        gb = [db[i * C:(i + 1) * C] if db is not None else None for i in range(3)]
        return (dx, *gw, *gb, None, None)


def linear_qkv(x: torch.Tensor, q_lin, k_lin, v_lin) -> torch.Tensor:
    """``cat([q_lin(x), k_lin(x), v_lin(x)], -1)`` as one GEMM over the fused parameter storage (no per-call ``cat``)."""
    wf, bf = fused_qkv_storage(q_lin, k_lin, v_lin)
    return _Linear3.apply(x, q_lin.weight, k_lin.weight, v_lin.weight, q_lin.bias, k_lin.bias, v_lin.bias, wf, bf)


def gemm_bias_act(x: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], gelu_act: bool = False):
    """``x @ weight^T + bias`` on this library's tcgen05 GEMM (csrc/gemm.cu); with ``gelu_act`` returns ``(gelu(u), u)``.
    bf16, ``weight.shape[0] % 256 == 0``, ``x.shape[-1] % 64 == 0``."""
    lib = _lib.load()
    _require_cuda(x, weight, bias)
    x2 = x.reshape(-1, x.shape[-1])
    if x2.stride(1) != 1 or x2.stride(0) % 8 != 0:
        x2 = x2.contiguous()
    w = weight if weight.stride(1) == 1 and weight.stride(0) % 8 == 0 else weight.contiguous()
    M, K = x2.shape
    N = w.shape[0]
    out = torch.empty((M, N), dtype=torch.bfloat16, device=x.device)
    aux = torch.empty_like(out) if gelu_act else None
    b = bias.contiguous() if bias is not None else None
    with torch.cuda.device(x.device):
        _lib.check(lib.mdt_gemm_bias_act(_lib.ptr(x2), x2.stride(0), _lib.ptr(w), w.stride(0), _lib.ptr(b), _lib.ptr(out), _lib.ptr(aux),
                                         M, N, K, 1 if gelu_act else 0, _lib.stream()), "mdt_gemm_bias_act")
    return (out, aux) if gelu_act else out


# ------------------------------------------------------------------------------------------
# N2  a whole post-LN transformer block of the fusion layers as ONE autograd node with a hand-written backward
#     (HF BertLayer called from multi_graphormer_fusion_layer.py:94-98 on packed tokens).
#     The arithmetic is the kernels above plus cuBLAS GEMMs; what the single node removes is the autograd glue between
#     them: the two gradient-accumulation adds over [total, C] per block (the residual gradients are accumulated by the
#     input-gradient GEMMs, beta = 1), the bias-gradient reductions of the two output dense layers (column sums inside the
#     LayerNorm backward), the GELU gradient written in place, and ~25 autograd nodes per block.
# ------------------------------------------------------------------------------------------
def _ln_fwd_raw(x2, r2, w, b, p, seed, eps):
    lib = _lib.load()
    rows, C = x2.shape
    y = torch.empty_like(x2)
    need_z = r2 is not None or p > 0.0
    z = torch.empty_like(x2) if need_z else None
    mean = torch.empty(rows, dtype=torch.float32, device=x2.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x2.device)
    with _timed("ln_fwd", rows, C, (2 if r2 is not None else 0) + (1 if need_z else 0)):
        _lib.check(lib.mdt_dropout_add_ln_fwd(_lib.ptr(x2), _lib.ptr(r2), _lib.ptr(w), _lib.ptr(b), _lib.ptr(y), _lib.ptr(z),
                                              _lib.ptr(mean), _lib.ptr(rstd), rows, C, float(p), int(seed), float(eps),
                                              _lib.stream()), "mdt_dropout_add_ln_fwd")
    return y, (z if need_z else x2), mean, rstd


def _ln_bwd_raw(gy2, z, mean, rstd, w, p, seed, dadd=None):
    """-> (dx, dres, dgamma f32, dbeta f32, colsum(dx) f32); without dropout dx IS dres (one tensor).  ``dadd`` joins d z."""
    lib = _lib.load()
    rows, C = gy2.shape
    dres = torch.empty_like(gy2)
    dx = torch.empty_like(gy2) if p > 0.0 else None
    acc = torch.zeros(3, C, dtype=torch.float32, device=gy2.device)
    with _timed("ln_bwd", rows, C, 2 if dx is not None else 1):
        _lib.check(lib.mdt_dropout_add_ln_bwd_colsum(_lib.ptr(gy2), _lib.ptr(z), _lib.ptr(mean), _lib.ptr(rstd), _lib.ptr(w),
                                                     _lib.ptr(dadd), _lib.ptr(dx), _lib.ptr(dres), _lib.ptr(acc[0]), _lib.ptr(acc[1]),
                                                     _lib.ptr(acc[2]), rows, C, float(p), int(seed), _lib.stream()),
                   "mdt_dropout_add_ln_bwd_colsum")
    return (dx if dx is not None else dres), dres, acc[0], acc[1], acc[2]


def _varlen_fwd_raw(qkv, plan, p, seed, scale):
    lib = _lib.load()
    total, _, H, _ = qkv.shape
    out = torch.empty((total, H, 64), dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty((H, total), dtype=torch.float32, device=qkv.device)
    rs = qkv.stride(0)
    with _timed("vl_fwd", total, H, plan.n_seq):
        _lib.check(lib.mdt_varlen_attn_fwd(_lib.ptr(qkv[:, 0]), _lib.ptr(qkv[:, 1]), _lib.ptr(qkv[:, 2]), rs, rs, rs, total, H,
                                           plan.n_seq, _lib.ptr(plan.data), _lib.ptr(out), H * 64, _lib.ptr(lse),
                                           float(scale), float(p), int(seed), _lib.stream()), "mdt_varlen_attn_fwd")
    return out, lse


def _varlen_bwd_raw(qkv, out, lse, plan, go, p, seed, scale):
    lib = _lib.load()
    total, _, H, _ = qkv.shape
    dqkv = torch.empty((total, 3, H, 64), dtype=qkv.dtype, device=qkv.device)  # every row is written (sequences tile [0, total))
    rs, gs = qkv.stride(0), 3 * H * 64
    with _timed("vl_bwd", total, H, plan.n_seq):
        _lib.check(lib.mdt_varlen_attn_bwd(_lib.ptr(qkv[:, 0]), _lib.ptr(qkv[:, 1]), _lib.ptr(qkv[:, 2]), rs, rs, rs,
                                           _lib.ptr(out), H * 64, _lib.ptr(go), H * 64, _lib.ptr(lse), total, H,
                                           plan.n_seq, _lib.ptr(plan.data), _lib.ptr(dqkv[:, 0]),
                                           _lib.ptr(dqkv[:, 1]), _lib.ptr(dqkv[:, 2]), gs, gs, gs, float(scale), float(p),
                                           int(seed), _lib.stream()), "mdt_varlen_attn_bwd")
    return dqkv


def _wgrad(dy2, x2, with_colsum=False):
    """``dy2.t() @ x2`` in the parameter dtype: this library's split-K kernel where it wins over cuBLAS (output columns <= 768,
    both output dimensions multiples of 256; benchmarks/wgrad_microbench.py), cuBLAS otherwise.  ``with_colsum``: returns
    ``(dW, dy2.sum(0) in fp32)``, the sums coming out of the same pass on the own kernel."""
    N, C = dy2.shape[1], x2.shape[1]
    if OWN_WGRAD and N % 256 == 0 and C % 256 == 0 and C <= 768 and dy2.shape[0] >= 32768:
        cs = torch.zeros(N, dtype=torch.float32, device=dy2.device) if with_colsum else None
        dw = wgrad_gemm(dy2, x2, colsum=cs).to(dy2.dtype)
        return (dw, cs) if with_colsum else dw
    if (OWN_WGRAD and not with_colsum and N % 256 == 0 and C % 256 == 0 and N <= 768 and dy2.shape[0] >= 131072):
        # wide outputs (fc2: 768 x 3072) as the transpose of the narrow product the kernel is good at (732 vs cuBLAS 822 us)
        return wgrad_gemm(x2, dy2).t().contiguous().to(dy2.dtype)
    dw = dy2.t() @ x2
    return (dw, dy2.sum(0, dtype=torch.float32)) if with_colsum else dw


#: MDT_OWN_GEMM=0 keeps cuBLAS + the separate GELU pass for dense -> GELU (A/B runs)
OWN_GEMM = __import__("os").environ.get("MDT_OWN_GEMM", "1") != "0"


def _dense_gelu(x2, w, b):
    """``(gelu(u), u)`` with ``u = x2 @ w.T + b``: on the CTA-pair tcgen05 GEMM with the activation in its epilogue (csrc/gemm2.cu;
    1191 vs 1543 us for cuBLAS + the GELU kernel at 238806 x 3072 x 768) when the shape qualifies, else cuBLAS + mdt_gelu_fwd."""
    N, K = w.shape
    if OWN_GEMM and b is not None and N % 256 == 0 and K % 64 == 0 and x2.shape[0] >= 16384 and x2.stride(1) == 1 and x2.stride(0) % 8 == 0:
        return gemm_bias_act(x2, w, b, gelu_act=True)
    lib = _lib.load()
    u = torch.nn.functional.linear(x2, w, b)
    g = torch.empty_like(u)
    _lib.check(lib.mdt_gelu_fwd(_lib.ptr(u), _lib.ptr(g), u.shape[0], u.shape[1], _lib.stream()), "mdt_gelu_fwd")
    return g, u


def _dgrad_gelu(df, w2, u):
    """``(du, db1 | None)`` with ``du = (df @ w2) * gelu'(u)``: on the CTA-pair GEMM with the derivative in its epilogue when the
    shape qualifies (then the bias gradient is left to the weight-gradient kernel's column sums: ``db1 = None``), else cuBLAS +
    ``mdt_gelu_bwd`` (which also yields ``db1``)."""
    lib = _lib.load()
    K, N = w2.shape
    if (OWN_GEMM and N % 256 == 0 and K % 64 == 0 and df.shape[0] >= 16384 and df.stride(1) == 1 and df.stride(0) % 8 == 0
            and w2.stride(1) == 1 and w2.stride(0) % 8 == 0 and u.is_contiguous()):
        du = torch.empty_like(u)
        _lib.check(lib.mdt_gemm_dgrad_gelu(_lib.ptr(df), df.stride(0), _lib.ptr(w2), w2.stride(0), _lib.ptr(u), _lib.ptr(du),
                                           df.shape[0], N, K, _lib.stream()), "mdt_gemm_dgrad_gelu")
        return du, None
    dgl = df @ w2
    db1 = torch.zeros(u.shape[1], dtype=torch.float32, device=u.device)
    _lib.check(lib.mdt_gelu_bwd(_lib.ptr(dgl), _lib.ptr(u), _lib.ptr(dgl), _lib.ptr(db1), u.shape[0], u.shape[1], _lib.stream()),
               "mdt_gelu_bwd")
    return dgl, db1


def _wgrad_bias(du, h, db):
    """``(dW, db)``: the weight gradient, and the bias gradient unless the caller already has it."""
    if db is not None:
        return _wgrad(du, h), db
    return _wgrad(du, h, with_colsum=True)


class _PackedPostLNLayer(torch.autograd.Function):
    """x -> LN1(drop(attn(x Wqkv + b) Wo + bo) + x) = h -> LN2(drop(gelu(h W1 + b1) W2 + b2) + h) on packed rows."""

    @staticmethod
    def forward(ctx, x, plan, cfg, wq, wk, wv, bq, bk, bv, wf, bf, wo, bo, g1, be1, w1, b1, w2, b2, g2, be2):
        H, p_attn, p_hid, eps1, eps2, seeds, take = cfg
        lib = _lib.load()
        lin = torch.nn.functional.linear
        M, C = x.shape
        with torch.cuda.device(x.device):
            qkv = lin(x, wf, bf).view(M, 3, H, 64)
            att, lse = _varlen_fwd_raw(qkv, plan, p_attn, seeds[0], 64 ** -0.5)
            att = att.view(M, C)
            h, z1, mean1, rstd1 = _ln_fwd_raw(lin(att, wo, bo), x, g1, be1, p_hid, seeds[1], eps1)
            g, u = _dense_gelu(h, w1, b1)
            y, z2, mean2, rstd2 = _ln_fwd_raw(lin(g, w2, b2), h, g2, be2, p_hid, seeds[2], eps2)
        ctx.save_for_backward(x, qkv, att, lse, z1, mean1, rstd1, h, u, g, z2, mean2, rstd2, wf, wo, g1, w1, w2, g2)
        ctx.plan, ctx.cfg = plan, cfg
        if take is not None:     # rows of y handed out as a second output: their gradient comes back as a second argument
            return y, y.index_select(0, take)
        return y

    @staticmethod
    def backward(ctx, dy, dtaken=None):
        x, qkv, att, lse, z1, mean1, rstd1, h, u, g, z2, mean2, rstd2, wf, wo, g1, w1, w2, g2 = ctx.saved_tensors
        H, p_attn, p_hid, eps1, eps2, seeds, take = ctx.cfg
        lib = _lib.load()
        M, C = x.shape
        need = ctx.needs_input_grad   # 0 x | 3-5 wq wk wv | 6-8 bq bk bv | 11 wo 12 bo 13 g1 14 be1 15 w1 16 b1 17 w2 18 b2 19 g2 20 be2
        bdt = g1.dtype
        with torch.cuda.device(x.device):
            dy = dy.contiguous()
            if dtaken is not None:
                # the handed-out rows' gradient joins dy in place (16 K rows) instead of autograd's zero-filled [total, C]
                # tensor + index_add + accumulation add; dy is the fresh dx of the block above unless it is a view
                if dy._base is not None:
                    dy = dy.clone()
                # dy[take] += dtaken on distinct rows: gather, add, scatter (ATen's index_add_ takes 230 us for these 16 K rows)
                t32 = _i32(take)
                tmp = torch.empty((t32.numel(), C), dtype=dy.dtype, device=dy.device)
                _move(False, dy, t32, tmp)
                tmp.add_(dtaken.to(dy.dtype))
                _move(True, tmp, t32, dy)
            # ---- feed-forward half (df may alias dh when there is no dropout: df is not read after dh is updated)
            df, dh, dg2, dbe2, db2 = _ln_bwd_raw(dy, z2, mean2, rstd2, g2, p_hid, seeds[2])
            dw2 = _wgrad(df, g) if need[17] else None
            dgl, db1 = _dgrad_gelu(df, w2, u)                          # [M, 4C] = (df W2) * gelu'(u)
            dw1, db1 = _wgrad_bias(dgl, h, db1)
            dh.addmm_(dgl, w1)                                         # residual gradient + input gradient: one GEMM, beta = 1
            del dgl, df
            # ---- attention half
            da, dx, dg1, dbe1, dbo = _ln_bwd_raw(dh, z1, mean1, rstd1, g1, p_hid, seeds[1])
            dwo = _wgrad(da, att) if need[11] else None
            datt = da @ wo
            dqkv = _varlen_bwd_raw(qkv, att.view(M, H, 64), lse, ctx.plan, datt.view(M, H, 64), p_attn, seeds[0], 64 ** -0.5)
            dqkv = dqkv.view(M, 3 * C)
            dwf, dbf = _wgrad(dqkv, x, with_colsum=True)
            dbf = dbf.to(bdt)
            if need[0]:
                dx.addmm_(dqkv, wf)
            else:
                dx = None
        gw = [dwf[i * C:(i + 1) * C] if dwf is not None else None for i in range(3)]
        gb = [dbf[i * C:(i + 1) * C] if dbf is not None else None for i in range(3)]
        return (dx, None, None, *gw, *gb, None, None, dwo, dbo.to(bdt), dg1.to(bdt), dbe1.to(bdt), dw1, db1.to(bdt), dw2,
                db2.to(bdt), dg2.to(bdt), dbe2.to(bdt))


def packed_post_ln_layer(x, plan, heads, p_attn, p_hidden, q_lin, k_lin, v_lin, o_lin, ln1, fc1, fc2, ln2, training,
                         take_rows=None):
    """One Hugging Face ``BertLayer`` (post-LN block) on packed rows ``x`` bf16 ``[total, C]`` (``C = 64 * heads`` in
    {256, 512, 768, 1024}) as a single autograd node; the modules only lend their parameters.  With ``take_rows`` (int64 row
    list) returns ``(y, y[take_rows])``: the rows' gradient re-enters the node directly."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 2 and x.shape[1] == 64 * heads):
        raise RuntimeError("ops.packed_post_ln_layer: bf16 CUDA rows [total, 64 * heads] (no other path)")
    wf, bf = fused_qkv_storage(q_lin, k_lin, v_lin)
    pa, ph = (float(p_attn), float(p_hidden)) if training else (0.0, 0.0)
    seeds = (next_seed() if pa > 0 else 0, next_seed() if ph > 0 else 0, next_seed() if ph > 0 else 0)
    cfg = (int(heads), pa, ph, float(ln1.eps), float(ln2.eps), seeds, take_rows)
    return _PackedPostLNLayer.apply(x.contiguous(), plan, cfg, q_lin.weight, k_lin.weight, v_lin.weight, q_lin.bias, k_lin.bias,
                                    v_lin.bias, wf, bf, o_lin.weight, o_lin.bias, ln1.weight, ln1.bias, fc1.weight, fc1.bias,
                                    fc2.weight, fc2.bias, ln2.weight, ln2.bias)


def wgrad_gemm(dy: torch.Tensor, x: torch.Tensor, out: Optional[torch.Tensor] = None, colsum: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``dy.t() @ x`` for bf16 ``dy [M, N]``, ``x [M, C]`` (N, C multiples of 256) on this library's split-K tcgen05 kernel
    (csrc/wgrad.cu) -> fp32 ``[N, C]``; with ``out`` the product is accumulated into it; ``colsum`` (fp32 ``[N]``) receives
    ``+= dy.sum(0)`` from the same pass."""
    lib = _lib.load()
    _require_cuda(dy, x)
    if dy.dtype != torch.bfloat16 or x.dtype != torch.bfloat16 or dy.dim() != 2 or x.dim() != 2 or dy.shape[0] != x.shape[0]:
        raise ValueError("wgrad_gemm: bf16 dy [M, N] and x [M, C]")
    if dy.stride(1) != 1 or dy.stride(0) % 8 != 0 or dy.data_ptr() % 16 != 0:
        dy = dy.contiguous()
    if x.stride(1) != 1 or x.stride(0) % 8 != 0 or x.data_ptr() % 16 != 0:
        x = x.contiguous()
    M, N = dy.shape
    C = x.shape[1]
    if out is None:
        out = torch.zeros((N, C), dtype=torch.float32, device=dy.device)
    with torch.cuda.device(dy.device):
        _lib.check(lib.mdt_wgrad_gemm(_lib.ptr(dy), dy.stride(0), _lib.ptr(x), x.stride(0), _lib.ptr(out), out.stride(0),
                                      _lib.ptr(colsum), M, N, C, _lib.stream()), "mdt_wgrad_gemm")
    return out


# ------------------------------------------------------------------------------------------
# N2  a whole pre-LN transformer block (Hugging Face ViTLayer, called from multi_graphormer_fusion_layer.py:73-100) as ONE
#     autograd node:  x2 = x + drop(Wo attn(LN1(x)));  y = x2 + drop(W2 gelu(W1 LN2(x2)))
#     What the single node removes from the step: the two residual adds and the two gradient-accumulation adds over
#     [rows, C] per block (LN2 is fused with the first dropout + add, the gradients join inside the LayerNorm backward through
#     its `dadd` input), five bias-gradient reductions, the separate dropout / masked-scale kernels and ~30 autograd nodes.
# ------------------------------------------------------------------------------------------
class _Ctx:
    """Stand-in for an autograd context: lets the static forward / backward of another Function be called directly."""
    needs_input_grad = (True,) * 16

    def save_for_backward(self, *t):
        self.saved_tensors = t


def _dropout_add_raw(x2, r2, p, seed, out_shape=None):
    lib = _lib.load()
    y = torch.empty_like(x2) if out_shape is None else torch.empty(out_shape, dtype=x2.dtype, device=x2.device)
    _lib.check(lib.mdt_dropout_add_fwd(_lib.ptr(x2), _lib.ptr(r2), _lib.ptr(y), x2.shape[0], x2.shape[1], float(p), int(seed),
                                       _lib.stream()), "mdt_dropout_add_fwd")
    return y


def _dropout_bwd_raw(dy2, p, seed, want_dx=True):
    """-> (dx (dy itself when p == 0), colsum(dx) f32)"""
    lib = _lib.load()
    dx = torch.empty_like(dy2) if (want_dx and p > 0.0) else None
    xsum = torch.zeros(dy2.shape[1], dtype=torch.float32, device=dy2.device)
    _lib.check(lib.mdt_dropout_bwd(_lib.ptr(dy2), _lib.ptr(dx), _lib.ptr(xsum), dy2.shape[0], dy2.shape[1], float(p), int(seed),
                                   _lib.stream()), "mdt_dropout_bwd")
    return (dx if dx is not None else dy2), xsum


class _VitPreLNLayer(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, cfg, wq, wk, wv, bq, bk, bv, wf, bf, wo, bo, g1, be1, w1, b1, w2, b2, g2, be2):
        H, p_attn, p_hid, eps1, eps2, seeds, take = cfg
        lib = _lib.load()
        lin = torch.nn.functional.linear
        Bn, T, C = x.shape
        x2d = x.reshape(Bn * T, C)
        with torch.cuda.device(x.device):
            h1, _, mean1, rstd1 = _ln_fwd_raw(x2d, None, g1, be1, 0.0, 0, eps1)
            # the kernels take q pre-scaled by d^-0.5 = 1/8 (multihead_attention.py:137): scale the q rows of the fused weight
            # (0.6 M elements, exact in bf16) instead of the activations (63 M)
            ws, bs = wf.clone(), bf.clone()
            ws[:C].mul_(0.125)
            bs[:C].mul_(0.125)
            qkv = lin(h1, ws, bs).view(Bn, T, 3, H, 64)
            actx = _Ctx()
            att = _BiasedAttention.forward(actx, qkv[:, :, 0], qkv[:, :, 1], qkv[:, :, 2], None, None, None, None, p_attn, seeds[0])
            att2 = att.reshape(Bn * T, C)
            # first residual fused with LN2: z2 = drop(o) + x is the block's inner state, h2 = LN2(z2)
            h2, z2, mean2, rstd2 = _ln_fwd_raw(lin(att2, wo, bo), x2d, g2, be2, p_hid, seeds[1], eps2)
            g, u = _dense_gelu(h2, w1, b1)
            y = _dropout_add_raw(lin(g, w2, b2), z2, p_hid, seeds[2], out_shape=(Bn, T, C))   # a base tensor, not a view:
            #                                                  the next layer rewrites its bottleneck slots in place
        ctx.save_for_backward(x2d, h1, mean1, rstd1, att2, z2, mean2, rstd2, h2, u, g, ws, wo, g1, w1, w2, g2, *actx.saved_tensors[:5])
        ctx.cfg, ctx.shape = cfg, (Bn, T, C)
        if take is not None:      # the first `take` tokens handed out as a second output (see _PackedPostLNLayer)
            return y, y[:, :take].clone()
        return y

    @staticmethod
    def backward(ctx, dy, dtaken=None):
        (x2d, h1, mean1, rstd1, att2, z2, mean2, rstd2, h2, u, g, wf, wo, g1, w1, w2, g2, q, k, v, aout, lse) = ctx.saved_tensors
        H, p_attn, p_hid, eps1, eps2, seeds, take = ctx.cfg
        Bn, T, C = ctx.shape
        lib = _lib.load()
        need = ctx.needs_input_grad   # 0 x | 2-4 wq wk wv | 5-7 bq bk bv | 10 wo 11 bo 12 g1 13 be1 14 w1 15 b1 16 w2 17 b2 18 g2 19 be2
        bdt = g1.dtype
        with torch.cuda.device(dy.device):
            if dtaken is not None:    # joins dy in place when dy is a whole tensor (the fresh dx of the block above)
                if dy._base is not None or not dy.is_contiguous():
                    dy = dy.contiguous().clone() if dy.is_contiguous() else dy.contiguous()
                dy[:, :take].add_(dtaken)
            dy2 = dy.reshape(Bn * T, C).contiguous()
            # ---- feed-forward half: y = drop(f) + z2
            df, db2 = _dropout_bwd_raw(dy2, p_hid, seeds[2])
            dw2 = _wgrad(df, g) if need[16] else None
            dgl, db1 = _dgrad_gelu(df, w2, u)
            del df
            dw1, db1 = _wgrad_bias(dgl, h2, db1)
            dh2 = dgl @ w1
            del dgl
            # LN2 backward; the direct path of the second residual (dy) joins d z2 inside the kernel
            do, dz2, dg2, dbe2, dbo = _ln_bwd_raw(dh2, z2, mean2, rstd2, g2, p_hid, seeds[1], dadd=dy2)
            # ---- attention half: z2 = drop(o) + x
            dwo = _wgrad(do, att2) if need[10] else None
            datt = (do @ wo).view(Bn, T, H, 64)
            actx = _Ctx()
            actx.saved_tensors = (q, k, v, aout, lse, None, None, None, None)
            actx.pitch, actx.drop = 0, (p_attn, seeds[0])
            dqkv = torch.empty((Bn, T, 3, H, 64), dtype=dy.dtype, device=dy.device)
            actx.grad_views = (dqkv[:, :, 0], dqkv[:, :, 1], dqkv[:, :, 2])
            _BiasedAttention.backward(actx, datt)
            dqkv = dqkv.view(Bn * T, 3 * C)      # gradient w.r.t. [q / 8 | k | v]: wf here is the weight with its q rows / 8
            dwf, dbf = _wgrad(dqkv, h1, with_colsum=True)
            dbf = dbf.to(bdt)
            dwf[:C].mul_(0.125)
            dbf[:C].mul_(0.125)
            dh1 = dqkv @ wf
            # LN1 backward (no dropout, no residual inside LN1: z = x); d z2 (the first residual's direct path) joins here
            dx, _, dg1, dbe1, _ = _ln_bwd_raw(dh1, x2d, mean1, rstd1, g1, 0.0, 0, dadd=dz2)
        gw = [dwf[i * C:(i + 1) * C] if dwf is not None else None for i in range(3)]
        gb = [dbf[i * C:(i + 1) * C] if dbf is not None else None for i in range(3)]
        return (dx.view(Bn, T, C) if need[0] else None, None, *gw, *gb, None, None, dwo, dbo.to(bdt), dg1.to(bdt), dbe1.to(bdt), dw1,
                db1.to(bdt), dw2, db2.to(bdt), dg2.to(bdt), dbe2.to(bdt))


def vit_pre_ln_layer(x, heads, p_attn, p_hidden, ln1, q_lin, k_lin, v_lin, o_lin, ln2, fc1, fc2, training, take_tokens=None):
    """One Hugging Face ``ViTLayer`` (pre-LN block) on ``x`` bf16 ``[B, T, C]`` (``C = 64 * heads`` in {256, 512, 768, 1024})
    as a single autograd node; the modules only lend their parameters.  With ``take_tokens = nb`` returns
    ``(y, y[:, :nb] copy)``: the copy's gradient re-enters the node directly."""
    if not (x.is_cuda and x.dtype == torch.bfloat16 and x.dim() == 3 and x.shape[2] == 64 * heads):
        raise RuntimeError("ops.vit_pre_ln_layer: bf16 CUDA [B, T, 64 * heads] (no other path)")
    wf, bf = fused_qkv_storage(q_lin, k_lin, v_lin)
    pa, ph = (float(p_attn), float(p_hidden)) if training else (0.0, 0.0)
    seeds = (next_seed() if pa > 0 else 0, next_seed() if ph > 0 else 0, next_seed() if ph > 0 else 0)
    cfg = (int(heads), pa, ph, float(ln1.eps), float(ln2.eps), seeds, take_tokens)
    return _VitPreLNLayer.apply(x.contiguous(), cfg, q_lin.weight, k_lin.weight, v_lin.weight, q_lin.bias, k_lin.bias, v_lin.bias,
                                wf, bf, o_lin.weight, o_lin.bias, ln1.weight, ln1.bias, fc1.weight, fc1.bias, fc2.weight,
                                fc2.bias, ln2.weight, ln2.bias)


class _ReplaceRows(torch.autograd.Function):
    """``x[rows] = src`` IN PLACE (``x`` is returned, marked dirty).  ``x.index_copy(0, rows, src)`` copies all of ``x`` -- 367 MB
    of packed text per fusion layer to rewrite 16 K bottleneck rows -- and its backward clones the whole gradient to zero
    those rows.  Legal here because nothing saves ``x`` for its backward before the rows are rewritten (the LayerNorm kernels
    save their pre-norm input, not their output); the backward zeroes the rows of the incoming gradient in place when that
    gradient is a whole, contiguous tensor (it is the fresh ``dx`` of the block above), and clones otherwise."""

    @staticmethod
    def forward(ctx, x, rows, src):
        own = x.is_cuda and x.dim() == 2 and x.stride(1) == 1 and x.dtype in (torch.bfloat16, torch.float32)
        if own:
            _move(True, src.contiguous(), _i32(rows), x)
        else:
            x.index_copy_(0, rows, src)
        ctx.mark_dirty(x)
        ctx.save_for_backward(rows)
        ctx.own = own
        return x

    @staticmethod
    def backward(ctx, g):
        (rows,) = ctx.saved_tensors
        g_src = None
        if ctx.needs_input_grad[2]:
            if ctx.own and g.stride(1) == 1:
                g_src = torch.empty((rows.numel(), g.shape[1]), dtype=g.dtype, device=g.device)
                _move(False, g, _i32(rows), g_src)
            else:
                g_src = g.index_select(0, rows)
        g_x = None
        if ctx.needs_input_grad[0]:
            g_x = g if (g._base is None and g.is_contiguous()) else g.clone()
            g_x.index_fill_(0, rows, 0)
        return g_x, None, g_src


def replace_rows(x: torch.Tensor, rows: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    """``x.index_copy(0, rows, src)`` without copying ``x`` (see ``_ReplaceRows``); falls back to the out-of-place form for a
    leaf that requires grad (autograd forbids in-place edits of those)."""
    if x.is_leaf and x.requires_grad:
        return x.index_copy(0, rows, src)
    return _ReplaceRows.apply(x, rows, src)


class _SplitTokens(torch.autograd.Function):
    """``(x[:, :nb], x[:, nb:])``: autograd turns each slice's gradient into a zero-filled tensor of ``x``'s size and adds the two
    (five passes over ``x``); the two gradients are disjoint, so the backward is one buffer and two copies."""

    @staticmethod
    def forward(ctx, x, nb):
        ctx.nb, ctx.shape = nb, x.shape
        return x[:, :nb], x[:, nb:]

    @staticmethod
    def backward(ctx, g0, g1):
        B, T, C = ctx.shape
        ref = g0 if g0 is not None else g1
        out = torch.empty((B, T, C), dtype=ref.dtype, device=ref.device)
        if g0 is not None:
            out[:, :ctx.nb].copy_(g0)
        else:
            out[:, :ctx.nb].zero_()
        if g1 is not None:
            out[:, ctx.nb:].copy_(g1)
        else:
            out[:, ctx.nb:].zero_()
        return out, None


def split_tokens(x: torch.Tensor, nb: int):
    """``x[:, :nb], x[:, nb:]`` of ``[B, T, C]`` with a single-buffer backward."""
    return _SplitTokens.apply(x, nb)


class _ReplaceTokens(torch.autograd.Function):
    """``x[:, :nb] = src`` IN PLACE on ``[B, T, C]`` (the bottleneck slots of the ViT buffer; see ``_ReplaceRows``)."""

    @staticmethod
    def forward(ctx, x, src):
        nb = src.shape[1]
        x[:, :nb].copy_(src)
        ctx.mark_dirty(x)
        ctx.nb = nb
        return x

    @staticmethod
    def backward(ctx, g):
        nb = ctx.nb
        g_src = g[:, :nb].clone() if ctx.needs_input_grad[1] else None
        g_x = None
        if ctx.needs_input_grad[0]:
            g_x = g if (g._base is None and g.is_contiguous()) else g.clone()
            g_x[:, :nb].zero_()
        return g_x, g_src


def replace_tokens(x: torch.Tensor, src: torch.Tensor) -> torch.Tensor:
    if x.is_leaf and x.requires_grad:
        return torch.cat([src, x[:, src.shape[1]:]], dim=1)
    return _ReplaceTokens.apply(x, src)
