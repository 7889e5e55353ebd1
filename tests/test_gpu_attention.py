"""GPU parity of the tcgen05 biased-attention kernels (A6, multihead_attention.py:168-199).

The checker is the oracle's formula in fp32 (``softmax(q k^T + bias + key padding) v``) on the
same bf16 inputs.  Tolerances (stated once): forward ``|o - o_ref| <= 2e-2`` absolute on
N(0,1)-scaled values; gradients ``<= 3e-2`` relative Frobenius error (bf16 P / dS operands,
fp32 accumulation and softmax).
"""

from __future__ import annotations

import math

import pytest
import torch

pytestmark = pytest.mark.gpu

FWD_ATOL = 2e-2
GRAD_RTOL = 3e-2
NEG = float("-inf")


@pytest.fixture(scope="module")
def ops():
    from multimodaldiscussiontransformer_b200 import ops as _ops

    return _ops


def make_qkv(B, T, H, seed, layout="bthd", scale=1.0):
    g = torch.Generator().manual_seed(seed)
    if layout == "tbhd":  # the encoder's [T,B,C] activations viewed as [B,T,H,64]
        mk = lambda: (torch.randn(T, B, H, 64, generator=g) * scale).bfloat16().cuda().permute(1, 0, 2, 3)
    elif layout == "fused":  # slices of one [T,B,3C] projection
        big = (torch.randn(T, B, 3, H, 64, generator=g) * scale).bfloat16().cuda()
        return tuple(big[:, :, i].permute(1, 0, 2, 3) for i in range(3))
    else:
        mk = lambda: (torch.randn(B, T, H, 64, generator=g) * scale).bfloat16().cuda()
    q, k, v = mk(), mk(), mk()
    return q * (64 ** -0.5), k, v


def make_bias(B, H, T, seed, frac_masked=0.3):
    g = torch.Generator().manual_seed(seed + 1000)
    bias = torch.randn(B, H, T, T, generator=g) * 0.5
    mask = torch.rand(B, 1, T, T, generator=g) < frac_masked
    idx = torch.arange(T)
    mask[:, :, idx, idx] = False  # the diagonal (distance 0) is never clipped: rows stay non-empty
    return bias.masked_fill(mask, NEG).cuda()


def make_code(B, H, T, seed, frac_masked=0.3):
    g = torch.Generator().manual_seed(seed + 2000)
    pitch = (T + 15) // 16 * 16
    code = torch.randint(0, 23, (B, T, pitch), generator=g, dtype=torch.int64)
    code[:, 0, :] = 254
    code[:, :, 0] = 254
    mask = torch.rand(B, T, pitch, generator=g) < frac_masked
    idx = torch.arange(T)
    mask[:, idx, idx] = False
    code[mask] = 255
    lut = torch.randn(H, 256, generator=g) * 0.5
    return code.to(torch.uint8).cuda(), lut.cuda()


def code_to_bias(code, lut, T):
    B = code.shape[0]
    H = lut.shape[0]
    c = code[:, :, :T].long()
    bias = lut[:, c.reshape(-1)].reshape(H, B, T, T).permute(1, 0, 2, 3).clone()
    bias.masked_fill_((c == 255).unsqueeze(1), NEG)
    return bias


def reference(q, k, v, bias=None, key_pad=None):
    s = torch.einsum("bthd,bshd->bhts", q.float(), k.float())
    if bias is not None:
        s = s + bias.float()
    if key_pad is not None:
        s = s.masked_fill(key_pad[:, None, None, :].bool(), NEG)
    p = torch.softmax(s, dim=-1)
    return torch.einsum("bhts,bshd->bthd", p, v.float())


def key_padding(B, T, seed):
    g = torch.Generator().manual_seed(seed + 3000)
    n_valid = torch.randint(max(1, T // 2), T + 1, (B,), generator=g)
    return (torch.arange(T)[None, :] >= n_valid[:, None]).cuda()


@pytest.mark.parametrize("T", [1, 7, 64, 65, 128, 129, 200, 257, 513, 1025])
@pytest.mark.parametrize("mode", ["none", "f32", "bf16", "code"])
def test_forward_modes(ops, T, mode):
    B, H = (2, 3) if T > 300 else (3, 4)
    q, k, v = make_qkv(B, T, H, seed=T)
    bias = code = lut = None
    if mode in ("f32", "bf16"):
        bias = make_bias(B, H, T, T)
        if mode == "bf16":
            bias = bias.bfloat16()
        ref_bias = bias
    elif mode == "code":
        code, lut = make_code(B, H, T, T)
        ref_bias = code_to_bias(code, lut, T)
    else:
        ref_bias = None
    o = ops.biased_attention(q, k, v, bias=bias, code=code, lut=lut)
    o_ref = reference(q, k, v, ref_bias)
    assert o.shape == (B, T, H, 64) and o.dtype == torch.bfloat16
    err = float((o.float() - o_ref).abs().max())
    assert err <= FWD_ATOL, f"T={T} mode={mode}: max abs err {err}"


@pytest.mark.parametrize("layout", ["tbhd", "fused"])
@pytest.mark.parametrize("T", [33, 129, 300])
def test_forward_strided_layouts_and_key_padding(ops, layout, T):
    B, H = 4, 12
    q, k, v = make_qkv(B, T, H, seed=5 + T, layout=layout)
    bias = make_bias(B, H, T, 7 + T).bfloat16()
    kp = key_padding(B, T, T)
    o = ops.biased_attention(q, k, v, bias=bias, key_padding_mask=kp)
    o_ref = reference(q, k, v, bias, kp)
    assert float((o.float() - o_ref).abs().max()) <= FWD_ATOL
    # memory order follows q: [T,B,H,64]
    assert o.permute(1, 0, 2, 3).is_contiguous()


def test_forward_large_scores_exercise_rescale(ops):
    """Scores far apart between key tiles force the lazy O rescale path."""
    B, T, H = 2, 400, 2
    q, k, v = make_qkv(B, T, H, seed=11, scale=3.0)
    bias = torch.zeros(B, H, T, T)
    bias[..., 256:] += 30.0  # later tiles dominate
    bias[..., 390:] -= 60.0
    bias = bias.cuda()
    o = ops.biased_attention(q, k, v, bias=bias)
    o_ref = reference(q, k, v, bias)
    assert bool(((o.float() - o_ref).abs() <= 2e-2 + 1e-2 * o_ref.abs()).all())


def test_forward_padding_invariance(ops):
    """Padded keys do not change the outputs of real queries (property test)."""
    B, H, T, Tp = 2, 4, 100, 150
    q, k, v = make_qkv(B, Tp, H, seed=3)
    kp = (torch.arange(Tp)[None, :] >= T).expand(B, Tp).cuda()
    o_pad = ops.biased_attention(q, k, v, key_padding_mask=kp)[:, :T]
    o = ops.biased_attention(q[:, :T].contiguous(), k[:, :T].contiguous(), v[:, :T].contiguous())
    assert float((o_pad.float() - o.float()).abs().max()) <= 1e-2


def test_fully_masked_row_is_zero_not_nan(ops):
    B, H, T = 1, 2, 40
    q, k, v = make_qkv(B, T, H, seed=4)
    bias = torch.zeros(B, H, T, T)
    bias[:, :, 5, :] = NEG
    o = ops.biased_attention(q, k, v, bias=bias.cuda())
    assert bool(torch.isfinite(o.float()).all())
    assert float(o[:, 5].float().abs().max()) == 0.0


def test_argument_errors(ops):
    q, k, v = make_qkv(1, 8, 2, seed=0)
    with pytest.raises(TypeError):
        ops.biased_attention(q.float(), k, v)
    with pytest.raises(ValueError):
        ops.biased_attention(q, k, v, bias=torch.zeros(1, 2, 8, 9).cuda())
    with pytest.raises(ValueError):
        ops.biased_attention(q, k, v, bias=torch.zeros(1, 2, 8, 8).cuda(), code=torch.zeros(1, 8, 16, dtype=torch.uint8).cuda(),
                             lut=torch.zeros(2, 256).cuda())
    with pytest.raises(RuntimeError):
        ops.biased_attention(q.cpu(), k.cpu(), v.cpu())


# ------------------------------------------------------------------------------------------ backward
def rel_err(a, b):
    """Relative Frobenius error; the 1e-3 RMS floor keeps exactly-zero references (T == 1) meaningful."""
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-3 * math.sqrt(b.numel())))


def run_backward(ops, q, k, v, bias=None, code=None, lut=None, key_pad=None, seed=0):
    """Returns kernel grads and fp32 autograd grads of the same bf16 inputs."""
    B, T, H, _ = q.shape
    qk, kk, vk = (t.detach().clone().requires_grad_(True) for t in (q, k, v))
    bias_k = bias.detach().clone().requires_grad_(True) if bias is not None else None
    lut_k = lut.detach().clone().requires_grad_(True) if lut is not None else None
    o = ops.biased_attention(qk, kk, vk, bias=bias_k, code=code, lut=lut_k, key_padding_mask=key_pad)
    g = torch.Generator().manual_seed(seed + 77)
    go = torch.randn(B, T, H, 64, generator=g).bfloat16().cuda()
    o.backward(go)

    qr, kr, vr = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    if code is not None:
        lut_r = lut.detach().clone().requires_grad_(True)
        c = code[:, :, :T].long()
        bias_r = lut_r[:, c.reshape(-1)].reshape(H, B, T, T).permute(1, 0, 2, 3).masked_fill((c == 255).unsqueeze(1), NEG)
        leaf = lut_r
    elif bias is not None:
        leaf = bias.detach().float().requires_grad_(True)
        bias_r = leaf
    else:
        leaf = bias_r = None
    o_ref = reference(qr, kr, vr, bias_r, key_pad)
    o_ref.backward(go.float())
    got = dict(dq=qk.grad, dk=kk.grad, dv=vk.grad)
    want = dict(dq=qr.grad, dk=kr.grad, dv=vr.grad)
    if bias is not None:
        got["dbias"], want["dbias"] = bias_k.grad, leaf.grad
    if code is not None:
        got["dlut"], want["dlut"] = lut_k.grad, leaf.grad
    return got, want


@pytest.mark.parametrize("T", [1, 7, 64, 65, 128, 129, 130, 136, 200, 257, 264, 513, 769, 775])   # >= 768: tail-key split with bwd2
@pytest.mark.parametrize("mode", ["none", "f32", "bf16", "code"])
def test_backward_modes(ops, T, mode):
    B, H = (2, 2) if T > 300 else (3, 4)
    q, k, v = make_qkv(B, T, H, seed=100 + T)
    bias = code = lut = None
    if mode in ("f32", "bf16"):
        bias = make_bias(B, H, T, T)
        if mode == "bf16":
            bias = bias.bfloat16()
    elif mode == "code":
        code, lut = make_code(B, H, T, T)
    got, want = run_backward(ops, q, k, v, bias=bias, code=code, lut=lut, seed=T)
    for name in want:
        tol = GRAD_RTOL if not (name == "dbias" and mode == "bf16") else 4e-2
        assert rel_err(got[name], want[name]) <= tol, f"T={T} mode={mode} {name}: {rel_err(got[name], want[name])}"
    if "dbias" in got:  # exactly zero gradient at masked positions
        assert float(got["dbias"].float()[torch.isneginf(bias.float())].abs().sum()) == 0.0
    if "dlut" in got:
        assert float(got["dlut"][:, 255].abs().sum()) == 0.0


@pytest.mark.parametrize("layout", ["tbhd", "fused"])
def test_backward_strided_layouts_and_key_padding(ops, layout):
    B, H, T = 4, 12, 129
    q, k, v = make_qkv(B, T, H, seed=9, layout=layout)
    code, lut = make_code(B, H, T, 5)
    kp = key_padding(B, T, 2)
    got, want = run_backward(ops, q, k, v, code=code, lut=lut, key_pad=kp)
    for name in want:
        assert rel_err(got[name], want[name]) <= GRAD_RTOL, name
    # padded keys receive exactly zero dK / dV
    pad = kp[:, :, None, None].expand_as(got["dk"])
    assert float(got["dk"].float()[pad].abs().max()) == 0.0
    assert float(got["dv"].float()[pad].abs().max()) == 0.0


@pytest.mark.parametrize("T", [100, 129, 300])
def test_wide_spatial_codes_forward_and_gradient_contract(ops, T):
    """include/mdt_b200.h: spatial codes 0..61 are legal in every kernel (the persistent forward indexes 64 entries);
    grad_lut is exact for codes 0..22 and 254, gradients of codes 23..61 are accumulated into entry 22."""
    B, H = 2, 4
    q, k, v = make_qkv(B, T, H, seed=300 + T)
    g = torch.Generator().manual_seed(T)
    pitch = (T + 15) // 16 * 16
    code = torch.randint(0, 62, (B, T, pitch), generator=g, dtype=torch.int64)
    code[:, 0, :] = 254
    code[:, :, 0] = 254
    mask = torch.rand(B, T, pitch, generator=g) < 0.3
    idx = torch.arange(T)
    mask[:, idx, idx] = False
    code[mask] = 255
    code = code.to(torch.uint8).cuda()
    lut = (torch.randn(H, 256, generator=g) * 0.5).cuda()
    o = ops.biased_attention(q, k, v, code=code, lut=lut)
    o_ref = reference(q, k, v, code_to_bias(code, lut, T))
    assert float((o.float() - o_ref).abs().max()) <= FWD_ATOL
    got, want = run_backward(ops, q, k, v, code=code, lut=lut, seed=T)
    for name in ("dq", "dk", "dv"):
        assert rel_err(got[name], want[name]) <= GRAD_RTOL, name
    folded = want["dlut"].clone()
    folded[:, 22] = want["dlut"][:, 22:62].sum(1)
    folded[:, 23:62] = 0
    assert rel_err(got["dlut"], folded) <= GRAD_RTOL


def test_backward_T1025(ops):
    B, H, T = 1, 2, 1025
    q, k, v = make_qkv(B, T, H, seed=21)
    code, lut = make_code(B, H, T, 21)
    got, want = run_backward(ops, q, k, v, code=code, lut=lut)
    for name in want:
        assert rel_err(got[name], want[name]) <= GRAD_RTOL, name


# ---------------------------------------------------------------------------------------------
# attention dropout (multihead_attention.py:196): the kernels' counter-hash mask, restated in torch
# ---------------------------------------------------------------------------------------------
_M32 = 0xFFFFFFFF


def _fmix(h):
    h = h ^ (h >> 16)
    h = (h * 0x85EBCA6B) & _M32
    h = h ^ (h >> 13)
    h = (h * 0xC2B2AE35) & _M32
    return h ^ (h >> 16)


def dropout_keep_mask(seed, p, B, H, T, device="cuda"):
    """bool [B,H,T,T]: csrc/attn_common.cuh drop_salt / drop_tile / drop_lane / drop_keep_w in 64-bit integer torch ops."""
    thr = min(65535, int(p * 65536.0 + 0.5)) << 16
    lo, hi = seed & _M32, (seed >> 32) & _M32
    mul = 0x9E3779B1
    bh = torch.arange(B * H, dtype=torch.int64, device=device)
    salt = ((_fmix((lo ^ ((bh * mul) & _M32)) & _M32) + hi) & _M32).view(B * H, 1)
    t = torch.arange(T, dtype=torch.int64, device=device).view(1, T)

    def tile(which):
        return _fmix(salt ^ (((2 * (t >> 6) + which) * mul) & _M32))

    def lane(c):
        return _fmix(((c + 1) * mul) & _M32)

    row = (tile(0) ^ lane(t & 63)).view(B * H, T, 1)
    col = (tile(1) ^ lane(64 + (t & 63))).view(B * H, 1, T)
    x = ((row ^ col) * mul) & _M32
    return (x >= thr).view(B, H, T, T)


def reference_dropout(q, k, v, bias, key_pad, keep, p):
    s = torch.einsum("bthd,bshd->bhts", q.float(), k.float())
    if bias is not None:
        s = s + bias.float()
    if key_pad is not None:
        s = s.masked_fill(key_pad[:, None, None, :].bool(), NEG)
    pr = torch.softmax(s, dim=-1) * keep.float() / (1.0 - p)
    return torch.einsum("bhts,bshd->bthd", pr, v.float())


@pytest.mark.parametrize("T", [64, 129, 200, 385, 770])
@pytest.mark.parametrize("mode", ["none", "bf16", "code"])
def test_attention_dropout_forward_backward(ops, T, mode):
    """Forward and gradients with dropout_p = 0.3 against the fp32 formula using the very same mask; T = 64 runs the
    single-tile kernel, 129 / 385 the persistent forward plus the tail-key backward kernel, 200 a ragged last tile."""
    B, H, p, seed = 2, 4, 0.3, 0x1234_5678_9ABC
    q, k, v = make_qkv(B, T, H, seed=11 + T)
    kp = key_padding(B, T, T) if T != 129 else None
    bias = code = lut = None
    if mode == "bf16":
        bias = make_bias(B, H, T, T).bfloat16().requires_grad_(True)
        ref_bias = bias.detach().float().requires_grad_(True)
    elif mode == "code":
        code, lut = make_code(B, H, T, T)
        lut.requires_grad_(True)
        lut_r = lut.detach().clone().requires_grad_(True)
        c = code[:, :, :T].long()
        ref_bias = lut_r[:, c.reshape(-1)].reshape(H, B, T, T).permute(1, 0, 2, 3).masked_fill((c == 255).unsqueeze(1), NEG)
    else:
        ref_bias = None
    keep = dropout_keep_mask(seed, p, B, H, T)
    assert abs(float(keep.float().mean()) - (1 - p)) < 2e-2
    qs, ks, vs = (t.detach().requires_grad_(True) for t in (q, k, v))
    o = ops.biased_attention(qs, ks, vs, bias=bias, code=code, lut=lut, key_padding_mask=kp, dropout_p=p, seed=seed)
    qr, kr, vr = (t.detach().float().requires_grad_(True) for t in (q, k, v))
    o_ref = reference_dropout(qr, kr, vr, ref_bias, kp, keep, p)
    assert float((o.float() - o_ref).abs().max()) <= 3e-2
    go = torch.randn(B, T, H, 64, generator=torch.Generator().manual_seed(5)).cuda()
    ins = [qs, ks, vs] + ([bias] if mode == "bf16" else []) + ([lut] if mode == "code" else [])
    ins_r = [qr, kr, vr] + ([ref_bias] if mode == "bf16" else []) + ([lut_r] if mode == "code" else [])
    got = torch.autograd.grad(o, ins, go.bfloat16())
    want = torch.autograd.grad(o_ref, ins_r, go.bfloat16().float())
    for name, a, b_ in zip(("dq", "dk", "dv", "dbias/dlut"), got, want):
        num, den = float((a.float() - b_).norm()), float(b_.norm())
        assert num <= GRAD_RTOL * max(den, 1e-6), (name, num / max(den, 1e-6))
    # dropout_p = 0 through the same entry leaves the deterministic result
    o0 = ops.biased_attention(q, k, v, bias=None if bias is None else bias.detach(), code=code,
                              lut=None if lut is None else lut.detach(), key_padding_mask=kp)
    o0_ref = reference(q, k, v, None if ref_bias is None else ref_bias.detach(), kp)
    assert float((o0.float() - o0_ref).abs().max()) <= FWD_ATOL
