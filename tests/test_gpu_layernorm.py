"""GPU parity of the fused dropout + residual + LayerNorm kernels (csrc/layernorm.cu) against the plain PyTorch fp32
formula ``LayerNorm(dropout(x) + residual)`` -- the tail of the post-LN blocks on the path (HF BertSelfOutput /
BertOutput inside the fusion layers; graphormer_graph_encoder_layer.py:127-130,139-141).

Tolerances: bf16 outputs ``<= 2e-2`` absolute on O(1) values; gradients ``<= 2e-2`` relative Frobenius (fp32 math inside,
bf16 rounding of the saved pre-norm sum and of the outputs).
"""

from __future__ import annotations

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    return float((a.float() - b.float()).norm() / b.float().norm().clamp_min(1e-12))


def reference(x, res, w, b, keep, scale, eps):
    z = x.float() * keep * scale + (res.float() if res is not None else 0.0)
    z = z.bfloat16().float()  # the kernel normalises the rounded sum it saves
    return F.layer_norm(z, (x.shape[-1],), w.float(), b.float(), eps)


@pytest.mark.parametrize("C", [256, 768, 1024])
@pytest.mark.parametrize("rows", [1, 37, 5000])
@pytest.mark.parametrize("with_res", [True, False])
def test_add_layer_norm_no_dropout(C, rows, with_res):
    from multimodaldiscussiontransformer_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(C + rows)
    x = torch.randn(rows, C, device="cuda", generator=g).bfloat16().requires_grad_(True)
    res = torch.randn(rows, C, device="cuda", generator=g).bfloat16().requires_grad_(True) if with_res else None
    w = (1 + 0.1 * torch.randn(C, device="cuda", generator=g)).bfloat16().requires_grad_(True)
    b = (0.1 * torch.randn(C, device="cuda", generator=g)).bfloat16().requires_grad_(True)
    go = torch.randn(rows, C, device="cuda", generator=g).bfloat16()
    y = ops.dropout_add_layer_norm(x, res, w, b, 0.3, training=False, eps=1e-12)
    ins = [t for t in (x, res, w, b) if t is not None]
    got = torch.autograd.grad(y, ins, go)
    xr, rr, wr, br = (t.detach().float().requires_grad_(True) if t is not None else None for t in (x, res, w, b))
    yr = reference(xr, rr, wr, br, 1.0, 1.0, 1e-12)
    want = torch.autograd.grad(yr, [t for t in (xr, rr, wr, br) if t is not None], go.float())
    assert float((y.float() - yr).abs().max()) <= 2e-2
    for a, b_ in zip(got, want):
        assert rel(a, b_) <= 2e-2


@pytest.mark.parametrize("p", [0.1, 0.4])
def test_dropout_add_layer_norm_training(p):
    from multimodaldiscussiontransformer_b200 import ops

    rows, C = 4096, 768
    g = torch.Generator(device="cuda").manual_seed(7)
    x = (torch.randn(rows, C, device="cuda", generator=g) + 3.0).bfloat16().requires_grad_(True)
    res = torch.randn(rows, C, device="cuda", generator=g).bfloat16().requires_grad_(True)
    w = (1 + 0.1 * torch.randn(C, device="cuda", generator=g)).bfloat16().requires_grad_(True)
    b = (0.1 * torch.randn(C, device="cuda", generator=g)).bfloat16().requires_grad_(True)
    go = torch.randn(rows, C, device="cuda", generator=g).bfloat16()
    torch.manual_seed(123)
    y = ops.dropout_add_layer_norm(x, res, w, b, p, training=True, eps=1e-12)
    gx, gr, gw, gb = torch.autograd.grad(y, [x, res, w, b], go)
    torch.manual_seed(123)
    y2 = ops.dropout_add_layer_norm(x, res, w, b, p, training=True, eps=1e-12)
    assert torch.equal(y, y2)                                    # same seed -> same mask
    keep = (gx != 0).float()                                     # d x is exactly 0 where the element was dropped
    assert abs(float(1 - keep.mean()) - p) < 5e-3                # drop rate
    assert abs(float(keep[:, 0].mean()) - float(keep[:, C - 1].mean())) < 5e-2
    xr, rr, wr, br = (t.detach().float().requires_grad_(True) for t in (x, res, w, b))
    yr = reference(xr, rr, wr, br, keep, 1.0 / (1.0 - p), 1e-12)
    want = torch.autograd.grad(yr, [xr, rr, wr, br], go.float())
    assert float((y.float() - yr).abs().max()) <= 3e-2           # the forward used the very mask the backward regenerates
    for a, b_ in zip((gx, gr, gw, gb), want):
        assert rel(a, b_) <= 2e-2


def test_layer_norm_rejects_other_dtypes_and_widths():
    from multimodaldiscussiontransformer_b200 import ops

    x = torch.randn(4, 768, device="cuda")
    w = torch.ones(768, device="cuda")
    with pytest.raises(RuntimeError):
        ops.dropout_add_layer_norm(x, None, w, w)
    xb = torch.randn(4, 640, device="cuda").bfloat16()
    wb = torch.ones(640, device="cuda").bfloat16()
    with pytest.raises(RuntimeError):
        ops.dropout_add_layer_norm(xb, None, wb, wb)


# ---------------------------------------------------------------------------------------------
# N2: GELU kernels (csrc/gelu.cu) and dense -> GELU with the bias gradient folded into the activation's backward
# ---------------------------------------------------------------------------------------------
@pytest.mark.parametrize("rows,C", [(1, 8), (37, 768), (1000, 3072), (4099, 96)])
def test_gelu_forward_backward_vs_torch(rows, C):
    """Exact-erf GELU in fp32 on bf16 tensors: equal to ``F.gelu(u.float()).bfloat16()`` up to one bf16 ulp of the result;
    gradient against fp32 autograd <= 1e-2 relative Frobenius."""
    from multimodaldiscussiontransformer_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(rows + C)
    u = (torch.randn(rows, C, device="cuda", generator=g) * 2.5).bfloat16().requires_grad_(True)
    y = ops.gelu(u)
    ref = torch.nn.functional.gelu(u.detach().float())
    assert float((y.float() - ref).abs().max()) <= 2e-2 and rel(y, ref) <= 4e-3
    go = torch.randn(rows, C, device="cuda", generator=g).bfloat16()
    (gu,) = torch.autograd.grad(y, u, go)
    ur = u.detach().float().requires_grad_(True)
    (gr,) = torch.autograd.grad(torch.nn.functional.gelu(ur), ur, go.float())
    assert rel(gu, gr) <= 1e-2
    # extreme arguments: no NaN, the right limits
    x = torch.tensor([[-50.0, -8.0, -0.0, 0.0, 8.0, 50.0, 1e-4, -1e-4]], device="cuda").bfloat16()
    out = ops.gelu(x).float()
    assert bool(torch.isfinite(out).all()) and float(out[0, 0]) == 0.0 and float(out[0, 5]) == 50.0 and float(out[0, 2]) == 0.0


def test_linear_gelu_and_linear_qkv_match_autograd():
    from multimodaldiscussiontransformer_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(3)
    rows, C, F_ = 777, 256, 1024
    x = torch.randn(rows, C, device="cuda", generator=g).bfloat16().requires_grad_(True)
    lin = torch.nn.Linear(C, F_).cuda().bfloat16()
    y = ops.linear_gelu(x, lin.weight, lin.bias)
    go = torch.randn(rows, F_, device="cuda", generator=g).bfloat16()
    gx, gw, gb = torch.autograd.grad(y, [x, lin.weight, lin.bias], go)
    xr = x.detach().float().requires_grad_(True)
    wr, br = lin.weight.detach().float().requires_grad_(True), lin.bias.detach().float().requires_grad_(True)
    yr = torch.nn.functional.gelu(torch.nn.functional.linear(xr, wr, br))
    rx, rw, rb = torch.autograd.grad(yr, [xr, wr, br], go.float())
    assert rel(y, yr) <= 6e-3
    for a, b in ((gx, rx), (gw, rw), (gb, rb)):
        assert rel(a, b) <= 1.5e-2
    # three projections over the fused parameter storage
    q, k, v = (torch.nn.Linear(C, C).cuda().bfloat16() for _ in range(3))
    keys_before = [p.data_ptr() for p in (q.weight, k.weight, v.weight)]
    want = torch.cat([q(x), k(x), v(x)], dim=-1)
    got = ops.linear_qkv(x, q, k, v)
    assert k.weight.data_ptr() == q.weight.data_ptr() + q.weight.numel() * 2          # adjacent now
    assert torch.equal(got, want)
    go3 = torch.randn_like(got)
    grads = torch.autograd.grad(got, [x, q.weight, k.weight, v.weight, q.bias, k.bias, v.bias], go3)
    ref = torch.autograd.grad(torch.cat([q(x), k(x), v(x)], dim=-1), [x, q.weight, k.weight, v.weight, q.bias, k.bias, v.bias], go3)
    for a, b in zip(grads, ref):
        assert rel(a, b) <= 1e-2
    q.float(), k.float(), v.float()                                                    # re-allocation breaks adjacency ...
    q.bfloat16(), k.bfloat16(), v.bfloat16()
    assert torch.equal(ops.linear_qkv(x, q, k, v), torch.cat([q(x), k(x), v(x)], dim=-1))   # ... and the view is rebuilt
    assert keys_before is not None


@pytest.mark.parametrize("M,N,K", [(1, 256, 64), (777, 768, 256), (5000, 3072, 768), (130, 256, 3072)])
def test_own_gemm_bias_and_gelu_epilogue(M, N, K):
    """``mdt_gemm_bias_act`` (tcgen05 GEMM, csrc/gemm.cu) against fp32 arithmetic: the bias-only output within one bf16
    rounding of ``x @ W^T + b`` (<= 2^-8 relative, the test measures bit equality with cuBLAS on most shapes), the GELU
    epilogue against ``gelu(round_bf16(u))`` with the pre-activation returned as second output; ragged M (masked last tile),
    a non-contiguous x (row stride > K)."""
    from multimodaldiscussiontransformer_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(M + N)
    wide = torch.randn(M, K + 8, device="cuda", generator=g).bfloat16()
    x = wide[:, :K]                                                   # row stride K + 8
    w = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
    b = torch.randn(N, device="cuda", generator=g).bfloat16()
    u_ref = x.float() @ w.float().t() + b.float()
    u = ops.gemm_bias_act(x, w, b)
    assert u.shape == (M, N) and u.dtype == torch.bfloat16
    err = (u.float() - u_ref).abs()
    assert bool((err <= 2 ** -8 * u_ref.abs() + 1e-3).all()), float(err.max())
    act, pre = ops.gemm_bias_act(x, w, b, gelu_act=True)
    assert torch.equal(pre, u)
    want = torch.nn.functional.gelu(u.float())
    assert float((act.float() - want).abs().max()) <= 2 ** -8 * float(want.abs().max()) + 1e-3
    with pytest.raises(RuntimeError):
        ops.gemm_bias_act(x, w[:100], b[:100])                        # N must be a multiple of 256


@pytest.mark.parametrize("M,N,C", [(1, 256, 256), (63, 256, 512), (4128, 768, 768), (50001, 2304, 768), (9000, 256, 1024)])
def test_own_weight_gradient_gemm(M, N, C):
    """``mdt_wgrad_gemm`` (split-K tcgen05, both operands read MN-major): ``dy.t() @ x`` against fp64, <= 2e-5 of the largest
    entry (fp32 accumulation and output; cuBLAS's bf16 result is 3e-3 away); ragged M (zero-filled last row block), strided
    operands, accumulation into an existing fp32 buffer."""
    from multimodaldiscussiontransformer_b200 import ops

    g = torch.Generator(device="cuda").manual_seed(M + N + C)
    wide = torch.randn(M, N + C + 16, device="cuda", generator=g).bfloat16()
    dy, x = wide[:, :N], wide[:, N + 16:]                           # row stride N + C + 16: views, no copies
    want = dy.double().t() @ x.double()
    got = ops.wgrad_gemm(dy, x)
    assert got.dtype == torch.float32 and got.shape == (N, C)
    scale = float(want.abs().max())
    assert float((got.double() - want).abs().max()) <= 2e-5 * scale
    base = torch.randn(N, C, device="cuda", generator=g)
    acc = base.clone()
    ops.wgrad_gemm(dy, x, out=acc)
    assert float((acc.double() - base.double() - want).abs().max()) <= 2e-5 * scale + 1e-5
    # the column sums of dy (bias gradient) from the same pass, accumulated
    cs = torch.full((N,), 0.5, dtype=torch.float32, device="cuda")
    got2 = ops.wgrad_gemm(dy, x, colsum=cs)
    assert torch.equal(got2, got) or float((got2 - got).abs().max()) <= 2e-5 * scale      # atomics: order may differ
    want_cs = dy.double().sum(0) + 0.5
    assert float((cs.double() - want_cs).abs().max()) <= 1e-5 * max(1.0, float(want_cs.abs().max()))
    with pytest.raises(RuntimeError):
        ops.wgrad_gemm(dy[:, :200], x)                                 # N must be a multiple of 256


@pytest.mark.parametrize("M,N,K", [(1, 256, 64), (777, 768, 256), (20000, 3072, 768), (300, 256, 3072)])
def test_own_dgrad_gemm_with_gelu_derivative(M, N, K):
    """``mdt_gemm_dgrad_gelu`` (CTA-pair GEMM, weight consumed MN-major): ``(dy @ w) * gelu'(u)`` against fp32 arithmetic,
    within one bf16 rounding (<= 2^-8 relative + 1e-3)."""
    from multimodaldiscussiontransformer_b200 import _lib

    g = torch.Generator(device="cuda").manual_seed(M + N + K)
    dy = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    w = (torch.randn(K, N, device="cuda", generator=g) * K ** -0.5).bfloat16()      # [K, N] row-major
    u = torch.randn(M, N, device="cuda", generator=g).bfloat16()
    du = torch.empty_like(u)
    lib = _lib.load()
    _lib.check(lib.mdt_gemm_dgrad_gelu(_lib.ptr(dy), dy.stride(0), _lib.ptr(w), w.stride(0), _lib.ptr(u), _lib.ptr(du), M, N, K,
                                       _lib.stream()), "mdt_gemm_dgrad_gelu")
    uf = u.float().requires_grad_(True)
    torch.nn.functional.gelu(uf).backward(dy.float() @ w.float())
    want = uf.grad
    err = (du.float() - want).abs()
    assert bool((err <= 2 ** -8 * want.abs() + 2e-3).all()), float(err.max())
