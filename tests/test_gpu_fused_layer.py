"""``ops.packed_post_ln_layer`` -- a Hugging Face ``BertLayer`` on packed rows as ONE autograd node with a hand-written
backward -- against (a) the same block as separate autograd nodes over the same kernels (``MDT_FUSED_TEXT_LAYER=0`` path of
``packed_bert_layer``), with and without dropout (same seeds => same masks), and (b) Hugging Face's own fp32 forward /
backward of the layer on padded sequences (dropout off).

Tolerances: (a) the two paths differ only in where bf16 rounding happens (residual gradients accumulated by the GEMM instead
of a separate add; bias gradients summed in fp32 inside the LayerNorm backward instead of a bf16 reduction): outputs equal
bit for bit, gradients <= 1e-2 relative Frobenius; (b) bf16 kernels vs fp32: output <= 6e-2 abs
(4 bf16 ulps at |y| in [2, 4); measured 3.2e-2), gradients <= 3e-2
relative Frobenius (the bar of tests/test_gpu_model.py's fp32-parameter cases is 3e-2 as well)."""

from __future__ import annotations

import pytest
import torch

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.float(), b.float()
    return float((a - b).norm() / (b.norm() + 1e-12))


def make_layer(p_drop, C=768, seed=0):
    from transformers import BertConfig
    from transformers.models.bert.modeling_bert import BertLayer

    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import use_fused_intermediates

    torch.manual_seed(seed)
    cfg = BertConfig(hidden_size=C, num_attention_heads=C // 64, intermediate_size=4 * C, hidden_dropout_prob=p_drop,
                     attention_probs_dropout_prob=p_drop)
    cfg._attn_implementation = "eager"
    layer = BertLayer(cfg)
    for n, p in layer.named_parameters():       # HF zero-initialises biases: make every gradient path visible
        if n.endswith("bias"):
            torch.nn.init.normal_(p, std=0.05)
    return use_fused_intermediates(layer)


def packed_inputs(lens, C, seed=1):
    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import _SeqLens

    g = torch.Generator().manual_seed(seed)
    total = sum(lens)
    x = torch.randn(total, C, generator=g).bfloat16().cuda()
    go = torch.randn(total, C, generator=g).bfloat16().cuda()
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32).cuda()
    return x, go, _SeqLens(cu, max(lens), total)


def run(layer, x, go, pack, fused, seed):
    import multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer as fl

    old = fl.FUSED_TEXT_LAYER
    fl.FUSED_TEXT_LAYER = fused
    try:
        torch.manual_seed(seed)
        xi = x.detach().clone().requires_grad_(True)
        layer.zero_grad(set_to_none=True)
        y = fl.packed_bert_layer(layer, xi, pack)
        y.backward(go)
        grads = {n: p.grad.detach().clone() for n, p in layer.named_parameters()}
        return y.detach(), xi.grad.detach(), grads
    finally:
        fl.FUSED_TEXT_LAYER = old


@pytest.mark.parametrize("p_drop", [0.0, 0.3])
def test_fused_layer_matches_separate_nodes(p_drop):
    from multimodaldiscussiontransformer_b200 import _lib

    lens = [128, 3, 77, 1, 64, 100, 9, 128, 31]
    layer = make_layer(p_drop).cuda().bfloat16().train()
    x, go, pack = packed_inputs(lens, 768)
    before = _lib.CALLS.get("mdt_dropout_add_ln_bwd_colsum", 0)
    y1, dx1, g1 = run(layer, x, go, pack, True, 123)
    assert _lib.CALLS.get("mdt_dropout_add_ln_bwd_colsum", 0) == before + 2      # the single node ran
    y0, dx0, g0 = run(layer, x, go, pack, False, 123)
    assert _lib.CALLS.get("mdt_dropout_add_ln_bwd_colsum", 0) == before + 2
    assert torch.equal(y1, y0)
    assert rel(dx1, dx0) <= 1e-2, rel(dx1, dx0)
    assert set(g1) == set(g0)
    for n in g0:
        assert rel(g1[n], g0[n]) <= 1e-2, (n, rel(g1[n], g0[n]))


def test_fused_layer_matches_hugging_face_fp32():
    lens = [40, 128, 5, 77]
    C = 256
    layer = make_layer(0.0, C=C).cuda()
    x, go, pack = packed_inputs(lens, C)
    # Hugging Face's layer in fp32 on the padded layout with an additive key mask
    ref = make_layer(0.0, C=C).cuda().float().train()
    ref.load_state_dict(layer.state_dict())
    n, L = len(lens), max(lens)
    xp = torch.zeros(n, L, C, device="cuda")
    gp = torch.zeros(n, L, C, device="cuda")
    mask = torch.full((n, 1, 1, L), float("-inf"), device="cuda")
    o = 0
    for i, l in enumerate(lens):
        xp[i, :l], gp[i, :l] = x[o:o + l].float(), go[o:o + l].float()
        mask[i, ..., :l] = 0.0
        o += l
    xp.requires_grad_(True)
    out = ref(xp, attention_mask=mask)
    out = out[0] if isinstance(out, (tuple, list)) else out
    out.backward(gp)
    y_ref = torch.cat([out[i, :l] for i, l in enumerate(lens)]).detach()
    dx_ref = torch.cat([xp.grad[i, :l] for i, l in enumerate(lens)])
    g_ref = {k: p.grad for k, p in ref.named_parameters()}

    layer = layer.bfloat16().train()
    y, dx, g = run(layer, x, go, pack, True, 5)
    assert float((y.float() - y_ref).abs().max()) <= 6e-2
    assert rel(dx, dx_ref) <= 3e-2, rel(dx, dx_ref)
    for k in g_ref:
        if k == "attention.self.key.bias":      # softmax is invariant to a key bias: the exact gradient is 0, bf16 leaves noise
            assert float(g[k].float().abs().max()) <= 2e-2
            continue
        assert rel(g[k], g_ref[k]) <= 3e-2, (k, rel(g[k], g_ref[k]))


# ---------------------------------------------------------------------------------------------------------------------
# pre-LN block (Hugging Face ViTLayer) as one autograd node: ops.vit_pre_ln_layer
# ---------------------------------------------------------------------------------------------------------------------
def make_vit_layer(p_hidden, p_attn, C=768, seed=0):
    from transformers import ViTConfig
    from transformers.models.vit.modeling_vit import ViTLayer

    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import use_fused_intermediates

    torch.manual_seed(seed)
    cfg = ViTConfig(hidden_size=C, num_attention_heads=C // 64, intermediate_size=4 * C, hidden_dropout_prob=p_hidden,
                    attention_probs_dropout_prob=p_attn)
    cfg._attn_implementation = "eager"
    layer = ViTLayer(cfg)
    for n, p in layer.named_parameters():
        if n.endswith("bias"):
            torch.nn.init.normal_(p, std=0.05)
        if "layernorm" in n and n.endswith("weight"):
            torch.nn.init.normal_(p, mean=1.0, std=0.1)
    return use_fused_intermediates(layer)


def vit_call(layer, x):
    from multimodaldiscussiontransformer_b200 import ops

    sa = layer.attention.attention
    return ops.vit_pre_ln_layer(x, sa.num_attention_heads, sa.dropout_prob, layer.output.dropout.p, layer.layernorm_before,
                                sa.query, sa.key, sa.value, layer.attention.output.dense, layer.layernorm_after,
                                layer.intermediate.dense, layer.output.dense, layer.training)


def vit_reference(layer, x, go, keep_attn=None, p_attn=0.0, m1=None, m2=None):
    """The block in fp32 torch with explicit masks (``keep_attn`` bool [B,H,T,T]; ``m1`` / ``m2`` = keep / (1 - p) [B*T, C])."""
    import torch.nn.functional as F

    ref = make_vit_layer(0.0, 0.0, C=x.shape[-1]).cuda().float()
    ref.load_state_dict({k: v.float() for k, v in layer.state_dict().items()})
    sa = ref.attention.attention
    B, T, C = x.shape
    H = sa.num_attention_heads
    xr = x.detach().float().requires_grad_(True)
    h1 = ref.layernorm_before(xr)
    q, k, v = (m(h1).view(B, T, H, 64).transpose(1, 2) for m in (sa.query, sa.key, sa.value))
    pr = torch.softmax(q @ k.transpose(-1, -2) * 64 ** -0.5, dim=-1)
    if keep_attn is not None:
        pr = pr * keep_attn.float() / (1.0 - p_attn)
    att = (pr @ v).transpose(1, 2).reshape(B, T, C)
    o = ref.attention.output.dense(att)
    z2 = (o if m1 is None else o * m1.view(B, T, C)) + xr
    f = ref.output.dense(F.gelu(ref.intermediate.dense(ref.layernorm_after(z2))))
    y = (f if m2 is None else f * m2.view(B, T, C)) + z2
    y.backward(go.float())
    return y.detach(), xr.grad, {n: p.grad for n, p in ref.named_parameters()}


def check_vit(got, want, tol_out, tol_grad):
    y, dx, g = got
    y_ref, dx_ref, g_ref = want
    assert float((y.float() - y_ref).abs().max()) <= tol_out * max(1.0, float(y_ref.abs().max()))
    assert rel(dx, dx_ref) <= tol_grad, ("dx", rel(dx, dx_ref))
    for k_ in g_ref:
        if k_.endswith("key.bias"):                 # exact gradient 0 (softmax is invariant to a key bias)
            assert float(g[k_].float().abs().max()) <= 5e-2
            continue
        assert rel(g[k_], g_ref[k_]) <= tol_grad, (k_, rel(g[k_], g_ref[k_]))


def run_vit(layer, x, go, seed):
    torch.manual_seed(seed)
    xi = x.detach().clone().requires_grad_(True)
    layer.zero_grad(set_to_none=True)
    y = vit_call(layer, xi)
    y.backward(go)
    return y.detach(), xi.grad.detach(), {n: p.grad.detach().clone() for n, p in layer.named_parameters()}


@pytest.mark.parametrize("B,T,C", [(3, 201, 768), (2, 64, 256), (5, 130, 512)])
def test_fused_vit_layer_matches_fp32_formula_without_dropout(B, T, C):
    """bf16 kernels vs the fp32 block: output <= 2e-2 of the largest entry, gradients <= 3e-2 relative Frobenius."""
    layer = make_vit_layer(0.0, 0.0, C=C).cuda().bfloat16().train()
    g = torch.Generator().manual_seed(B * T)
    x = torch.randn(B, T, C, generator=g).bfloat16().cuda()
    go = torch.randn(B, T, C, generator=g).bfloat16().cuda()
    check_vit(run_vit(layer, x, go, 1), vit_reference(layer, x, go), 2e-2, 3e-2)


def test_fused_vit_layer_with_dropout_against_the_same_masks():
    """Dropout 0.3 (hidden) / 0.2 (attention): the reference applies the very masks the kernels generate -- the attention mask
    restated in torch (tests/test_gpu_attention.py), the two hidden masks read back from ``mdt_dropout_add_fwd`` on ones --
    so every dropout site and its backward is checked, not just the statistics."""
    from test_gpu_attention import dropout_keep_mask

    from multimodaldiscussiontransformer_b200 import _lib, ops

    B, T, C, ph, pa = 3, 201, 768, 0.3, 0.2
    layer = make_vit_layer(ph, pa, C=C).cuda().bfloat16().train()
    g = torch.Generator().manual_seed(7)
    x = torch.randn(B, T, C, generator=g).bfloat16().cuda()
    go = torch.randn(B, T, C, generator=g).bfloat16().cuda()
    torch.manual_seed(99)
    seeds = [ops.next_seed() for _ in range(3)]                       # the order vit_pre_ln_layer draws them in
    lib = _lib.load()
    ones = torch.ones(B * T, C, dtype=torch.bfloat16, device="cuda")
    masks = []
    for s_ in seeds[1:]:
        m = torch.empty_like(ones)
        _lib.check(lib.mdt_dropout_add_fwd(_lib.ptr(ones), None, _lib.ptr(m), B * T, C, ph, s_, _lib.stream()), "mdt_dropout_add_fwd")
        masks.append(m.float())
        assert abs(float((m > 0).float().mean()) - (1 - ph)) < 1e-2
    keep = dropout_keep_mask(seeds[0], pa, B, C // 64, T)
    got = run_vit(layer, x, go, 99)
    want = vit_reference(layer, x, go, keep_attn=keep, p_attn=pa, m1=masks[0], m2=masks[1])
    check_vit(got, want, 2e-2, 3e-2)


def test_fusion_layer_routes_vit_through_the_single_node():
    from multimodaldiscussiontransformer_b200 import _lib
    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import GraphFusionLayer

    fl = GraphFusionLayer(make_layer(0.0), make_vit_layer(0.1, 0.1), 4).cuda().bfloat16().train()
    x = torch.randn(2, 201, 768, device="cuda").bfloat16().requires_grad_(True)
    before = _lib.CALLS.get("mdt_dropout_bwd", 0)
    fl.vit_forward(x).sum().backward()
    assert _lib.CALLS.get("mdt_dropout_bwd", 0) == before + 1
    assert x.grad is not None and bool(torch.isfinite(x.grad).all())


def test_patch_embedding_as_gemm_matches_the_convolution():
    from transformers import ViTConfig
    from transformers.models.vit.modeling_vit import ViTPatchEmbeddings

    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import use_patch_gemm

    torch.manual_seed(0)
    pe = ViTPatchEmbeddings(ViTConfig()).cuda().bfloat16()
    x = torch.randn(5, 3, 224, 224, device="cuda").bfloat16()
    want = pe(x).float()
    keys = list(pe.state_dict().keys())
    got = use_patch_gemm(pe)(x).float()
    assert list(pe.state_dict().keys()) == keys and got.shape == want.shape == (5, 196, 768)
    assert float((got - want).abs().max()) <= 2e-2 * max(1.0, float(want.abs().max()))
    want32 = torch.nn.functional.conv2d(x.float(), pe.projection.weight.float(), pe.projection.bias.float(), stride=16)
    assert float((got - want32.flatten(2).transpose(1, 2)).abs().max()) <= 2e-2 * max(1.0, float(want32.abs().max()))
