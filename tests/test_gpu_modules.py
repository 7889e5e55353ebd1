"""GPU parity of the drop-in modules (A6-A11) against the golden fixtures produced by the
unmodified reference (tests/golden/make_golden.py) -- fp32 parameters, bf16 attention core.

Tolerances: outputs / logits ``<= 2e-2`` absolute, gradients ``<= 3e-2`` relative Frobenius
(bf16 kernel vs fp32 reference)."""

from __future__ import annotations

import math
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
G = os.path.join(HERE, "golden")
for p in (G, os.path.join(HERE, "..", "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import cases  # noqa: E402
import mdt_oracle as orc  # noqa: E402  (checker only)

OUT_ATOL = 2e-2
GRAD_RTOL = 3e-2


def load(name):
    return np.load(os.path.join(G, name))


def rel(a, b):
    a = torch.as_tensor(a).float().cpu()
    b = torch.as_tensor(b).float().cpu()
    return float((a - b).norm() / (b.norm() + 1e-6 * math.sqrt(b.numel())))


def cuda(d):
    return {k: (v.cuda() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}


def test_multihead_attention_module_vs_reference_golden():
    from multimodaldiscussiontransformer_b200.modules import MultiheadAttention

    gold = load("graph_modules.npz")
    inp = cuda(cases.graph_module_inputs())
    C, H = inp["x"].shape[-1], 4
    mha = MultiheadAttention(C, H, self_attention=True)
    orc.seed_parameters(mha, 33, std=0.08)
    mha.cuda().eval()
    bias = torch.from_numpy(gold["gab.out"]).cuda().requires_grad_(True)
    q = inp["query"].clone().requires_grad_(True)
    o, w = mha(q, q, q, bias, inp["pad_mask"], need_weights=False)
    assert w is None and o.shape == q.shape
    assert float((o.detach().cpu() - torch.from_numpy(gold["mha.out"])).abs().max()) <= OUT_ATOL
    o.backward(inp["grad_out"])
    assert rel(q.grad, gold["mha.dq"]) <= GRAD_RTOL
    assert rel(bias.grad, gold["mha.dbias"]) <= GRAD_RTOL
    for n_, p_ in mha.named_parameters():
        ref = gold[f"mha.grad.{n_}"]
        if np.abs(ref).max() < 1e-6:  # k_proj.bias: softmax is shift invariant, gradient ~ 0
            assert float(p_.grad.abs().max()) < 1e-2
        else:
            assert rel(p_.grad, ref) <= GRAD_RTOL, n_
    o2, _ = mha(inp["query"], None, None, None, need_weights=False)
    assert float((o2.detach().cpu() - torch.from_numpy(gold["mha.out_nobias"])).abs().max()) <= OUT_ATOL
    with pytest.raises(NotImplementedError):
        mha(q, q, q, bias, need_weights=True)


@pytest.mark.parametrize("tag,pre_ln", [("postln", False), ("preln", True)])
def test_encoder_layer_vs_reference_golden(tag, pre_ln):
    from multimodaldiscussiontransformer_b200.modules import GraphEncoderStack, GraphormerGraphEncoderLayer

    gold = load("graph_modules.npz")
    inp = cuda(cases.graph_module_inputs())
    C, H = inp["x"].shape[-1], 4
    bias = torch.from_numpy(gold["gab.out"]).cuda()
    ly = GraphormerGraphEncoderLayer(C, 96, H, 0.0, 0.0, 0.0, "gelu", pre_layernorm=pre_ln)
    orc.seed_parameters(ly, 34, std=0.08)
    ly.cuda().eval()
    q = inp["query"].clone().requires_grad_(True)
    o, _ = ly(q, bias, None, inp["pad_mask"])
    assert float((o.detach().cpu() - torch.from_numpy(gold[f"layer_{tag}.out"])).abs().max()) <= OUT_ATOL
    o.backward(inp["grad_out"])
    assert rel(q.grad, gold[f"layer_{tag}.dx"]) <= GRAD_RTOL
    assert rel(ly.fc1.weight.grad, gold[f"layer_{tag}.d_fc1"]) <= GRAD_RTOL
    if not pre_ln:
        st = GraphEncoderStack(2, C, 96, H, 0.0, 0.0, 0.0, "relu")
        orc.seed_parameters(st, 35, std=0.08)
        st.cuda().eval()
        x, _ = st(inp["query"], bias, None, inp["pad_mask"])
        assert float((x.detach().cpu() - torch.from_numpy(gold["stack_relu.out"])).abs().max()) <= 3e-2


def build_encoder(name):
    from transformers import BertConfig, ViTConfig

    from multimodaldiscussiontransformer_b200.modules import MultiGraphormerGraphEncoder

    c = cases.ENCODER_CASES[name]
    kw = dict(cases.COMMON_MODEL)
    m = dict(c["model"])
    bc = BertConfig(num_labels=2, num_hidden_layers=m.pop("bert_layers"), hidden_dropout_prob=0.0,
                    attention_probs_dropout_prob=0.0)
    vc = ViTConfig(num_hidden_layers=m.pop("vit_layers"), hidden_dropout_prob=0.0, attention_probs_dropout_prob=0.0)
    kw.update(m)
    enc = MultiGraphormerGraphEncoder(multi_hop_max_dist=5, dropout=0.0, attention_dropout=0.0, activation_dropout=0.0,
                                      bert_config=bc, vit_config=vc, **kw)
    orc.seed_parameters(enc, 41)
    return enc.cuda().eval()


def head(enc, text, bn):
    for layer in (enc.text_pooler, enc.text_dropout, enc.node_classifier):
        text, bn = layer(text), layer(bn)
    return (text + bn) / 2


@pytest.mark.parametrize("fused", [True, False])
@pytest.mark.parametrize("name", ["tiny", "tiny_stack2", "tiny_noimg"])
def test_encoder_logits_loss_grads_vs_reference_golden(name, fused):
    from multimodaldiscussiontransformer_b200.data.collator import collator

    gold = load("encoder.npz")
    enc = build_encoder(name)
    enc.fused_bias = fused
    items, spmax = cases.encoder_items(name)
    batch = collator(items, spmax)
    text, bn, glob = enc(batch)
    out_all = head(enc, text, bn)
    loss, n = orc.node_cross_entropy(out_all, batch["y"], batch["y_mask"])
    assert float((out_all.detach().cpu() - torch.from_numpy(gold[f"{name}.out_all"])).abs().max()) <= OUT_ATOL
    assert float((glob.detach().cpu() - torch.from_numpy(gold[f"{name}.global"])).abs().max()) <= 5e-2
    assert abs(float(loss) - float(gold[f"{name}.loss"])) <= 2e-2 * max(1.0, abs(float(gold[f"{name}.loss"])))
    loss.backward()
    gab, gnf = enc.graph_attn_bias, enc.graph_node_feature
    checks = {
        "spatial_pos_encoder": (gab.spatial_pos_encoder.weight.grad[1:24], gold[f"{name}.grad.spatial_pos_encoder"][1:]),
        "virtual_distance": (gab.graph_token_virtual_distance.weight.grad, gold[f"{name}.grad.virtual_distance"]),
        "bottle_neck": (enc.bottle_neck.weight.grad, gold[f"{name}.grad.bottle_neck"]),
        "in_degree_encoder": (gnf.in_degree_encoder.weight.grad[1:8], gold[f"{name}.grad.in_degree_encoder"][1:]),
        "q_proj0": (enc.layers[0].layers[0].self_attn.q_proj.weight.grad[:4], gold[f"{name}.grad.q_proj0"]),
    }
    for k, (got, want) in checks.items():
        assert rel(got, want) <= 5e-2, (k, rel(got, want))
    # padding rows never receive gradient (nn.Embedding padding_idx semantics)
    assert float(gab.spatial_pos_encoder.weight.grad[0].abs().sum()) == 0.0
    assert float(gnf.in_degree_encoder.weight.grad[0].abs().sum()) == 0.0
    # the graph stack that the reference never runs stays without gradient
    dead = enc.layers[len(enc.fusion_layers) - 1]
    assert all(p.grad is None or float(p.grad.abs().sum()) == 0.0 for p in dead.parameters())


def test_reference_format_batch_matches_fast_path():
    """A batch in the reference's own format (bool masks, materialised attn_bias) gives the same outputs."""
    from multimodaldiscussiontransformer_b200.data.collator import collator

    enc = build_encoder("tiny")
    items, spmax = cases.encoder_items("tiny")
    fast = collator(items, spmax)
    ref_fmt = {k: v for k, v in fast.items() if k not in ("bias_code", "node_index", "n_nodes", "x_nodes", "x_image_rows")}
    with torch.no_grad():
        a = enc(fast)
        b = enc(ref_fmt)
    for x, y in zip(a, b):
        assert float((x - y).abs().max()) <= 2e-2


@pytest.mark.parametrize("name", ["tiny", "tiny_noimg"])
def test_packed_text_fusion_matches_padded_path(name):
    """the BERT side of the fusion layers on packed (unpadded) tokens gives the same logits, bottleneck
    tokens, graph embedding, real-token hidden states and gradients as the padded Hugging Face path (bf16 model)."""
    from multimodaldiscussiontransformer_b200.data.collator import collator

    enc = build_encoder(name).to(torch.bfloat16)
    items, spmax = cases.encoder_items(name)
    batch = collator(items, spmax)
    assert batch["x_text_pack"] is not None
    res = {}
    for packed in (False, True):
        enc.pack_text = packed
        enc.zero_grad(set_to_none=True)
        text, bn, glob = enc(batch)
        out_all = head(enc, text, bn)
        loss, _ = orc.node_cross_entropy(out_all.float(), batch["y"], batch["y_mask"])
        loss.backward()
        res[packed] = dict(text=text.float(), bn=bn.float(), glob=glob.float(), out=out_all.float(), loss=float(loss),
                           g_bn=enc.bottle_neck.weight.grad.float().clone(),
                           g_q=enc.fusion_layers[0].fusion_layers[0].bert_encoder.attention.self.query.weight.grad.float().clone(),
                           g_sp=enc.graph_attn_bias.spatial_pos_encoder.weight.grad.float().clone())
    a, b = res[False], res[True]
    att = batch["x_nodes"][2].bool()
    assert float((a["out"] - b["out"]).abs().max()) <= 3e-2
    assert float((a["bn"] - b["bn"]).abs().max()) <= 6e-2
    assert float((a["glob"] - b["glob"]).abs().max()) <= 6e-2
    # real tokens agree: two bf16 evaluation orders of the same layers (SDPA vs flash-attn, ATen vs fused LayerNorm);
    # hidden states reach |x| ~ 10 where one bf16 ulp is 0.0625 => relative Frobenius 1e-2, worst element 0.13
    assert rel(b["text"][att], a["text"][att]) <= 1e-2
    assert float((a["text"][att] - b["text"][att]).abs().max()) <= 0.13
    assert float(b["text"][~att].abs().max()) == 0.0                            # padding positions come back as zeros
    assert abs(a["loss"] - b["loss"]) <= 2e-2 * max(1.0, abs(a["loss"]))
    for k in ("g_bn", "g_q", "g_sp"):
        assert rel(b[k], a[k]) <= 8e-2, (k, rel(b[k], a[k]))
