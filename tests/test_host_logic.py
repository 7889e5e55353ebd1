"""CPU checks of the host-side mirror of the reference interface (no kernel launches)."""

from __future__ import annotations

import json
import os
import sys
from argparse import Namespace

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
G = os.path.join(HERE, "golden")
for p in (G, os.path.join(HERE, "..", "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import cases  # noqa: E402
import mdt_oracle as orc  # noqa: E402


def small_encoder():
    from transformers import BertConfig, ViTConfig

    from multimodaldiscussiontransformer_b200.modules import MultiGraphormerGraphEncoder

    kw = dict(cases.COMMON_MODEL)
    kw.update(num_fusion_layers=1, num_fusion_stack=1, num_graph_stack=1)
    return MultiGraphormerGraphEncoder(multi_hop_max_dist=5, bert_config=BertConfig(num_labels=2, num_hidden_layers=3),
                                       vit_config=ViTConfig(num_hidden_layers=3), **kw)


def test_state_dict_keys_and_shapes_match_reference():
    keys = json.load(open(os.path.join(G, "encoder_state_keys.json")))
    mine = {k: list(v.shape) for k, v in small_encoder().state_dict().items()}
    assert mine == keys


def test_registry_names_match_reference():
    import multimodaldiscussiontransformer_b200.criterions  # noqa: F401
    import multimodaldiscussiontransformer_b200.models  # noqa: F401
    from multimodaldiscussiontransformer_b200._registry import REGISTRY

    assert "multi_graphormer" in REGISTRY["model"]
    assert {"multi_graphormer", "multi_graphormer_base"} <= set(REGISTRY["arch"])
    assert "node_cross_entropy" in REGISTRY["criterion"]


def test_arch_defaults():
    from multimodaldiscussiontransformer_b200.models import base_architecture, graphormer_base_architecture

    a = Namespace()
    graphormer_base_architecture(a)
    assert (a.encoder_embed_dim, a.encoder_ffn_embed_dim, a.encoder_layers, a.encoder_attention_heads) == (80, 80, 12, 8)
    assert a.num_bottleneck_tokens == 4 and a.num_fusion_layers == 4 and a.activation_fn == "gelu"
    b = Namespace()
    base_architecture(b)
    assert (b.encoder_embed_dim, b.encoder_ffn_embed_dim, b.dropout, b.act_dropout) == (1024, 4096, 0.1, 0.0)


def test_fused_bias_lut_equals_reference_bias():
    """code + lut (what the attention kernel consumes) reproduces GraphAttnBias.forward exactly."""
    from multimodaldiscussiontransformer_b200.modules import GraphAttnBias

    items, spmax = cases.collate_items("ragged_sp5")
    ref = orc.batch_to_torch(orc.collator(items, spmax))
    gab = GraphAttnBias(12, 512 * 9, 512 * 3, 512, 128, 768, "multi_hop", 5, 2)
    want = orc.graph_attn_bias(ref["attn_bias"], ref["spatial_pos"], gab.spatial_pos_encoder.weight,
                               gab.graph_token_virtual_distance.weight)
    # byte codes as the collate kernel defines them (include/mdt_b200.h): spatial_pos value, 254 virtual, 255 masked
    B, T, _ = ref["attn_bias"].shape
    code = torch.full((B, T, T), 254, dtype=torch.int64)
    code[:, 1:, 1:] = ref["spatial_pos"].long()
    code[torch.isneginf(ref["attn_bias"])] = 255
    fused = gab.forward_fused({"bias_code": code.to(torch.uint8)})
    got = fused.materialise()
    assert torch.equal(torch.isneginf(got), torch.isneginf(want))
    fin = ~torch.isneginf(want)
    assert torch.allclose(got[fin], want[fin].float(), atol=1e-7)
    got[fin].sum().backward()
    assert float(gab.spatial_pos_encoder.weight.grad[0].abs().sum()) == 0.0  # padding_idx row: no gradient


def test_host_batch_packing_matches_reference_collator():
    from multimodaldiscussiontransformer_b200.data.collator import HostBatch

    items, spmax = cases.collate_items("ragged_sp5")
    ref = orc.collator(items, spmax)
    hb = HostBatch(items, pin=False)
    assert (hb.B, hb.N) == ref["x"].shape[:2]
    mask = np.asarray(ref["x_token_mask"])
    assert hb.n_real == int(mask.sum())
    text = hb.tensors["text"].numpy()
    np.testing.assert_array_equal(text[0], np.asarray(ref["x"])[mask])
    np.testing.assert_array_equal(text[2], np.asarray(ref["x_attention_mask"])[mask])
    np.testing.assert_array_equal(hb.tensors["flags"][0].numpy(), np.asarray(ref["x_image_indexes"]))
    np.testing.assert_array_equal(hb.tensors["y"].numpy(), np.asarray(ref["y"]))
    assert hb.nbytes() > 0


def test_no_cpu_fallback():
    from multimodaldiscussiontransformer_b200 import ops
    from multimodaldiscussiontransformer_b200.data import algos

    q = torch.zeros(1, 4, 2, 64, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError):
        ops.biased_attention(q, q, q)
    with pytest.raises(RuntimeError):
        algos.floyd_warshall_batched(torch.zeros(1, 4, 4, dtype=torch.uint8))


def test_criterion_matches_oracle_loss():
    from multimodaldiscussiontransformer_b200.criterions import GraphPredictionNodeCrossEntropy

    g = torch.Generator().manual_seed(0)
    logits = torch.randn(9, 2, generator=g)
    y = torch.tensor([1.0, 0.0, 1.0])
    ym = torch.zeros(9, dtype=torch.bool)
    ym[[1, 4, 7]] = True
    crit = GraphPredictionNodeCrossEntropy(None, 1.5, 1.0)
    model = lambda batched_data: (logits, None)
    loss, n, log = crit(model, {"net_input": {"batched_data": {"x": torch.zeros(1, 9, 1), "y": y, "y_mask": ym}}})
    want, n2 = orc.node_cross_entropy(logits, y, ym, 1.5, 1.0)
    assert n == n2 == 3 and abs(float(loss) - float(want)) < 1e-6
    agg = crit.aggregate([log])
    assert 0.0 <= agg["accuracy"] <= 1.0 and "f1" in agg


def test_text_pack_round_trip_and_host_lengths():
    """the packed (unpadded) token layout of the fusion layers -- index construction is plain torch, so it
    is checked here on CPU: every real token and bottleneck slot lands in exactly one packed row and comes back."""
    from multimodaldiscussiontransformer_b200.data import synthetic as syn
    from multimodaldiscussiontransformer_b200.data.collator import HostBatch
    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import TextPack

    items = syn.synthetic_items([5, 9, 3], 11, with_images=False, as_torch=False)
    hb = HostBatch(items, pin=False)
    att = hb.tensors["text"][2] != 0                                  # [n_real, L]
    lens = att.sum(1).to(torch.int32)
    assert hb.text_prefix and hb.text_tokens == int(lens.sum()) and hb.text_max == int(lens.max())
    assert torch.equal(hb.tensors["ints"][-hb.n_real:], lens)         # lengths ride at the end of the int32 staging buffer
    n, L, nb, C = hb.n_real, hb.L, 4, 8
    pack = TextPack(lens, hb.text_tokens, hb.text_max, nb, L)
    assert pack.total == hb.text_tokens + nb * n and pack.max_seqlen == hb.text_max + nb
    assert torch.equal(pack.cu_seqlens[1:] - pack.cu_seqlens[:-1], lens + nb)
    assert torch.equal(pack.text_only.cu_seqlens[1:] - pack.text_only.cu_seqlens[:-1], lens)
    rows = torch.cat([pack.text_rows, pack.bn_rows])
    assert rows.numel() == pack.total and torch.equal(rows.sort().values, torch.arange(pack.total))
    g = torch.Generator().manual_seed(0)
    text = torch.randn(n, L, C, generator=g) * att[:, :, None]
    bn = torch.randn(n, nb, C, generator=g)
    x = pack.pack_tokens(pack.gather_text(text), bn)
    assert torch.equal(pack.unpack_text(x), text)                     # padding positions come back as zeros
    assert torch.equal(x.index_select(0, pack.bn_rows).view(n, nb, C), bn)
    # sequence i occupies [cu[i], cu[i+1]): bottleneck slots first, then its tokens in order
    i = 1
    lo = int(pack.cu_seqlens[i])
    assert torch.equal(x[lo:lo + nb], bn[i]) and torch.equal(x[lo + nb:lo + nb + int(lens[i])], text[i, :int(lens[i])])
    # a mask with a hole is not a prefix: the collator then does not offer the packed path
    items[0].x["attention_mask"][0, 1] = 0
    assert not HostBatch(items, pin=False).text_prefix


def test_prepend_tokens_equals_cat_with_gradients():
    """The strided-copy prepend of the bottleneck tokens (fusion layer, ViT side) is ``torch.cat`` value- and gradient-wise
    (multi_graphormer_fusion_layer.py:57 of the reference concatenates the same two tensors)."""
    from multimodaldiscussiontransformer_b200.modules.multi_graphormer_fusion_layer import _PrependTokens, prepend_tokens

    g = torch.Generator().manual_seed(3)
    bn = torch.randn(5, 4, 16, generator=g, requires_grad=True)
    full = torch.randn(5, 4 + 9, 16, generator=g, requires_grad=True)
    hidden = full[:, 4:]  # the non-contiguous view a fusion layer hands to the next one
    got, ref = _PrependTokens.apply(bn, hidden), torch.cat([bn, hidden], dim=1)
    assert torch.equal(got, ref) and torch.equal(prepend_tokens(bn, hidden), ref)
    go = torch.randn(ref.shape, generator=g)
    for a, b in zip(torch.autograd.grad(got, (bn, full), go), torch.autograd.grad(ref, (bn, full), go)):
        assert torch.equal(a, b)


def test_check_discussion_tree_refuses_non_trees():
    """the device path re-derives hop pairs from a tree rooted at node 0; everything else must raise."""
    from multimodaldiscussiontransformer_b200.data import synthetic as syn
    from multimodaldiscussiontransformer_b200.data.collator import HostBatch, check_discussion_tree

    items = syn.synthetic_items([6, 4, 1], 3, with_images=False)
    HostBatch(items, pin=False)                                 # trees (and a lone post) pass
    e = np.asarray(items[0].edge_index).copy()                  # [parent->child x5 | child->parent x5]
    check_discussion_tree(e, 6)
    with pytest.raises(ValueError):                             # a cycle: one edge too many
        check_discussion_tree(np.concatenate([e, np.array([[1, 5], [5, 1]])], axis=1), 6)
    with pytest.raises(ValueError):                             # a forest: one reply edge missing
        check_discussion_tree(np.delete(e, [4, 9], axis=1), 6)
    with pytest.raises(ValueError):                             # one direction listed twice, the other never
        check_discussion_tree(np.concatenate([e[:, :5], e[:, :5]], axis=1), 6)
    bad = e.copy()
    bad[0, 0] = 7
    with pytest.raises(ValueError):                             # endpoint outside the graph
        check_discussion_tree(bad, 6)
    # right edge count, symmetric, but a cycle plus a disconnected node
    cyc = np.array([[0, 1, 2, 0, 1, 2, 3, 4], [1, 2, 0, 2, 0, 1, 4, 3]])
    with pytest.raises(ValueError):
        check_discussion_tree(cyc, 5)
    items[1].edge_index = np.delete(np.asarray(items[1].edge_index), [2, 5], axis=1)
    with pytest.raises(ValueError):
        HostBatch(items, pin=False)
