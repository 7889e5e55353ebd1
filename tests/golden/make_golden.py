"""Generate the golden fixtures by running the UNMODIFIED reference in this container.

    python tests/golden/make_golden.py [--only apsp,hops,collate,modules,encoder,encoder_big] [--skip-cfg1]

Needs ``/root/reference`` (absent on the GPU box), so the outputs are committed:
``tests/golden/*.npz`` / ``*.json``.  Inputs are rebuilt from the seeds in ``cases.py``.
See ``ref_harness.py`` for the (minimal) shims the reference needs to import here.
"""

from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))

import cases  # noqa: E402
import ref_harness as rh  # noqa: E402
from mdt_oracle import seed_parameters  # noqa: E402  (name-keyed seeding helper only)


# tests/test_live_reference.py sets this to a dict: fixtures are then kept in memory (name -> arrays / object)
# instead of being written, and compared with the committed files.
CAPTURE = None


def save(name, **arrays):
    if CAPTURE is not None:
        CAPTURE[name] = {k: np.asarray(v) for k, v in arrays.items()}
        return
    path = os.path.join(HERE, name)
    np.savez_compressed(path, **arrays)
    print(f"  wrote {name}  ({os.path.getsize(path) / 1024:.1f} KiB)")


def save_json(name, obj, indent=1):
    if CAPTURE is not None:
        CAPTURE[name] = obj
        return
    json.dump(obj, open(os.path.join(HERE, name), "w"), indent=indent)
    print(f"  wrote {name}")


def gen_apsp():
    algos = rh.ref_algos()
    out = {}
    for name, adj in cases.apsp_cases().items():
        M, path = algos.floyd_warshall(adj)
        out[name + ".M"] = np.asarray(M)
        out[name + ".path"] = np.asarray(path)
    # gen_edge_input on two of them
    rng = np.random.default_rng(99)
    for name in ("tree16", "cyclic20"):
        adj = cases.apsp_cases()[name]
        n = adj.shape[0]
        M, path = algos.floyd_warshall(adj)
        ef = rng.integers(1, 50, size=(n, n, 2)).astype(np.int64)
        max_dist = int(np.asarray(M)[np.asarray(M) < 510].max())
        out[name + ".edge_input"] = np.asarray(algos.gen_edge_input(max_dist, np.asarray(path), ef))
        out[name + ".edge_feat"] = ef
    save("apsp_small.npz", **out)
    dig = {}
    for name, adj in cases.apsp_digest_cases().items():
        t = time.time()
        M, path = algos.floyd_warshall(adj)
        dig[name] = {"M": cases.digest(np.asarray(M)), "path": cases.digest(np.asarray(path)),
                     "n": int(adj.shape[0]), "ref_seconds": round(time.time() - t, 4)}
    save_json("apsp_digests.json", dig)


def gen_hops():
    H = rh.ref_hop_functions()
    out = {}
    rng = np.random.default_rng(77)
    for n in (1, 4, 11, 30):
        parent = cases.syn.random_tree(n, rng)
        nodes = rh.nested_tree_from_parent(parent)
        H.get_relative_depth(nodes[0], {})
        H.spread_downwards(nodes[0], {})
        dm = np.array([[nodes[i]["distances"][j] for j in range(n)] for i in range(n)], dtype=np.int32)
        out[f"n{n}.parent"] = parent
        out[f"n{n}.hops"] = dm.reshape(n, n, 2)
    save("hop_pairs.npz", **out)


def _ref_collate(items, spmax):
    pre = rh.ref_pre_processing()
    col = rh.ref_collator()
    tup = []
    for it in items:
        it = pre.preprocess_item(it)
        tup.append((it.idx, it.attn_bias, it.spatial_pos, it.in_degree, it.x, it.x_image_index, it.x_images,
                    it.distance, it.y))
    batch = col.collator(tup, spmax)
    batch["y_mask"] = torch.cat([it.y_mask for it in items]).bool()          # dataset.py:210-213
    return batch


def gen_collate():
    out = {}
    pre = rh.ref_pre_processing()
    # the Cantor LUT as the reference builds it (pre_processing.py:31-37)
    res = []
    for i in range(6):
        for k in range(6):
            res += [pre.cantor([i, k])]
    res = list(set(res))
    mapping = {val: i for i, val in enumerate(res)}
    out["cantor_lut"] = np.array([[mapping[pre.cantor([i, k])] for k in range(6)] for i in range(6)], dtype=np.int32)
    for name in cases.COLLATE_CASES:
        items, spmax = cases.collate_items(name, list_distance=True)
        # per-item preprocess outputs of the first item (dtype probe)
        it0 = pre.preprocess_item(cases.collate_items(name, list_distance=True)[0][0])
        out[f"{name}.item0.spatial_pos"] = it0.spatial_pos.numpy()
        out[f"{name}.item0.distance"] = it0.distance.numpy()
        out[f"{name}.item0.in_degree"] = it0.in_degree.numpy()
        out[f"{name}.item0.attn_bias"] = it0.attn_bias.numpy()
        batch = _ref_collate(items, spmax)
        for k, v in batch.items():
            if k == "x_images":
                # images are inputs; keep only their count and checksum
                out[f"{name}.x_images.shape"] = np.array(v.shape if v is not None else (0,), dtype=np.int64)
                out[f"{name}.x_images.sum"] = np.array(float(v.double().sum()) if v is not None else 0.0)
            else:
                out[f"{name}.{k}"] = v.numpy()
    save("collate.npz", **out)


def gen_modules():
    mods = rh.ref_modules()
    out = {}
    inp = cases.graph_module_inputs()
    C, H = inp["x"].shape[-1], 4
    # ---- GraphNodeFeature
    gnf = mods.GraphNodeFeature(num_heads=H, num_atoms=16, num_in_degree=16, num_out_degree=16, hidden_dim=C, n_layers=2)
    seed_parameters(gnf, 31)
    x = inp["x"].clone().requires_grad_(True)
    y = gnf(x, inp["in_degree"], inp["in_degree"])
    g = torch.Generator().manual_seed(1)
    gy = torch.randn(y.shape, generator=g)
    y.backward(gy)
    out["gnf.out"] = y.detach().numpy()
    out["gnf.dx"] = x.grad.numpy()
    out["gnf.d_in"] = gnf.in_degree_encoder.weight.grad.numpy()
    out["gnf.d_out"] = gnf.out_degree_encoder.weight.grad.numpy()
    out["gnf.d_tok"] = gnf.graph_token.weight.grad.numpy()
    # ---- GraphAttnBias
    gab = mods.GraphAttnBias(num_heads=H, num_atoms=16, num_edges=16, num_spatial=32, num_edge_dis=8, hidden_dim=C,
                             edge_type="multi_hop", multi_hop_max_dist=5, n_layers=2)
    seed_parameters(gab, 32)
    bias = gab(dict(attn_bias=inp["attn_bias"], spatial_pos=inp["spatial_pos"], x=None))
    gb = torch.randn(bias.shape, generator=g)
    finite = torch.isfinite(bias)
    (bias.masked_fill(~finite, 0.0) * gb).sum().backward()
    out["gab.out"] = bias.detach().numpy()
    out["gab.d_spatial"] = gab.spatial_pos_encoder.weight.grad.numpy()
    out["gab.d_virtual"] = gab.graph_token_virtual_distance.weight.grad.numpy()
    # ---- MultiheadAttention
    mha = mods.MultiheadAttention(C, H, dropout=0.0, self_attention=True)
    seed_parameters(mha, 33, std=0.08)
    mha.eval()
    q = inp["query"].clone().requires_grad_(True)
    bias_in = bias.detach().clone().requires_grad_(True)
    o, w = mha(q, q, q, bias_in, key_padding_mask=inp["pad_mask"], need_weights=False)
    assert w is None
    o.backward(inp["grad_out"])
    out["mha.out"] = o.detach().numpy()
    out["mha.dq"] = q.grad.numpy()
    out["mha.dbias"] = bias_in.grad.numpy()
    for n_, p_ in mha.named_parameters():
        out[f"mha.grad.{n_}"] = p_.grad.numpy()
    o2, _ = mha(inp["query"], None, None, None, key_padding_mask=None, need_weights=False)   # no bias, no mask
    out["mha.out_nobias"] = o2.detach().numpy()
    # ---- encoder layer (post-LN and pre-LN) and a 2-layer stack
    enc_mod = sys.modules["_mdt_ref_modules.graphormer_graph_encoder_layer"]
    for tag, pre_ln in (("postln", False), ("preln", True)):
        layer = mods.GraphormerGraphEncoderLayer(embedding_dim=C, ffn_embedding_dim=96, num_attention_heads=H,
                                                 dropout=0.0, attention_dropout=0.0, activation_dropout=0.0,
                                                 activation_fn="gelu", pre_layernorm=pre_ln)
        seed_parameters(layer, 34, std=0.08)
        layer.eval()
        q = inp["query"].clone().requires_grad_(True)
        o, _ = layer(q, self_attn_bias=bias.detach(), self_attn_padding_mask=inp["pad_mask"])
        o.backward(inp["grad_out"])
        out[f"layer_{tag}.out"] = o.detach().numpy()
        out[f"layer_{tag}.dx"] = q.grad.numpy()
        out[f"layer_{tag}.d_fc1"] = layer.fc1.weight.grad.numpy()
    stack = enc_mod.GraphEncoderStack(2, embedding_dim=C, ffn_embedding_dim=96, num_attention_heads=H, dropout=0.0,
                                      attention_dropout=0.0, activation_dropout=0.0, activation_fn="relu")
    seed_parameters(stack, 35, std=0.08)
    stack.eval()
    o, _ = stack(inp["query"], bias.detach(), None, inp["pad_mask"])
    out["stack_relu.out"] = o.detach().numpy()
    save("graph_modules.npz", **out)


def _encoder_case(name, c, out, meta):
    t0 = time.time()
    items, spmax = cases.encoder_items(name, list_distance=True)
    batch = _ref_collate(items, spmax)
    m = dict(c["model"])
    bl, vl = m.pop("bert_layers"), m.pop("vit_layers")
    with rh.reference_runtime(bert_layers=bl, vit_layers=vl) as mods:
        kw = dict(cases.COMMON_MODEL)
        kw.update(m)
        enc = mods.MultiGraphormerGraphEncoder(multi_hop_max_dist=5, dropout=0.0, attention_dropout=0.0,
                                               activation_dropout=0.0, **kw)
        seed_parameters(enc, 41)
        enc.eval()
        t1 = time.time()
        out_all, glob, loss = rh.reference_head_and_loss(enc, batch)
        t2 = time.time()
        loss.backward()
        t3 = time.time()
    out[f"{name}.out_all"] = out_all.detach().numpy()
    out[f"{name}.global"] = glob.detach().numpy()
    out[f"{name}.loss"] = np.array(float(loss))
    norms = {}
    for n_, p_ in enc.named_parameters():
        norms[n_] = None if p_.grad is None else float(p_.grad.double().norm())
    out[f"{name}.grad.spatial_pos_encoder"] = enc.graph_attn_bias.spatial_pos_encoder.weight.grad[:24].numpy()
    out[f"{name}.grad.virtual_distance"] = enc.graph_attn_bias.graph_token_virtual_distance.weight.grad.numpy()
    out[f"{name}.grad.bottle_neck"] = enc.bottle_neck.weight.grad.numpy()
    out[f"{name}.grad.in_degree_encoder"] = enc.graph_node_feature.in_degree_encoder.weight.grad[:8].numpy()
    out[f"{name}.grad.q_proj0"] = enc.layers[0].layers[0].self_attn.q_proj.weight.grad[:4].numpy()
    # criterion counters of hatespeech_loss.py:102-110 on the reference logits (for the product-criterion GPU test)
    logits = out_all.detach()[batch["y_mask"]].float()
    targets = torch.flatten(batch["y"]).long()
    pred = logits.argmax(-1)
    out[f"{name}.counters"] = np.array([logits.shape[0], int((pred == targets).sum()), int(((pred == targets) & (pred == 1)).sum()),
                                        int((targets == 1).sum()), int((pred == 1).sum())], dtype=np.int64)
    f0 = enc.fusion_layers[0].fusion_layers[0].bert_encoder
    out[f"{name}.grad.fusion0_query"] = f0.attention.self.query.weight.grad[:4].numpy()
    out[f"{name}.grad.fusion0_inter"] = f0.intermediate.dense.weight.grad[:4].numpy()
    out[f"{name}.grad.classifier"] = enc.node_classifier.weight.grad.numpy()
    meta[name] = dict(grad_norms=norms, n_params=sum(p.numel() for p in enc.parameters()),
                      ref_fwd_s=round(t2 - t1, 3), ref_bwd_s=round(t3 - t2, 3), build_s=round(t1 - t0, 3),
                      state_keys=len(enc.state_dict()))
    print(f"  encoder case {name}: loss={float(loss):.6f} fwd={t2 - t1:.2f}s bwd={t3 - t2:.2f}s")


def gen_encoder_big():
    meta, out = {}, {}
    for name, c in cases.ENCODER_BIG_CASES.items():
        _encoder_case(name, c, out, meta)
    save("encoder_big.npz", **out)
    save_json("encoder_big_meta.json", meta)


def gen_encoder(skip_cfg1=False, only=None):
    meta = {}
    out = {}
    for name, c in cases.ENCODER_CASES.items():
        if (name == "cfg1" and skip_cfg1) or (only is not None and name not in only):
            continue
        _encoder_case(name, c, out, meta)
    save("encoder.npz", **out)
    save_json("encoder_meta.json", meta)
    # reference state-dict key list (drop-in checkpoint compatibility test)
    with rh.reference_runtime(bert_layers=3, vit_layers=3) as mods:
        kw = dict(cases.COMMON_MODEL)
        kw.update(num_fusion_layers=1, num_fusion_stack=1, num_graph_stack=1)
        enc = mods.MultiGraphormerGraphEncoder(multi_hop_max_dist=5, **kw)
        keys = {k: list(v.shape) for k, v in enc.state_dict().items()}
    save_json("encoder_state_keys.json", keys, indent=0)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="apsp,hops,collate,modules,encoder,encoder_big")
    ap.add_argument("--skip-cfg1", action="store_true")
    a = ap.parse_args()
    assert rh.reference_available(), "needs /root/reference"
    torch.manual_seed(0)
    torch.set_num_threads(os.cpu_count())
    for part in a.only.split(","):
        print(f"[{part}]")
        {"apsp": gen_apsp, "hops": gen_hops, "collate": gen_collate, "modules": gen_modules,
         "encoder": lambda: gen_encoder(a.skip_cfg1), "encoder_big": gen_encoder_big}[part]()
