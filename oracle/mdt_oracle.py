"""ORACLE -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

CPU restatement of mDT's graph-transformer hot path.
numpy for the integer work, torch fp32 for the floating-point modules; no CUDA.

Who may import this: ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs -- as the checker or the timed CPU
baseline, never as the product.  ``multimodaldiscussiontransformer_b200`` does not
import it and fails loudly without its CUDA library.

Parity pin (the reference has no tests or golden vectors):
every function here is checked against fixtures under ``tests/golden/`` that
``tests/golden/make_golden.py`` produced by running the *unmodified* reference files
(``/root/reference/mDT/src/modules/*.py``, ``data/collator.py``,
``data/pyg_datasets/pre_processing.py``, compiled ``data/algos.pyx``) in the build
container, and -- when ``/root/reference`` is present -- against the live reference
(``tests/test_live_reference.py``).

Third-party arithmetic restated by name only (not under /root/reference):
  * fairseq (version unpinned by the reference): ``utils.softmax`` = softmax in fp32,
    ``get_activation_fn('gelu')`` = ``F.gelu(x.float()).type_as(x)``, ``LayerNorm`` =
    ``torch.nn.LayerNorm(eps=1e-5)``, ``FairseqDropout`` = ``F.dropout``.
  * transformers 5.5.0 ``BertLayer`` / ``ViTLayer`` / ``BertModel`` / ``ViTModel`` are
    called as libraries (their layer math is third-party for the reference too).
"""

from __future__ import annotations

import ctypes
import math
import os
from typing import Any, Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

UNREACHABLE = 510  # algos.pyx:24-30
_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


# ======================================================================================
# A1 / A1b -- Floyd-Warshall + edge-input expansion (C restatement, oracle/apsp_oracle.c)
# ======================================================================================
def _lib():
    global _LIB
    if _LIB is None:
        path = os.path.join(_HERE, "_build", "libmdt_oracle.so")
        if not os.path.isfile(path):
            import subprocess

            subprocess.check_call(["make", "-C", _HERE, "-s"])
        lib = ctypes.CDLL(path)
        p64 = ctypes.POINTER(ctypes.c_int64)
        lib.mdt_oracle_floyd_warshall.argtypes = [p64, p64, p64, ctypes.c_int64]
        lib.mdt_oracle_floyd_warshall.restype = None
        lib.mdt_oracle_floyd_warshall_batch.argtypes = [p64, p64, p64, ctypes.c_int64, ctypes.c_int64]
        lib.mdt_oracle_floyd_warshall_batch.restype = None
        lib.mdt_oracle_gen_edge_input.argtypes = [ctypes.c_int64, p64, p64, p64, ctypes.c_int64, ctypes.c_int64]
        lib.mdt_oracle_gen_edge_input.restype = ctypes.c_int
        _LIB = lib
    return _LIB


def _p64(a: np.ndarray):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))


def floyd_warshall(adjacency_matrix: np.ndarray):
    """``algos.floyd_warshall`` (algos.pyx:7-52) -> ``(M, path)`` int64 ``[n, n]``."""
    nrows, ncols = adjacency_matrix.shape
    assert nrows == ncols
    adj = np.ascontiguousarray(adjacency_matrix.astype(np.int64, order="C", casting="safe", copy=True))
    M = np.empty_like(adj)
    path = np.empty_like(adj)
    _lib().mdt_oracle_floyd_warshall(_p64(adj), _p64(M), _p64(path), nrows)
    return M, path


def floyd_warshall_batch(adj: np.ndarray):
    """``[B, n, n]`` int64 in -> ``(M, path)``; the loop the CPU-baseline timer measures."""
    B, n, _ = adj.shape
    adj = np.ascontiguousarray(adj.astype(np.int64))
    M = np.empty_like(adj)
    path = np.empty_like(adj)
    _lib().mdt_oracle_floyd_warshall_batch(_p64(adj), _p64(M), _p64(path), n, B)
    return M, path


def gen_edge_input(max_dist: int, path: np.ndarray, edge_feat: np.ndarray) -> np.ndarray:
    """``algos.gen_edge_input`` (algos.pyx:63-89) -> int64 ``[n, n, max_dist, F]``."""
    n = path.shape[0]
    assert path.shape[1] == n
    p = np.ascontiguousarray(path.astype(np.int64))
    ef = np.ascontiguousarray(edge_feat.astype(np.int64))
    Fdim = ef.shape[-1]
    out = np.empty((n, n, max_dist, Fdim), dtype=np.int64)
    rc = _lib().mdt_oracle_gen_edge_input(max_dist, _p64(p), _p64(ef), _p64(out), n, Fdim)
    if rc != 0:
        raise IndexError("shortest path longer than max_dist")
    return out


# ======================================================================================
# A2 -- preprocess_item (pre_processing.py:18-69)
# ======================================================================================
HOP_MAX = 5          # range(6) in pre_processing.py:32-33
NUM_SPATIAL_LIVE = 21


def cantor_pair(a: int, b: int) -> int:
    """pre_processing.py:18-20 (sorted pair)."""
    lo, hi = (a, b) if a <= b else (b, a)
    return (lo + hi) * (lo + hi + 1) // 2 + lo


def cantor_lut() -> np.ndarray:
    """``[6, 6]`` table: rank of cantor(i, k) among the 21 distinct in-range values.

    pre_processing.py:31-37 builds ``mapping`` by enumerating ``list(set(res))``; the
    values are small non-negative integer-valued floats, which CPython hashes to
    themselves, so the set iterates in ascending order (checked against the reference
    in tests/golden).  Out-of-range pairs map to ``mapping[cantor(5, 5)] == 20``
    (pre_processing.py:44-47).
    """
    vals = sorted({cantor_pair(i, k) for i in range(HOP_MAX + 1) for k in range(HOP_MAX + 1)})
    rank = {v: r for r, v in enumerate(vals)}
    lut = np.zeros((HOP_MAX + 1, HOP_MAX + 1), dtype=np.int32)
    for i in range(HOP_MAX + 1):
        for k in range(HOP_MAX + 1):
            lut[i, k] = rank[cantor_pair(i, k)]
    return lut


def spatial_from_hops(hops: np.ndarray) -> np.ndarray:
    """``[n, n, 2]`` hop pairs -> spatial index 0..20 (pre_processing.py:39-53)."""
    lut = cantor_lut()
    up = hops[..., 0].astype(np.int64)
    down = hops[..., 1].astype(np.int64)
    inside = (up >= 0) & (down >= 0) & (up <= HOP_MAX) & (down <= HOP_MAX)
    idx = np.where(inside, lut[np.clip(up, 0, HOP_MAX), np.clip(down, 0, HOP_MAX)], lut[HOP_MAX, HOP_MAX])
    return idx.astype(np.int64)


def preprocess_item(item) -> Dict[str, np.ndarray]:
    """``preprocess_item`` (pre_processing.py:23-69) as plain arrays.

    Returns ``spatial_pos`` f32 ``[N,N]``, ``distance`` f32 ``[N,N]``, ``in_degree`` /
    ``out_degree`` i64 ``[N]``, ``attn_bias`` f32 zeros ``[N+1,N+1]``.
    """
    ids = item.x["input_ids"]
    N = int(ids.shape[0])
    ei = np.asarray(item.edge_index)
    adj = np.zeros((N, N), dtype=bool)                       # :28-29
    adj[ei[0], ei[1]] = True
    hops = np.asarray(item.distance_matrix, dtype=np.int64).reshape(N, N, 2)
    spatial = spatial_from_hops(hops).astype(np.float32)      # :39-53,64
    distance = hops.sum(-1).astype(np.float32)                # :54-56,65
    deg = adj.astype(np.int64).sum(axis=1)                    # :66
    return dict(
        spatial_pos=spatial,
        distance=distance,
        in_degree=deg,
        out_degree=deg,                                       # :67
        attn_bias=np.zeros((N + 1, N + 1), dtype=np.float32), # :57-59
    )


# ======================================================================================
# A3 -- collator (collator.py:69-179) + y_mask (dataset.py:210-213)
# ======================================================================================
def collator(items: Sequence[Any], spatial_pos_max: int = 10) -> Dict[str, Any]:
    """Collate preprocessed items; numpy restatement of collator.py:69-179.

    ``items`` are objects with the item fields (``idx, x, x_image_index, x_images, y,
    y_mask``) plus the arrays ``preprocess_item`` returns (as attributes or via
    ``item.extra``).  Output dtypes/shapes follow SURVEY.md section 8a row A3.
    """
    pre = [it.extra["pre"] if "pre" in getattr(it, "extra", {}) else preprocess_item(it) for it in items]
    B = len(items)
    ns = [int(np.asarray(it.x["input_ids"]).shape[0]) for it in items]
    N = max(ns)                                                # :127
    T = N + 1
    L = int(np.asarray(items[0].x["input_ids"]).shape[1])
    attn_bias = np.full((B, T, T), -np.inf, dtype=np.float32)
    spatial_pos = np.zeros((B, N, N), dtype=np.int32)
    in_degree = np.zeros((B, N), dtype=np.int64)
    xs = {k: np.zeros((B, N, L), dtype=np.int64) for k in ("input_ids", "token_type_ids", "attention_mask")}
    for b, (it, p, n) in enumerate(zip(items, pre, ns)):
        ab = p["attn_bias"].copy()
        ab[1:, 1:][p["distance"] >= spatial_pos_max] = -np.inf    # :122-126
        if n < N:                                                 # pad_attn_bias_unsqueeze :38-47
            attn_bias[b, : n + 1, : n + 1] = ab
            attn_bias[b, n + 1 :, : n + 1] = 0
        else:
            attn_bias[b] = ab
        spatial_pos[b, :n, :n] = (p["spatial_pos"] + 1).astype(np.int32)   # :59-66,159-161
        in_degree[b, :n] = p["in_degree"] + 1                              # :8-20,162-164
        for k in xs:
            xs[k][b, :n] = np.asarray(it.x[k])                             # :131-139
    token_mask = ~(xs["input_ids"] == 0).all(axis=2)                       # :141
    imgs = [np.asarray(it.x_images) for it in items]
    imgs = [im for im in imgs if not np.all(im == 0)]                      # :144
    x_images = np.concatenate(imgs) if imgs else None                      # :145-148
    x_image_indexes = np.concatenate([np.asarray(it.x_image_index) for it in items]).astype(bool)  # :149-154
    y = np.concatenate([np.asarray(it.y) for it in items]).astype(np.float32)  # :129
    out = dict(
        idx=np.asarray([it.idx for it in items], dtype=np.int64),
        attn_bias=attn_bias,
        spatial_pos=spatial_pos,
        in_degree=in_degree,
        out_degree=in_degree,                                              # :171
        x_token_mask=token_mask,
        x=xs["input_ids"],
        x_token_type_ids=xs["token_type_ids"],
        x_attention_mask=xs["attention_mask"],
        x_images=x_images,
        x_image_indexes=x_image_indexes,
        y=y,
    )
    if all(hasattr(it, "y_mask") and it.y_mask is not None for it in items):
        out["y_mask"] = np.concatenate([np.asarray(it.y_mask) for it in items]).astype(bool)  # dataset.py:210-213
    return out


def batch_to_torch(batch: Dict[str, Any]) -> Dict[str, Any]:
    return {k: (torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) else v) for k, v in batch.items()}


# ======================================================================================
# A4 / A5 -- GraphNodeFeature, GraphAttnBias (graphormer_layers.py:39-50, 86-110)
# ======================================================================================
def graph_node_feature(x, in_degree, out_degree, w_in, w_out, graph_token):
    """``[B,N,C]`` -> ``[B,N+1,C]``: x + E_in[in] + E_out[out], graph token prepended."""
    node = x + w_in[in_degree] + w_out[out_degree]                        # :42-44
    tok = graph_token.unsqueeze(0).expand(x.shape[0], -1, -1)             # :46
    return torch.cat([tok.to(node.dtype), node], dim=1)                   # :48


def graph_attn_bias(attn_bias, spatial_pos, w_spatial, w_virtual):
    """``[B,T,T]``, ``[B,N,N]`` -> ``[B,H,T,T]``.

    out = 2*attn_bias (clone :93 + 'reset' :108) + E_sp[spatial_pos] on [1:,1:] (:100-101)
    + t on (i>=1, j=0) (:105) and on (i=0, all j) (:106).  Row 0 of E_sp is gathered like
    any other row (init_params overwrites the padding row, :12-13).
    """
    H = w_spatial.shape[1]
    g = attn_bias.unsqueeze(1).repeat(1, H, 1, 1)
    sp = w_spatial[spatial_pos.long()].permute(0, 3, 1, 2)
    g[:, :, 1:, 1:] = g[:, :, 1:, 1:] + sp
    t = w_virtual.view(1, H, 1)
    g[:, :, 1:, 0] = g[:, :, 1:, 0] + t
    g[:, :, 0, :] = g[:, :, 0, :] + t
    return g + attn_bias.unsqueeze(1)


# ======================================================================================
# A6 -- MultiheadAttention (multihead_attention.py:91-214)
# ======================================================================================
def multihead_attention(query, p: Dict[str, torch.Tensor], prefix: str, num_heads: int,
                        attn_bias=None, key_padding_mask=None, attn_mask=None,
                        dropout_p: float = 0.0, training: bool = False):
    """``query [T,B,C]`` -> ``[T,B,C]``.  key/value arguments do not exist: the reference
    projects ``query`` three times (:134-136)."""
    T, B, C = query.shape
    d = C // num_heads
    q = F.linear(query, p[prefix + "q_proj.weight"], p.get(prefix + "q_proj.bias"))
    k = F.linear(query, p[prefix + "k_proj.weight"], p.get(prefix + "k_proj.bias"))
    v = F.linear(query, p[prefix + "v_proj.weight"], p.get(prefix + "v_proj.bias"))
    q = q * (d ** -0.5)                                                    # :137
    q = q.contiguous().view(T, B * num_heads, d).transpose(0, 1)
    k = k.contiguous().view(T, B * num_heads, d).transpose(0, 1)
    v = v.contiguous().view(T, B * num_heads, d).transpose(0, 1)
    s = torch.bmm(q, k.transpose(1, 2))                                    # :168
    if attn_bias is not None:
        s = s + attn_bias.reshape(B * num_heads, T, T)                     # :173-174
    if attn_mask is not None:
        s = s + attn_mask.unsqueeze(0)                                     # :176-178
    if key_padding_mask is not None:
        s = s.view(B, num_heads, T, T).masked_fill(
            key_padding_mask.unsqueeze(1).unsqueeze(2).to(torch.bool), float("-inf")
        ).view(B * num_heads, T, T)                                        # :180-187
    pr = F.softmax(s, dim=-1, dtype=torch.float32).type_as(s)              # :192-195
    pr = F.dropout(pr, p=dropout_p, training=training)                     # :196
    o = torch.bmm(pr, v)                                                   # :199
    o = o.transpose(0, 1).contiguous().view(T, B, C)                       # :202
    return F.linear(o, p[prefix + "out_proj.weight"], p.get(prefix + "out_proj.bias"))  # :203


# ======================================================================================
# A7 -- GraphormerGraphEncoderLayer / GraphEncoderStack
# ======================================================================================
def _act(name: str):
    if name == "gelu":
        return lambda x: F.gelu(x.float()).type_as(x)
    if name == "relu":
        return F.relu
    raise ValueError(name)


def graph_encoder_layer(x, p, prefix, num_heads, attn_bias, padding_mask, attn_mask=None,
                        pre_layernorm=False, activation_fn="gelu",
                        dropout=0.0, attention_dropout=0.0, activation_dropout=0.0, training=False):
    """graphormer_graph_encoder_layer.py:103-142 (post-LN default, pre-LN optional)."""
    C = x.shape[-1]

    def ln(t, name):
        return F.layer_norm(t, (C,), p[prefix + name + ".weight"], p[prefix + name + ".bias"], 1e-5)

    res = x
    if pre_layernorm:
        x = ln(x, "self_attn_layer_norm")
    x = multihead_attention(x, p, prefix + "self_attn.", num_heads, attn_bias, padding_mask, attn_mask,
                            attention_dropout, training)
    x = F.dropout(x, dropout, training)
    x = res + x
    if not pre_layernorm:
        x = ln(x, "self_attn_layer_norm")
    res = x
    if pre_layernorm:
        x = ln(x, "final_layer_norm")
    x = _act(activation_fn)(F.linear(x, p[prefix + "fc1.weight"], p[prefix + "fc1.bias"]))
    x = F.dropout(x, activation_dropout, training)
    x = F.linear(x, p[prefix + "fc2.weight"], p[prefix + "fc2.bias"])
    x = F.dropout(x, dropout, training)
    x = res + x
    if not pre_layernorm:
        x = ln(x, "final_layer_norm")
    return x


# ======================================================================================
# A8 / A9 / A10 / A11 -- encoder assembly, fusion, head, loss
# ======================================================================================
class OracleMDT(nn.Module):
    """fp32 CPU restatement of ``MultiGraphormerGraphEncoder`` (+ the model head).

    Parameter *names* equal the reference's so a state dict moves
    between the reference, this oracle and the CUDA product unchanged.  The modules
    registered here are containers only; the arithmetic is the functional code above
    plus HF ``BertLayer`` / ``ViTLayer`` called as libraries.
    """

    def __init__(self, num_in_degree=512, num_out_degree=512, num_edges=512 * 3, num_spatial=512,
                 num_edge_dis=128, num_atoms=512 * 9, num_bottle_neck=4, num_fusion_layers=1,
                 num_fusion_stack=1, num_graph_stack=1, embedding_dim=768, ffn_embedding_dim=768,
                 num_attention_heads=12, encoder_normalize_before=True, pre_layernorm=False,
                 activation_fn="gelu", edge_type="multi_hop", bert_layers=12, vit_layers=12,
                 dropout=0.0, attention_dropout=0.0, activation_dropout=0.0):
        super().__init__()
        from transformers import BertConfig, BertForSequenceClassification, ViTConfig, ViTModel

        C, H = embedding_dim, num_attention_heads
        self.cfg = dict(C=C, H=H, nb=num_bottle_neck, pre_ln=pre_layernorm, act=activation_fn,
                        dropout=dropout, attention_dropout=attention_dropout, activation_dropout=activation_dropout)

        gnf = nn.Module()
        gnf.atom_encoder = nn.Embedding(num_atoms + 1, C, padding_idx=0)
        gnf.in_degree_encoder = nn.Embedding(num_in_degree, C, padding_idx=0)
        gnf.out_degree_encoder = nn.Embedding(num_out_degree, C, padding_idx=0)
        gnf.graph_token = nn.Embedding(1, C)
        self.graph_node_feature = gnf
        gab = nn.Module()
        gab.edge_encoder = nn.Embedding(num_edges + 1, H, padding_idx=0)
        if edge_type == "multi_hop":
            gab.edge_dis_encoder = nn.Embedding(num_edge_dis * H * H, 1)
        gab.spatial_pos_encoder = nn.Embedding(num_spatial, H, padding_idx=0)
        gab.graph_token_virtual_distance = nn.Embedding(1, H)
        self.graph_attn_bias = gab
        self.emb_layer_norm = nn.LayerNorm(C) if encoder_normalize_before else None
        if pre_layernorm:
            self.final_layer_norm = nn.LayerNorm(C)

        bc = BertConfig(num_labels=2, num_hidden_layers=bert_layers, hidden_dropout_prob=activation_dropout,
                        attention_probs_dropout_prob=attention_dropout)
        vc = ViTConfig(num_hidden_layers=vit_layers, hidden_dropout_prob=activation_dropout,
                       attention_probs_dropout_prob=attention_dropout)
        bc._attn_implementation = "eager"
        vc._attn_implementation = "eager"
        vit = ViTModel(vc)
        bert = BertForSequenceClassification(bc)
        peel = num_fusion_layers + 1                                   # encoder.py:141, 249-260
        vit_other = list(vit.encoder.layer[-peel:])
        vit.encoder.layer = vit.encoder.layer[:-peel]
        bert_other = list(bert.bert.encoder.layer[-peel:])
        bert.bert.encoder.layer = bert.bert.encoder.layer[:-peel]
        self.vit_model = vit
        self.vit_pooler = vit.pooler
        self.text_model = bert.bert
        self.text_pooler = bert.bert.pooler
        self.node_classifier = bert.classifier
        self.text_dropout = bert.dropout

        def chunks(xs, k):
            return [xs[i * k:(i + 1) * k] for i in range((len(xs) + k - 1) // k)]

        self.fusion_layers = nn.ModuleList()
        for bl, vl in zip(chunks(bert_other, num_fusion_stack), chunks(vit_other, num_fusion_stack)):
            stack = nn.Module()
            stack.fusion_layers = nn.ModuleList()
            for b_layer, v_layer in zip(bl, vl):
                fl = nn.Module()
                fl.bert_encoder = b_layer
                fl.vit_encoder = v_layer
                fl.bert_projection = nn.Linear(768, 768)               # use_projection=True, dead
                fl.vit_projection = nn.Linear(768, 768)
                stack.fusion_layers.append(fl)
            self.fusion_layers.append(stack)

        self.layers = nn.ModuleList()
        for _ in range(len(self.fusion_layers) + 1):                   # encoder.py:173-191 (one stack is never run)
            st = nn.Module()
            st.layers = nn.ModuleList()
            for _ in range(num_graph_stack):
                ly = nn.Module()
                sa = nn.Module()
                for nm in ("k_proj", "v_proj", "q_proj", "out_proj"):
                    setattr(sa, nm, nn.Linear(C, C))
                ly.self_attn = sa
                ly.self_attn_layer_norm = nn.LayerNorm(C)
                ly.fc1 = nn.Linear(C, ffn_embedding_dim)
                ly.fc2 = nn.Linear(ffn_embedding_dim, C)
                ly.final_layer_norm = nn.LayerNorm(C)
                st.layers.append(ly)
            self.layers.append(st)
        self.bottle_neck = nn.Embedding(num_bottle_neck, 768)

    # ---------------------------------------------------------------------------------
    def _graph_stack(self, x, idx, p, attn_bias, padding_mask):
        c = self.cfg
        for j in range(len(self.layers[idx].layers)):
            x = graph_encoder_layer(
                x, p, f"layers.{idx}.layers.{j}.", c["H"], attn_bias, padding_mask, None, c["pre_ln"], c["act"],
                c["dropout"], c["attention_dropout"], c["activation_dropout"], self.training)
        return x

    def _fusion_stack(self, idx, bert_h, vit_h, bn, mask, img_idx):
        """multi_graphormer_fusion_layer.py:29-71 for every layer of fusion stack ``idx``."""
        nb = self.cfg["nb"]
        for fl in self.fusion_layers[idx].fusion_layers:
            b_out = fl.bert_encoder(torch.cat([bn, bert_h], dim=1), mask)
            bert_h, bn_out = b_out[:, nb:], b_out[:, :nb]
            if vit_h is not None:
                v_out = fl.vit_encoder(torch.cat([bn[img_idx], vit_h], dim=1))   # ViT sees the *input* bottleneck
                vit_h = v_out[:, nb:]
                avg = (v_out[:, :nb] + bn_out[img_idx]) / 2
                bn_out = bn_out.index_put((torch.nonzero(img_idx).squeeze(1),), avg)
            bn = bn_out
        return bert_h, vit_h, bn

    def forward(self, batched_data):
        """encoder.py:310-464 -> (bert_out [n_real,100,768], bottle_neck [n_real,nb,768], global [B,768])."""
        c = self.cfg
        p = dict(self.named_parameters())
        mask = batched_data["x_token_mask"]
        node_idx = torch.nonzero(mask.reshape(-1)).squeeze(1)
        ids = batched_data["x"][mask]
        bert_h = self.text_model(input_ids=ids, token_type_ids=batched_data["x_token_type_ids"][mask],
                                 attention_mask=batched_data["x_attention_mask"][mask]).last_hidden_state
        n_real = bert_h.shape[0]
        vit_h = None
        if batched_data["x_images"] is not None:
            vit_h = self.vit_model(batched_data["x_images"]).last_hidden_state
        bn = self.bottle_neck.weight.unsqueeze(0).repeat(n_real, 1, 1)               # :339
        am = torch.cat([torch.ones(n_real, c["nb"]), batched_data["x_attention_mask"][mask].float()], dim=1)
        ext = ((1.0 - am[:, None, None, :].half()) * torch.finfo(torch.half).min)     # :341-354 (fp16-typed mask)
        ext = ext.float()
        img_idx = batched_data["x_image_indexes"]
        bert_h, vit_h, bn = self._fusion_stack(0, bert_h, vit_h, bn, ext, img_idx)    # :356-362
        B, N = batched_data["x"].shape[:2]
        graph = torch.zeros(B * N, 768, dtype=bn.dtype).index_put((node_idx,), bn[:, 0, :]).view(B, N, 768)  # :363-371
        padding_mask = torch.cat([torch.zeros(B, 1, dtype=torch.bool), graph[:, :, 0].eq(0)], dim=1)         # :376-382
        x = graph_node_feature(graph, batched_data["in_degree"], batched_data["out_degree"],
                               p["graph_node_feature.in_degree_encoder.weight"],
                               p["graph_node_feature.out_degree_encoder.weight"],
                               p["graph_node_feature.graph_token.weight"])              # :386-388
        bias = graph_attn_bias(batched_data["attn_bias"], batched_data["spatial_pos"],
                               p["graph_attn_bias.spatial_pos_encoder.weight"],
                               p["graph_attn_bias.graph_token_virtual_distance.weight"])  # :392
        if self.emb_layer_norm is not None:
            x = self.emb_layer_norm(x)                                                 # :400-401
        x = F.dropout(x, c["dropout"], self.training)                                  # :403
        x = x.transpose(0, 1)                                                          # :407  [T,B,C]
        n_f = len(self.fusion_layers)
        tok_idx = (node_idx // N) * (N + 1) + (node_idx % N) + 1                        # mask with the CLS column, :383
        for g in range(n_f - 1):                                                       # zip(layers, fusion_layers[1:]) :413
            x = self._graph_stack(x, g, p, bias, padding_mask)
            xb = x.transpose(0, 1).reshape(B * (N + 1), 768)
            bn = torch.cat([xb[tok_idx].unsqueeze(1), bn[:, 1:]], dim=1)               # :425
            bert_h, vit_h, bn = self._fusion_stack(g + 1, bert_h, vit_h, bn, ext, img_idx)  # :427-433
            xb = xb.index_put((tok_idx,), bn[:, 0, :])                                 # :435
            x = xb.view(B, N + 1, 768).transpose(0, 1)
        x = self._graph_stack(x, len(self.layers) - 1, p, bias, padding_mask)          # layers[-1], :441-446
        return bert_h, bn, x[0]                                                        # :463-464

    def head(self, text, bn):
        """multi_modal_discussion_transformer.py:256-276: pooler -> dropout -> classifier on both, averaged."""
        for layer in (self.text_pooler, self.text_dropout, self.node_classifier):
            text = layer(text)
            bn = layer(bn)
        return (text + bn) / 2


def node_cross_entropy(out_all, y, y_mask, positive_weight=1.5, negative_weight=1.0):
    """hatespeech_loss.py:95-118: weighted CE, reduction='sum', over labelled comments."""
    logits = out_all[y_mask].float()
    targets = torch.flatten(y).long()
    w = torch.tensor([negative_weight, positive_weight], dtype=torch.float32, device=logits.device)
    loss = F.cross_entropy(logits, targets, reduction="sum", weight=w)
    return loss, logits.shape[0]


def seed_parameters(module: nn.Module, seed: int, std: float = 0.02) -> None:
    """Deterministic, name-keyed weights shared by reference, oracle and product.

    Every floating tensor of ``state_dict()`` (sorted by name, aliases filled once) is
    drawn from a per-name generator: N(0, std) -- LayerNorm weights get 1 + N(0, std).
    Padding rows of embeddings are therefore non-zero (the parity trap of SURVEY.md
    section 8a row A5).
    """
    import zlib

    seen = set()
    with torch.no_grad():
        for name, t in sorted(module.state_dict().items()):
            if not t.is_floating_point() or t.data_ptr() in seen:
                continue
            seen.add(t.data_ptr())
            g = torch.Generator().manual_seed((seed * 1000003 + zlib.crc32(name.encode())) % (2 ** 31))
            v = torch.randn(t.shape, generator=g, dtype=torch.float32) * std
            lname = name.lower()
            if ("layernorm" in lname or "layer_norm" in lname) and name.endswith("weight"):
                v = v + 1.0
            t.copy_(v.to(t.dtype))
