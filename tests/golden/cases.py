"""Seeded test cases shared by ``make_golden.py`` (reference side) and the tests
(oracle / CUDA side).  Inputs are never stored in the fixtures -- both sides rebuild
them from these seeds -- only the reference's outputs are.
"""

from __future__ import annotations

import os
import sys

import numpy as np
import torch

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from multimodaldiscussiontransformer_b200.data import synthetic as syn  # noqa: E402

GOLDEN_DIR = os.path.dirname(os.path.abspath(__file__))


# ------------------------------------------------------------------ APSP (config 2 shaped)
def apsp_cases():
    """name -> adjacency (any integer dtype).  Small ones are stored in full, large ones by digest."""
    rng = np.random.default_rng(1234)
    cases = {}
    for n in (1, 2, 5, 16, 33, 64):
        cases[f"tree{n}"] = syn.tree_adjacency(syn.random_tree(n, rng))
    cases["chain40"] = syn.tree_adjacency(syn.chain_tree(40))
    cases["forest24"] = syn.forest_adjacency(24, 3, rng)
    w = rng.integers(0, 4, size=(12, 12)).astype(np.int64)          # weighted, directed, zeros = no edge
    cases["weighted12"] = w
    g = (rng.random((20, 20)) < 0.15).astype(np.int32)               # general undirected graph with cycles
    cases["cyclic20"] = np.maximum(g, g.T)
    cases["bool9"] = syn.tree_adjacency(syn.random_tree(9, rng)).astype(bool)
    return cases


def apsp_digest_cases():
    rng = np.random.default_rng(4321)
    cases = {}
    for n in (128, 256):
        cases[f"tree{n}"] = syn.tree_adjacency(syn.random_tree(n, rng))
    cases["chain512"] = syn.tree_adjacency(syn.chain_tree(512))      # distances hit the 510 clamp
    cases["forest300"] = syn.forest_adjacency(300, 4, rng)
    return cases


def digest(a: np.ndarray) -> str:
    import hashlib

    return hashlib.sha256(np.ascontiguousarray(a.astype(np.int64)).tobytes()).hexdigest()


# ------------------------------------------------------------------ collate
COLLATE_CASES = {
    # name: (sizes, seed, spatial_pos_max, kwargs)
    "ragged_sp5": ([5, 12, 9], 11, 5, dict(n_images=1)),
    "ragged_sp3": ([7, 3, 14, 1], 12, 3, dict(image_frac=0.3)),
    "equal_sp10": ([8, 8], 13, 10, dict(with_images=False)),
    "deep_sp5": ([40], 14, 5, dict(n_images=2)),
}


def collate_items(name, list_distance=False):
    sizes, seed, spmax, kw = COLLATE_CASES[name]
    return syn.synthetic_items(sizes, seed, list_distance=list_distance, **kw), spmax


# ------------------------------------------------------------------ small graph modules
def graph_module_inputs(B=3, N=8, C=64, H=4, seed=5, n_pad=(0, 3, 5)):
    """Inputs for GraphNodeFeature / GraphAttnBias / MHA / encoder-layer parity at toy size."""
    g = torch.Generator().manual_seed(seed)
    T = N + 1
    x = torch.randn(B, N, C, generator=g)
    in_deg = torch.randint(1, 6, (B, N), generator=g)
    spatial = torch.randint(1, 22, (B, N, N), generator=g).int()
    attn_bias = torch.zeros(B, T, T)
    far = torch.rand(B, N, N, generator=g) < 0.25
    far = far & ~torch.eye(N, dtype=torch.bool)
    attn_bias[:, 1:, 1:][far] = float("-inf")
    pad_mask = torch.zeros(B, T, dtype=torch.bool)
    for b, k in enumerate(n_pad):
        if k:
            n = N - k
            x[b, n:] = 0
            in_deg[b, n:] = 0
            spatial[b, n:, :] = 0
            spatial[b, :, n:] = 0
            attn_bias[b, n + 1:, :] = float("-inf")
            attn_bias[b, :, n + 1:] = float("-inf")
            attn_bias[b, n + 1:, : n + 1] = 0
            pad_mask[b, n + 1:] = True
    q = torch.randn(T, B, C, generator=g)
    go = torch.randn(T, B, C, generator=g)
    return dict(x=x, in_degree=in_deg, spatial_pos=spatial, attn_bias=attn_bias, pad_mask=pad_mask, query=q, grad_out=go)


# ------------------------------------------------------------------ full encoder
ENCODER_CASES = {
    # name: dict(sizes, seed, n_images per item, model kwargs)
    "tiny": dict(sizes=[6, 9], seed=21, n_images=1, spmax=5,
                 model=dict(num_fusion_layers=1, num_fusion_stack=1, num_graph_stack=1, bert_layers=3, vit_layers=3)),
    "tiny_stack2": dict(sizes=[5, 4, 7], seed=22, n_images=1, spmax=3,
                        model=dict(num_fusion_layers=3, num_fusion_stack=2, num_graph_stack=2, bert_layers=5, vit_layers=5)),
    "tiny_noimg": dict(sizes=[6, 3], seed=23, n_images=0, spmax=5,
                       model=dict(num_fusion_layers=1, num_fusion_stack=1, num_graph_stack=1, bert_layers=3, vit_layers=3)),
    # BASELINE.json configs[0]: one 32-comment tree, BERT-base, ViT-base, 2 executed graph layers, 3 images
    "cfg1": dict(sizes=[32], seed=24, n_images=3, spmax=5,
                 model=dict(num_fusion_layers=1, num_fusion_stack=1, num_graph_stack=1, bert_layers=12, vit_layers=12)),
}

# Separate fixture file (encoder_big.npz): graphs past one 128-wide attention tile, so an encoder-level check runs the
# persistent forward (attn_fwd3_kernel), the tail-key backward and the n > 128 collate path against the reference.
ENCODER_BIG_CASES = {
    "big2": dict(sizes=[134, 131], seed=25, n_images=2, spmax=5,
                 model=dict(num_fusion_layers=1, num_fusion_stack=1, num_graph_stack=2, bert_layers=3, vit_layers=3)),
}
ENCODER_CASES_ALL = dict(ENCODER_CASES, **ENCODER_BIG_CASES)

COMMON_MODEL = dict(num_in_degree=512, num_out_degree=512, num_edges=512 * 3, num_spatial=512, num_edge_dis=128,
                    num_atoms=512 * 9, num_bottle_neck=4, embedding_dim=768, ffn_embedding_dim=768,
                    num_attention_heads=12, encoder_normalize_before=True, pre_layernorm=False,
                    activation_fn="gelu", edge_type="multi_hop")


def encoder_items(name, list_distance=False):
    c = ENCODER_CASES_ALL[name]
    kw = dict(n_images=c["n_images"]) if c["n_images"] else dict(with_images=False)
    return syn.synthetic_items(c["sizes"], c["seed"], list_distance=list_distance, **kw), c["spmax"]
