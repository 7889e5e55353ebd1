"""Seeded synthetic discussion trees and mDT items (numpy only, no CUDA).

There is no network on the build or GPU boxes, so every test and the bench run on
synthetic HatefulDiscussions-shaped data.  The shapes follow the reference dataset
builder (``mDT/experiments/hateful_discussions/datasets/hateful_discussions.py``):

* trees are pruned Reddit threads (``Pre-Processing/3-prune-trees.py:16-39``): the
  post is node 0, every comment has at most two children and the tree is at most
  12 deep;
* ``distance_matrix[i][j] = [hops up from i to lca(i,j), hops down from lca to j]``
  (``hateful_discussions.py:155-159,242-264``);
* the graph is undirected, both directions are listed in ``edge_index``
  (``hateful_discussions.py:116-148``);
* text is tokenised to 100 tokens (``:160-166``), images are 3x224x224 (``:172-182``),
  exactly one labelled comment per saved graph (``:202-223``).
"""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional

import numpy as np

TEXT_LEN = 100
VOCAB = 30522
IMG_SHAPE = (3, 224, 224)


def random_tree(n: int, rng: np.random.Generator, max_depth: int = 12, max_children: int = 2) -> np.ndarray:
    """Parent array of a pruned discussion tree: ``parent[0] == -1``, ``parent[i] < i``.

    A comment attaches to a uniformly drawn earlier node whose depth is below
    ``max_depth`` and which has fewer than ``max_children`` children; the post
    (node 0) takes any number of top-level comments.
    """
    parent = np.full(n, -1, dtype=np.int32)
    depth = np.zeros(n, dtype=np.int32)
    nchild = np.zeros(n, dtype=np.int32)
    for i in range(1, n):
        ok = np.flatnonzero((depth[:i] < max_depth) & ((nchild[:i] < max_children) | (np.arange(i) == 0)))
        p = int(ok[rng.integers(len(ok))])
        parent[i] = p
        depth[i] = depth[p] + 1
        nchild[p] += 1
    return parent


def chain_tree(n: int) -> np.ndarray:
    """A path 0-1-2-...: distances reach n-1 (hits the 510 clamp at n=512)."""
    return np.arange(-1, n - 1, dtype=np.int32)


def tree_depth(parent: np.ndarray) -> np.ndarray:
    n = len(parent)
    depth = np.zeros(n, dtype=np.int32)
    for i in range(n):
        if parent[i] >= 0:
            depth[i] = depth[parent[i]] + 1
    return depth


def hop_pairs(parent: np.ndarray) -> np.ndarray:
    """``[n, n, 2]`` int32: (hops up from i to the LCA, hops down from the LCA to j).

    Restates ``get_relative_depth`` + ``spread_downwards``
    (``hateful_discussions.py:242-264``) for an index-ordered tree.
    """
    n = len(parent)
    depth = tree_depth(parent)
    # ancestor bitmap: anc[i, a] = 1 when a is an ancestor-or-self of i
    anc = np.zeros((n, n), dtype=bool)
    for i in range(n):
        if parent[i] >= 0:
            anc[i] = anc[parent[i]]
        anc[i, i] = True
    # depth of the lowest common ancestor = (number of common ancestors) - 1
    common = anc.astype(np.int32) @ anc.astype(np.int32).T
    lca_depth = common - 1
    up = depth[:, None] - lca_depth
    down = depth[None, :] - lca_depth
    return np.stack([up, down], axis=-1).astype(np.int32)


def tree_edge_index(parent: np.ndarray) -> np.ndarray:
    """``[2, 2(n-1)]`` int64: parent->child and child->parent, as ``from_networkx`` lists them."""
    child = np.arange(1, len(parent), dtype=np.int64)
    par = parent[1:].astype(np.int64)
    src = np.concatenate([par, child])
    dst = np.concatenate([child, par])
    return np.stack([src, dst])


def tree_adjacency(parent: np.ndarray, dtype=np.uint8) -> np.ndarray:
    n = len(parent)
    adj = np.zeros((n, n), dtype=dtype)
    e = tree_edge_index(parent)
    adj[e[0], e[1]] = 1
    return adj


def forest_adjacency(n: int, n_components: int, rng: np.random.Generator, dtype=np.uint8) -> np.ndarray:
    """Several disjoint trees in one adjacency matrix (unreachable pairs -> 510)."""
    adj = np.zeros((n, n), dtype=dtype)
    bounds = np.linspace(0, n, n_components + 1).astype(int)
    for a, b in zip(bounds[:-1], bounds[1:]):
        if b - a > 1:
            adj[a:b, a:b] = tree_adjacency(random_tree(b - a, rng), dtype)
    return adj


@dataclass
class DiscussionItem:
    """One discussion graph in the reference's item layout (``pyg_dataset.py:108-118``)."""

    idx: int
    parent: np.ndarray                 # [n] int32 (synthetic only; -1 for the post)
    x: Dict[str, Any]                  # input_ids / token_type_ids / attention_mask  [n, 100] int64
    edge_index: Any                    # [2, E] int64
    edge_attr: Any                     # unused by mDT
    distance_matrix: Any               # [n][n][2] hop pairs (python list in the reference)
    x_image_index: Any                 # [n] float 0/1
    x_images: Any                      # [n_img, 3, 224, 224] float32, or zeros[1,3,224,224] placeholder
    y: Any                             # [n_labelled] float
    y_mask: Any                        # [n] bool
    extra: Dict[str, Any] = field(default_factory=dict)

    @property
    def num_nodes(self) -> int:
        return int(len(self.parent))


def synthetic_item(
    n: int,
    seed: int,
    idx: int = 0,
    image_frac: float = 0.10,
    n_images: Optional[int] = None,
    with_images: bool = True,
    parent: Optional[np.ndarray] = None,
    as_torch: bool = True,
    list_distance: bool = False,
) -> DiscussionItem:
    """A seeded HatefulDiscussions-shaped item (SURVEY.md section 8d 'Synthetic batch')."""
    rng = np.random.default_rng(seed)
    if parent is None:
        parent = random_tree(n, rng)
    n = len(parent)
    lens = rng.integers(8, TEXT_LEN + 1, size=n)
    ids = rng.integers(1, VOCAB, size=(n, TEXT_LEN), dtype=np.int64)
    pos = np.arange(TEXT_LEN)[None, :]
    att = (pos < lens[:, None]).astype(np.int64)
    ids = ids * att
    tti = np.zeros_like(ids)
    has_img = np.zeros(n, dtype=bool)
    if with_images:
        if n_images is None:
            has_img = rng.random(n) < image_frac
        else:
            has_img[rng.choice(n, size=min(n_images, n), replace=False)] = True
    k = int(has_img.sum())
    if k:
        imgs = rng.standard_normal((k,) + IMG_SHAPE, dtype=np.float32)
    else:
        imgs = np.zeros((1,) + IMG_SHAPE, dtype=np.float32)   # placeholder, hateful_discussions.py:182
    lab = int(rng.integers(n))
    y = np.array([float(rng.integers(2))], dtype=np.float32)
    y_mask = np.zeros(n, dtype=bool)
    y_mask[lab] = True
    dm = hop_pairs(parent)
    ei = tree_edge_index(parent)
    x = {"input_ids": ids, "token_type_ids": tti, "attention_mask": att}
    ximg_idx = has_img.astype(np.float32)
    if as_torch:
        import torch

        x = {k_: torch.from_numpy(v) for k_, v in x.items()}
        ei = torch.from_numpy(ei)
        ximg_idx = torch.from_numpy(ximg_idx)
        imgs = torch.from_numpy(imgs)
        y = torch.from_numpy(y)
        y_mask = torch.from_numpy(y_mask)
    return DiscussionItem(
        idx=idx,
        parent=parent,
        x=x,
        edge_index=ei,
        edge_attr=None,
        distance_matrix=dm.tolist() if list_distance else dm,
        x_image_index=ximg_idx,
        x_images=imgs,
        y=y,
        y_mask=y_mask,
    )


def synthetic_items(sizes: List[int], seed: int, **kw) -> List[DiscussionItem]:
    return [synthetic_item(n, seed * 100003 + i, idx=i, **kw) for i, n in enumerate(sizes)]
