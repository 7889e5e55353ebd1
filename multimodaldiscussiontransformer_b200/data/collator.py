"""``mDT/src/data/collator.py`` + ``pyg_datasets/pre_processing.py`` on the B200 (items are objects with the
reference item's fields, ``edge_index`` instead of ``distance_matrix``; discussion trees only -- see
``check_discussion_tree``).

The reference builds ``spatial_pos`` / ``distance`` / ``attn_bias`` / ``in_degree`` per item on
the CPU with Python ``map`` over N^2 pairs (pre_processing.py:39-56) and pads them per batch
in Python loops (collator.py:122-164).  Here ``preprocess_item`` only keeps the raw edge list
(dataloader workers are forked and must not touch CUDA, src/__init__.py:8) and ``collator``
ships edge lists + token ids to the device, where ONE kernel launch (``mdt_collate_graphs``)
produces every padded graph tensor of the batch.  Output keys, shapes and dtypes are those of
collator.py:166-179; extra keys (``bias_code``, ``node_index``, ``n_nodes``) feed the fused
attention and the sync-free gather/scatter.
"""

from __future__ import annotations

from typing import Any, Dict, List, Sequence

import numpy as np
import torch

from .. import _lib


def _np(a) -> np.ndarray:
    if isinstance(a, torch.Tensor):
        return a.detach().cpu().numpy()
    return np.asarray(a)


def check_discussion_tree(edge_index, n: int, where: str = "item") -> None:
    """Raise unless ``edge_index`` ([2, E] int) is an undirected tree over nodes ``0..n-1`` listed in both directions.

    The reference takes ``spatial_pos`` / ``distance`` from the dataset's precomputed ``distance_matrix`` of (hops up,
    hops down) pairs (hateful_discussions.py:151-159,242-264; pre_processing.py:39-56).  This package re-derives the
    pairs on the device from shortest-path lengths and depths (depth = distance to node 0, the post), which equals the
    reference's construction exactly when the graph is a tree rooted at node 0 -- what a discussion is
    (hateful_discussions.py:116-148).  Anything else (endpoints out of range, self loops, a missing reverse edge,
    a cycle, several components) would silently give different numbers, so it is refused here, on the host, before any
    copy is issued."""
    e = _np(edge_index).reshape(2, -1).astype(np.int64, copy=False)
    E = e.shape[1]
    if n < 1:
        raise ValueError(f"{where}: a discussion needs at least its post (n = {n})")
    if E != 2 * (n - 1):
        raise ValueError(f"{where}: a tree over {n} comments has {2 * (n - 1)} directed edges, edge_index has {E}")
    if E == 0:
        return
    src, dst = e[0], e[1]
    if src.min() < 0 or dst.min() < 0 or src.max() >= n or dst.max() >= n:
        raise ValueError(f"{where}: edge endpoints must lie in [0, {n})")
    if np.any(src == dst):
        raise ValueError(f"{where}: self loops are not part of a discussion tree")
    fwd, rev = np.sort(src * n + dst), np.sort(dst * n + src)
    if np.any(fwd[1:] == fwd[:-1]) or not np.array_equal(fwd, rev):
        raise ValueError(f"{where}: edge_index must list every reply edge once in each direction")
    # n - 1 undirected edges: connected <=> acyclic <=> tree.  Union-find over the edges.
    parent = list(range(n))

    def find(a):
        while parent[a] != a:
            parent[a] = parent[parent[a]]
            a = parent[a]
        return a

    comps = n
    for a, b in zip(src.tolist(), dst.tolist()):
        if a < b:
            ra, rb = find(a), find(b)
            if ra != rb:
                parent[ra] = rb
                comps -= 1
    if comps != 1:
        raise ValueError(f"{where}: the reply graph is not connected to the post (node 0): {comps} components")


def preprocess_item(item):
    """``preprocess_item`` (pre_processing.py:23-69): nothing is computed on the host any more.

    The item keeps ``edge_index`` and ``x``; the O(N^2) tensors are built on the device by
    ``collator``.  Returned unchanged so dataset wrappers can keep calling it.  A precomputed
    ``item.distance_matrix`` is not read: the hop pairs are re-derived from ``edge_index``, which
    is exact for trees rooted at node 0 only -- ``HostBatch`` / ``collator`` check that
    (``check_discussion_tree``) and raise for anything else.
    """
    return item


def collate_graph_tensors(edge_index: Sequence[Any], n_nodes: Sequence[int], spatial_pos_max: int, device,
                          flavour: int = _lib.SPATIAL_CANTOR, bias_dtype=torch.float32, with_code: bool = True,
                          validate: bool = True):
    """Edge lists of B graphs -> padded ``spatial_pos``, ``attn_bias``, ``in_degree`` (+ ``bias_code``).

    ``flavour=SPATIAL_CANTOR`` (mDT) needs rooted trees (``check_discussion_tree``; ``validate=False`` skips the check
    for callers that have already made it).  ``flavour=SPATIAL_SPD`` (upstream Graphormer: ``spatial_pos`` = shortest-path
    length + 1) takes any undirected graph, but endpoints must still lie inside each graph.  Bias codes are one byte:
    unmasked positions carry ``spatial_pos <= spatial_pos_max``, and the fused attention kernels index a 62-entry
    table and bin LUT gradients for codes 0..22 (include/mdt_b200.h), so ``with_code`` needs ``spatial_pos_max <= 22``
    in the SPD flavour (the Cantor flavour never exceeds 21)."""
    lib = _lib.load()
    B = len(n_nodes)
    if with_code and flavour == _lib.SPATIAL_SPD and spatial_pos_max > _lib.CODE_GRAD_MAX:
        raise ValueError(f"bias codes of the SPD flavour need spatial_pos_max <= {_lib.CODE_GRAD_MAX} (got {spatial_pos_max}): "
                         "pass with_code=False and use the materialised GraphAttnBias.forward")
    for b, (ei, n) in enumerate(zip(edge_index, n_nodes)):
        if flavour == _lib.SPATIAL_CANTOR and validate:
            check_discussion_tree(ei, int(n), f"graph {b}")
        else:
            e = _np(ei).reshape(2, -1)
            if e.size and (e.min() < 0 or e.max() >= int(n)):
                raise ValueError(f"graph {b}: edge endpoints must lie in [0, {int(n)})")
    N = int(max(n_nodes))
    T = N + 1
    srcs, dsts, offs = [], [], [0]
    for ei in edge_index:
        e = _np(ei).astype(np.int32, copy=False).reshape(2, -1)
        srcs.append(e[0])
        dsts.append(e[1])
        offs.append(offs[-1] + e.shape[1])
    E = offs[-1]
    # one pinned staging buffer, one H2D copy: [src | dst | offsets | n_nodes]
    host = torch.empty(2 * E + (B + 1) + B, dtype=torch.int32).pin_memory()
    hv = host.numpy()
    if E:
        hv[:E] = np.concatenate(srcs)
        hv[E:2 * E] = np.concatenate(dsts)
    hv[2 * E:2 * E + B + 1] = offs
    hv[2 * E + B + 1:] = np.asarray(n_nodes, dtype=np.int32)
    dev = host.to(device, non_blocking=True)
    src_d, dst_d = dev[:E], dev[E:2 * E]
    off_d, n_d = dev[2 * E:2 * E + B + 1], dev[2 * E + B + 1:]
    spatial_pos = torch.empty((B, N, N), dtype=torch.int32, device=device)
    attn_bias = torch.empty((B, T, T), dtype=bias_dtype, device=device)
    in_degree = torch.empty((B, N), dtype=torch.int64, device=device)
    pitch = (T + 15) // 16 * 16
    code = torch.empty((B, T, pitch), dtype=torch.uint8, device=device) if with_code else None
    n8 = (N + 7) // 8 * 8
    ws = torch.empty((B, n8, n8), dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        _lib.check(lib.mdt_collate_graphs(_lib.ptr(src_d), _lib.ptr(dst_d), _lib.ptr(off_d), _lib.ptr(n_d), B, N,
                                          int(spatial_pos_max), flavour, _lib.ptr(spatial_pos), _lib.ptr(attn_bias),
                                          _lib.dtype_code(bias_dtype), _lib.ptr(in_degree), _lib.ptr(code), pitch,
                                          _lib.ptr(ws), _lib.stream()), "mdt_collate_graphs")
    return dict(spatial_pos=spatial_pos, attn_bias=attn_bias, in_degree=in_degree, bias_code=code, n_nodes=n_d)


class HostBatch:
    """One batch packed into pinned host buffers by the dataloader side (no CUDA calls; safe in forked workers
    when ``pin=False``).  ``tensors`` is what crosses PCIe each step."""

    def __init__(self, items: List[Any], pin: bool = True, image_dtype: torch.dtype = torch.float32):
        """``image_dtype``: dtype the images cross PCIe in.  The reference ships fp32 (collator.py:144-148); the model casts
        them to its compute dtype before the first ViT layer anyway (multigraphormer_graph_encoder.py: ``x_images.to(dtype)``),
        so a bf16 model loses nothing by staging them as bf16 -- and 97 % of a batch's bytes are images."""
        items = [it for it in items if it is not None]                      # dataset.py:160,195
        self.B = len(items)
        self.sizes = [int(_np(it.x["input_ids"]).shape[0]) for it in items]
        self.N = max(self.sizes)                                             # collator.py:127
        self.L = int(_np(items[0].x["input_ids"]).shape[1])
        B, N, ns = self.B, self.N, self.sizes
        srcs, dsts, offs = [], [], [0]
        for b, it in enumerate(items):
            check_discussion_tree(it.edge_index, ns[b], f"item {b} (idx {int(it.idx)})")
            e = _np(it.edge_index).astype(np.int32, copy=False).reshape(2, -1)
            srcs.append(e[0])
            dsts.append(e[1])
            offs.append(offs[-1] + e.shape[1])
        self.E = E = offs[-1]
        self.n_real = n_real = sum(ns)
        node_index = np.concatenate([b * N + np.arange(n, dtype=np.int32) for b, n in enumerate(ns)])
        has_img = np.concatenate([_np(it.x_image_index).reshape(-1) for it in items]).astype(bool)
        img_rows = np.nonzero(has_img)[0].astype(np.int32)
        # real text tokens per comment (attention_mask is a prefix of ones for tokenizer output): lets the fusion
        # layers run on packed tokens without a device->host sync
        att_all = np.concatenate([_np(it.x["attention_mask"]).reshape(n, -1) for it, n in zip(items, ns)]) != 0
        text_lens = att_all.sum(1).astype(np.int32)
        self.text_prefix = bool((att_all == (np.arange(att_all.shape[1])[None, :] < text_lens[:, None])).all())
        self.text_tokens, self.text_max = int(text_lens.sum()), int(text_lens.max()) if n_real else 0
        # int32 staging buffer: [src | dst | offsets | n_nodes | node_index | image rows | idx | text lengths]
        ints = torch.empty(2 * E + (B + 1) + B + n_real + len(img_rows) + B + n_real, dtype=torch.int32)
        iv = ints.numpy()
        o = 0
        for arr in (np.concatenate(srcs) if E else np.zeros(0, np.int32), np.concatenate(dsts) if E else np.zeros(0, np.int32),
                    np.asarray(offs, np.int32), np.asarray(ns, np.int32), node_index, img_rows,
                    np.asarray([int(it.idx) for it in items], np.int32), text_lens):
            iv[o:o + len(arr)] = arr
            o += len(arr)
        self.n_img_rows = len(img_rows)
        text = torch.empty((3, n_real, self.L), dtype=torch.int64)          # node-major text
        tv = text.numpy()
        o = 0
        for it, n in zip(items, ns):
            for k, key in enumerate(("input_ids", "token_type_ids", "attention_mask")):
                tv[k, o:o + n] = _np(it.x[key])
            o += n
        imgs = [it.x_images if isinstance(it.x_images, torch.Tensor) else torch.from_numpy(_np(it.x_images)) for it in items]
        imgs = [im for im in imgs if not bool(torch.all(im.eq(0)))]          # collator.py:144
        images = torch.cat(imgs).to(image_dtype) if imgs else None
        flags = torch.from_numpy(np.stack([has_img, np.concatenate(
            [_np(getattr(it, "y_mask", None) if getattr(it, "y_mask", None) is not None else np.zeros(n, bool)).reshape(-1)
             for it, n in zip(items, ns)]).astype(bool)]))                    # [2, n_real] bool: image flag, y_mask
        self.has_y_mask = all(getattr(it, "y_mask", None) is not None for it in items)
        y = torch.from_numpy(np.concatenate([_np(it.y).reshape(-1) for it in items]).astype(np.float32))
        self.tensors = dict(ints=ints, text=text, images=images, flags=flags, y=y)
        if pin:
            self.tensors = {k: (v.pin_memory() if v is not None else None) for k, v in self.tensors.items()}

    def nbytes(self) -> int:
        return sum(v.numel() * v.element_size() for v in self.tensors.values() if v is not None)


class RawBatch:
    """The same buffers resident in HBM (``HostBatch.tensors`` after one async copy each).

    With ``stream`` the copies are issued on that (copy) stream, so the batch of step i+1 crosses PCIe under the
    kernels of step i; ``wait()`` makes the consumer's current stream wait for them and tells the caching allocator
    that the buffers are in use there."""

    def __init__(self, host: HostBatch, device, stream=None):
        self.host = host
        self.device = device
        self.ready = None
        if stream is None:
            self.t = {k: (v.to(device, non_blocking=True) if v is not None else None) for k, v in host.tensors.items()}
        else:
            with torch.cuda.stream(stream):
                self.t = {k: (v.to(device, non_blocking=True) if v is not None else None) for k, v in host.tensors.items()}
                self.ready = torch.cuda.Event()
                self.ready.record(stream)

    def wait(self) -> "RawBatch":
        if self.ready is not None:
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self.ready)
            for v in self.t.values():
                if v is not None:
                    v.record_stream(cur)
            self.ready = None
        return self


def collate_device(raw: RawBatch, spatial_pos_max: int = 10) -> Dict[str, Any]:
    """Device half of ``collator``: every padded tensor of collator.py:166-179 from the raw buffers, sync-free."""
    lib = _lib.load()
    hb, t, device = raw.host, raw.t, raw.device
    B, N, E, L, n_real = hb.B, hb.N, hb.E, hb.L, hb.n_real
    T = N + 1
    ints = t["ints"]
    o = 0
    src_d, dst_d = ints[o:o + E], ints[o + E:o + 2 * E]
    o += 2 * E
    off_d = ints[o:o + B + 1]
    o += B + 1
    n_d = ints[o:o + B]
    o += B
    node_index = ints[o:o + n_real]
    o += n_real
    img_rows = ints[o:o + hb.n_img_rows]
    o += hb.n_img_rows
    idx = ints[o:o + B]
    o += B
    text_lens = ints[o:o + n_real]
    spatial_pos = torch.empty((B, N, N), dtype=torch.int32, device=device)
    attn_bias = torch.empty((B, T, T), dtype=torch.float32, device=device)
    in_degree = torch.empty((B, N), dtype=torch.int64, device=device)
    pitch = (T + 15) // 16 * 16
    code = torch.empty((B, T, pitch), dtype=torch.uint8, device=device)
    n8 = (N + 7) // 8 * 8
    ws = torch.empty((B, n8, n8), dtype=torch.uint8, device=device)
    with torch.cuda.device(device):
        _lib.check(lib.mdt_collate_graphs(_lib.ptr(src_d), _lib.ptr(dst_d), _lib.ptr(off_d), _lib.ptr(n_d), B, N,
                                          int(spatial_pos_max), _lib.SPATIAL_CANTOR, _lib.ptr(spatial_pos),
                                          _lib.ptr(attn_bias), _lib.F32, _lib.ptr(in_degree), _lib.ptr(code), pitch,
                                          _lib.ptr(ws), _lib.stream()), "mdt_collate_graphs")
    text_d = t["text"]
    padded = torch.zeros((3, B * N, L), dtype=torch.int64, device=device)
    padded.index_copy_(1, node_index.long(), text_d)                        # collator.py:131-139
    padded = padded.view(3, B, N, L)
    token_mask = ~padded[0].eq(0).all(dim=2)                                # collator.py:141
    out = dict(
        idx=idx.long(),
        attn_bias=attn_bias,
        spatial_pos=spatial_pos,
        in_degree=in_degree,
        out_degree=in_degree,                                               # collator.py:171
        x_token_mask=token_mask,
        x=padded[0],
        x_token_type_ids=padded[1],
        x_attention_mask=padded[2],
        x_images=t["images"],
        x_image_indexes=t["flags"][0],
        y=t["y"],
        # extras of the B200 path: in-kernel bias synthesis and sync-free row moves
        bias_code=code,
        node_index=node_index,
        n_nodes=n_d,
        x_nodes=text_d,
        x_image_rows=img_rows,
        # (lengths on the device, total and longest on the host) when every attention_mask is a prefix of ones
        x_text_pack=(text_lens, hb.text_tokens, hb.text_max) if hb.text_prefix else None,
    )
    if hb.has_y_mask:
        out["y_mask"] = t["flags"][1]                                       # dataset.py:210-213
    return out


def collator(items: List[Any], spatial_pos_max: int = 10, device=None) -> Dict[str, Any]:
    """``collator(items, spatial_pos_max)`` (collator.py:69-179) producing CUDA tensors.

    ``items`` carry ``idx, x{input_ids,token_type_ids,attention_mask}, edge_index, x_image_index,
    x_images, y`` (and optionally ``y_mask``, dataset.py:210-213).  Host packing -> one async copy per
    buffer -> one kernel launch for all graph tensors.
    """
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device())
    return collate_device(RawBatch(HostBatch(items), device), spatial_pos_max)
