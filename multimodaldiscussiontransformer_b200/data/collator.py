"""Drop-in for ``mDT/src/data/collator.py`` + ``pyg_datasets/pre_processing.py`` on the B200.

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


def preprocess_item(item):
    """``preprocess_item`` (pre_processing.py:23-69): nothing is computed on the host any more.

    The item keeps ``edge_index`` and ``x``; the O(N^2) tensors are built on the device by
    ``collator``.  Returned unchanged so dataset wrappers can keep calling it.
    """
    return item


def collate_graph_tensors(edge_index: Sequence[Any], n_nodes: Sequence[int], spatial_pos_max: int, device,
                          flavour: int = _lib.SPATIAL_CANTOR, bias_dtype=torch.float32, with_code: bool = True):
    """Edge lists of B graphs -> padded ``spatial_pos``, ``attn_bias``, ``in_degree`` (+ ``bias_code``)."""
    lib = _lib.load()
    B = len(n_nodes)
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

    def __init__(self, items: List[Any], pin: bool = True):
        items = [it for it in items if it is not None]                      # dataset.py:160,195
        self.B = len(items)
        self.sizes = [int(_np(it.x["input_ids"]).shape[0]) for it in items]
        self.N = max(self.sizes)                                             # collator.py:127
        self.L = int(_np(items[0].x["input_ids"]).shape[1])
        B, N, ns = self.B, self.N, self.sizes
        srcs, dsts, offs = [], [], [0]
        for it in items:
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
        images = torch.cat(imgs) if imgs else None
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
    """The same buffers resident in HBM (``HostBatch.tensors`` after one async copy each)."""

    def __init__(self, host: HostBatch, device):
        self.host = host
        self.device = device
        self.t = {k: (v.to(device, non_blocking=True) if v is not None else None) for k, v in host.tensors.items()}


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
