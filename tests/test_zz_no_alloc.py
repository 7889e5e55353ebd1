"""``mdt_collate_graphs_static`` -- the collate entry point that allocates nothing (adjacency scratch in a file-scope
device array of ``MDT_COLLATE_STATIC_MAX_BYTES``) -- against ``mdt_collate_graphs`` with a caller-provided workspace.

Both are called through the C ABI on the same device buffers of edges; every output is integer / 0 / -inf, so the bar is
bit equality (the bar of the other collate tests).  Two calls in a row share the static array: the second batch is smaller
and differently shaped, so bytes left by the first would corrupt it if the used part were not cleared.  The file sorts
last on purpose."""

from __future__ import annotations

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_MAX_BYTES = 8 << 20   # MDT_COLLATE_STATIC_MAX_BYTES in include/mdt_b200.h


def _edges_on_device(sizes, seed):
    """Random trees rooted at node 0, both edge directions, as the int32 device arrays the entry points take."""
    rng = np.random.default_rng(seed)
    srcs, dsts, offs = [], [], [0]
    for n in sizes:
        child = np.arange(1, n, dtype=np.int32)
        parent = np.array([rng.integers(0, c) for c in child], dtype=np.int32) if n > 1 else np.zeros(0, np.int32)
        srcs.append(np.concatenate([parent, child]))
        dsts.append(np.concatenate([child, parent]))
        offs.append(offs[-1] + 2 * (n - 1))
    dev = torch.device("cuda", 0)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.int32)).to(dev)
    return t(np.concatenate(srcs)), t(np.concatenate(dsts)), t(np.array(offs)), t(np.array(sizes))


def _outputs(B, N, dev):
    T = N + 1
    pitch = (T + 15) // 16 * 16
    # filled with a pattern: whatever an entry point leaves unwritten shows up as a difference
    return dict(spatial_pos=torch.full((B, N, N), -7, dtype=torch.int32, device=dev),
                attn_bias=torch.full((B, T, T), 3.0, dtype=torch.bfloat16, device=dev),
                in_degree=torch.full((B, N), -7, dtype=torch.int64, device=dev),
                code=torch.full((B, T, pitch), 99, dtype=torch.uint8, device=dev)), pitch


def _run_both(sizes, spmax, seed):
    from multimodaldiscussiontransformer_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda", 0)
    B, N = len(sizes), max(sizes)
    src, dst, off, nn = _edges_on_device(sizes, seed)
    n8 = (N + 7) // 8 * 8
    want, pitch = _outputs(B, N, dev)
    got, _ = _outputs(B, N, dev)
    ws = torch.empty((B, n8, n8), dtype=torch.uint8, device=dev)
    bf16 = _lib.dtype_code(torch.bfloat16)
    with torch.cuda.device(dev):
        _lib.check(lib.mdt_collate_graphs(_lib.ptr(src), _lib.ptr(dst), _lib.ptr(off), _lib.ptr(nn), B, N, spmax, 0,
                                          _lib.ptr(want["spatial_pos"]), _lib.ptr(want["attn_bias"]), bf16,
                                          _lib.ptr(want["in_degree"]), _lib.ptr(want["code"]), pitch, _lib.ptr(ws),
                                          _lib.stream()), "mdt_collate_graphs")
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        _lib.check(lib.mdt_collate_graphs_static(_lib.ptr(src), _lib.ptr(dst), _lib.ptr(off), _lib.ptr(nn), B, N, spmax, 0,
                                                 _lib.ptr(got["spatial_pos"]), _lib.ptr(got["attn_bias"]), bf16,
                                                 _lib.ptr(got["in_degree"]), _lib.ptr(got["code"]), pitch,
                                                 _lib.stream()), "mdt_collate_graphs_static")
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before      # nothing went through torch's allocator either
    for k in want:
        a, b = got[k], want[k]
        if a.dtype == torch.bfloat16:                           # 0 / -inf: compare the bit patterns
            a, b = a.view(torch.int16), b.view(torch.int16)
        assert torch.equal(a, b), k
    assert int((want["spatial_pos"] > 0).sum()) > 0             # the reference entry did produce something


def test_static_collate_matches_workspace_entry_on_two_calls_in_a_row():
    _run_both([128] * 8 + [77, 5, 1, 300], 5, seed=1)          # first call dirties 12 * 304 * 304 bytes of the array
    _run_both([40, 33, 64, 2, 17], 3, seed=2)                   # smaller, other pitch: stale bytes must not leak in
    _run_both([128] * 8 + [77, 5, 1, 300], 5, seed=1)          # and the first batch again after it


def test_static_collate_rejects_sizes_beyond_the_array():
    from multimodaldiscussiontransformer_b200 import _lib

    lib = _lib.load()
    dev = torch.device("cuda", 0)
    B, N = 33, 512                                              # 33 * 512 * 512 bytes > 8 MiB
    assert B * N * N > STATIC_MAX_BYTES
    dummy = torch.zeros(64, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        rc = lib.mdt_collate_graphs_static(_lib.ptr(dummy), _lib.ptr(dummy), _lib.ptr(dummy), _lib.ptr(dummy), B, N, 5, 0,
                                           _lib.ptr(dummy), _lib.ptr(dummy), _lib.dtype_code(torch.bfloat16), _lib.ptr(dummy),
                                           None, 0, _lib.stream())
    assert rc == -1                                             # MDT_EINVAL, nothing launched
    msg = lib.mdt_last_error_string().decode()
    assert "static array holds" in msg and "mdt_collate_graphs" in msg
    torch.cuda.synchronize()
    _run_both([16, 9], 5, seed=3)                               # the entry point still works after the refusal
