"""Drop-in for ``mDT/src/data/algos.pyx`` running on the B200 (``libmdt_b200.so``).

``floyd_warshall(adjacency_matrix) -> (M, path)`` and ``gen_edge_input(max_dist, path,
edge_feat)`` keep the Cython signatures (algos.pyx:7,63): numpy in, numpy int64 out,
``AssertionError`` on a non-square matrix, ``IndexError`` when a path is longer than
``max_dist``.  The batched device entry points (``floyd_warshall_batched``,
``gen_edge_input_batched``) take and return CUDA tensors and never touch the host.
"""

from __future__ import annotations

import numpy as np
import torch

from .. import _lib

UNREACHABLE = 510


def _adj_tensor(adj: torch.Tensor):
    if adj.dtype in (torch.uint8, torch.bool):
        return adj.contiguous().view(torch.uint8), 1
    if adj.dtype == torch.int32:
        return adj.contiguous(), 4
    if adj.dtype == torch.int64:
        return adj.contiguous(), 8
    if adj.dtype in (torch.int8, torch.int16):
        return adj.to(torch.int32).contiguous(), 4
    raise TypeError(f"adjacency must be an integer or bool tensor, got {adj.dtype}")  # casting='safe', algos.pyx:13


def floyd_warshall_batched(adj: torch.Tensor, need_path: bool = False):
    """``adj [B,n,n]`` CUDA integer tensor -> ``(M, path|None)`` int64 ``[B,n,n]`` (bit-exact, algos.pyx:7-52)."""
    assert adj.dim() == 3 and adj.shape[1] == adj.shape[2], "adjacency must be [B, n, n]"
    if not adj.is_cuda:
        raise RuntimeError("floyd_warshall_batched needs a CUDA tensor (there is no CPU path)")
    lib = _lib.load()
    B, n, _ = adj.shape
    a, eb = _adj_tensor(adj)
    M = torch.empty((B, n, n), dtype=torch.int64, device=adj.device)
    path = ws = None
    if B == 0 or n == 0:  # nothing to launch (data_ptr() of an empty tensor is NULL)
        return M, (torch.empty_like(M) if need_path else None)
    if need_path:
        path = torch.empty_like(M)
        ws = torch.empty(max(1, lib.mdt_apsp_workspace_bytes(B, n)), dtype=torch.uint8, device=adj.device)
    with torch.cuda.device(adj.device):
        _lib.check(lib.mdt_apsp_batched(_lib.ptr(a), eb, B, n, _lib.ptr(M), _lib.ptr(path), _lib.ptr(ws), _lib.stream()),
                   "mdt_apsp_batched")
    return M, path


def gen_edge_input_batched(max_dist: int, path: torch.Tensor, edge_feat: torch.Tensor) -> torch.Tensor:
    """``path [B,n,n]``, ``edge_feat [B,n,n,F]`` int64 CUDA -> ``[B,n,n,max_dist,F]`` (algos.pyx:63-89)."""
    lib = _lib.load()
    B, n, _ = path.shape
    F = edge_feat.shape[-1]
    path = path.contiguous()
    edge_feat = edge_feat.to(torch.int64).contiguous()
    out = torch.empty((B, n, n, max_dist, F), dtype=torch.int64, device=path.device)
    status = torch.empty(B, dtype=torch.int32, device=path.device)
    with torch.cuda.device(path.device):
        _lib.check(lib.mdt_gen_edge_input(_lib.ptr(path), _lib.ptr(edge_feat), B, n, F, max_dist, _lib.ptr(out),
                                          _lib.ptr(status), _lib.stream()), "mdt_gen_edge_input")
    if bool(status.any().item()):
        raise IndexError("shortest path longer than max_dist")  # numpy would raise at algos.pyx:87
    return out


# ---------------------------------------------------------------- Cython-signature wrappers
def floyd_warshall(adjacency_matrix):
    """``algos.floyd_warshall`` (algos.pyx:7-52): numpy ``[n,n]`` -> ``(M, path)`` numpy int64."""
    adjacency_matrix = np.asarray(adjacency_matrix)
    (nrows, ncols) = adjacency_matrix.shape
    assert nrows == ncols
    adj = adjacency_matrix.astype(np.int64, order="C", casting="safe", copy=True)
    t = torch.from_numpy(adj).unsqueeze(0).cuda()
    M, path = floyd_warshall_batched(t, need_path=True)
    return M[0].cpu().numpy(), path[0].cpu().numpy()


def gen_edge_input(max_dist, path, edge_feat):
    """``algos.gen_edge_input`` (algos.pyx:63-89): numpy in, numpy int64 ``[n,n,max_dist,F]`` out."""
    path = np.asarray(path)
    edge_feat = np.asarray(edge_feat)
    (nrows, ncols) = path.shape
    assert nrows == ncols
    p = torch.from_numpy(path.astype(np.int64, order="C", casting="safe", copy=True)).unsqueeze(0).cuda()
    e = torch.from_numpy(edge_feat.astype(np.int64, order="C", casting="safe", copy=True)).unsqueeze(0).cuda()
    return gen_edge_input_batched(int(max_dist), p, e)[0].cpu().numpy()
