#!/usr/bin/env python
"""Preprocessing kernels alone: the Floyd-Warshall sweep of BASELINE.json configs[1] and GraphAttnBias.

    python benchmarks/prep_microbench.py [--batch 256] [--n 16 32 64 128 256 512] [--iters 10] [--json out.json]

* APSP (``mdt_apsp_batched``, csrc/apsp.cu; replaces algos.pyx:7-52): synthetic pruned discussion trees
  (SURVEY §8d ``tree(n, seed)``, seeds 0..batch-1), uint8 adjacency resident in HBM, int64 ``M`` out (the
  Cython signature's dtype), with and without the predecessor matrix.  Algorithmic work per graph: ``n^3``
  min-plus relaxations; algorithmic bytes ``n^2`` (adjacency) + ``8 n^2`` (``M``) [+ ``8 n^2`` for ``path``].
  ``alu_frac`` is against ONE stated integer-ALU bound, ``148 SMs x 128 lanes x 2 relaxations x max SM clock``
  = 74 T relaxations/s: one ``VIADDMNMX.U16x2`` (packed add + min of two uint16 pairs = 2 relaxations) per lane per
  clock, nothing else issued (the same denominator as DESIGN.md section 4; round 1 quoted 37 T/s here and 74 T/s
  there); ``hbm_frac`` against the measured copy bandwidth.  Size-independent checks after every launch: ``M`` symmetric, zero diagonal, ``M[0, i]`` = tree
  depth, no entry above 510.  (Bit-exactness against the oracle is ``tests/test_gpu_graph_kernels.py``.)
* GraphAttnBias (``mdt_attn_bias_fwd/bwd``, csrc/attn_bias.cu; graphormer_layers.py:86-110): algorithmic bytes
  forward ``B N^2 4 + B T^2 e`` read + ``B H T^2 e`` written, backward ``B H T^2 e + B N^2 4`` read.

Times are CUDA events on the launching stream, L2 flushed (512 MiB write) before every timed call.
``bench.py --prep`` runs this file's sweep and adds the CPU oracle beside it.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from multimodaldiscussiontransformer_b200 import ops  # noqa: E402
from multimodaldiscussiontransformer_b200.data import algos, synthetic  # noqa: E402

SM_COUNT, LANES = 148, 128


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        d = json.load(open(path))
        return d["hbm_gbs"], d.get("sm_max_mhz", 1965.0), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


class Timer:
    def __init__(self):
        self.flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")

    def __call__(self, fn, iters, warmup=3):
        for _ in range(warmup):
            fn()
        ts = []
        for _ in range(iters):
            self.flush.fill_(1)
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            fn()
            e.record()
            torch.cuda.synchronize()
            ts.append(s.elapsed_time(e))
        ts.sort()
        return ts[len(ts) // 2]  # median, ms


def tree_batch(n: int, batch: int):
    parents = [synthetic.random_tree(n, np.random.default_rng(seed)) for seed in range(batch)]
    adj = np.stack([synthetic.tree_adjacency(p) for p in parents])
    depth = np.stack([synthetic.tree_depth(p) for p in parents])
    return adj, depth


def check_apsp(M: torch.Tensor, depth: np.ndarray):
    assert bool((M == M.transpose(1, 2)).all()), "M not symmetric"
    assert bool((M.diagonal(dim1=1, dim2=2) == 0).all()), "diagonal not zero"
    assert int(M.max()) <= algos.UNREACHABLE
    assert bool((M[:, 0, :].cpu() == torch.from_numpy(depth).long()).all()), "M[0, i] != depth(i)"


def apsp_sweep(sizes, batch, iters, timer, cpu_fn=None):
    hbm, mhz, src = peaks()
    alu_peak = SM_COUNT * LANES * 2 * mhz * 1e6  # relaxations / s: one VIADDMNMX.U16x2 (2 relaxations) per lane-clock
    rows = []
    for n in sizes:
        adj_np, depth = tree_batch(n, batch)
        adj = torch.from_numpy(adj_np).cuda()
        for need_path in (False, True):
            M, path = algos.floyd_warshall_batched(adj, need_path)
            check_apsp(M, depth)
            ms = timer(lambda: algos.floyd_warshall_batched(adj, need_path), iters)
            relax = float(batch) * n ** 3
            nbytes = float(batch) * n * n * (1 + 8 + (8 if need_path else 0))
            row = dict(kernel="apsp", n=n, batch=batch, path=need_path, us=ms * 1e3, graphs_per_s=batch / ms * 1e3,
                       grelax_per_s=relax / ms / 1e6, alu_frac=relax / (ms * 1e-3) / alu_peak,
                       gbs=nbytes / ms / 1e6, hbm_frac=nbytes / ms / 1e6 / hbm, peak_source=src)
            if cpu_fn is not None and not need_path:
                row["cpu"] = cpu_fn(adj_np)
            rows.append(row)
            print(json.dumps(row), flush=True)
    return rows


def attn_bias_sweep(shapes, iters, timer, H=12, num_spatial=512, spatial_pos_max=5):
    hbm, _, src = peaks()
    rows = []
    for B, N in shapes:
        T = N + 1
        g = torch.Generator(device="cuda").manual_seed(N)
        sp = torch.randint(0, 22, (B, N, N), device="cuda", dtype=torch.int32, generator=g)
        for dtype in (torch.bfloat16, torch.float32):
            e = 2 if dtype == torch.bfloat16 else 4
            ab = torch.zeros(B, T, T, device="cuda", dtype=dtype)
            ab[:, 1:, 1:][sp > 2 * spatial_pos_max] = float("-inf")
            table = (torch.randn(num_spatial, H, device="cuda", generator=g) * 0.02).requires_grad_(True)
            virt = (torch.randn(1, H, device="cuda", generator=g) * 0.02).requires_grad_(True)
            out = ops.attn_bias(ab, sp, table, virt)
            go = torch.randn(out.shape, device="cuda", dtype=dtype, generator=g)
            f_ms = timer(lambda: ops.attn_bias(ab, sp, table, virt), iters)

            def bwd():
                torch.autograd.grad(out, (table, virt), go, retain_graph=True)

            b_ms = timer(bwd, iters)
            fb = B * N * N * 4.0 + B * T * T * e + float(B) * H * T * T * e
            bb = float(B) * H * T * T * e + B * N * N * 4.0
            row = dict(kernel="attn_bias", B=B, N=N, H=H, dtype=str(dtype).split(".")[-1], fwd_us=f_ms * 1e3,
                       bwd_us=b_ms * 1e3, fwd_gbs=fb / f_ms / 1e6, bwd_gbs=bb / b_ms / 1e6,
                       fwd_frac=fb / f_ms / 1e6 / hbm, bwd_frac=bb / b_ms / 1e6 / hbm, fwd_mb=fb / 1e6, bwd_mb=bb / 1e6,
                       peak_gbs=hbm, peak_source=src,
                       note="bwd time includes the two torch.zeros fills of the gradient tables")
            rows.append(row)
            print(json.dumps(row), flush=True)
            del out, go, ab
    return rows


def main(cpu_fn=None, argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=256)
    ap.add_argument("--n", type=int, nargs="+", default=[16, 32, 64, 128, 256, 512])
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", default=None)
    a, _ = ap.parse_known_args(argv)
    timer = Timer()
    rows = apsp_sweep(a.n, a.batch, a.iters, timer, cpu_fn)
    rows += attn_bias_sweep([(32, 128), (32, 256), (32, 512)], a.iters, timer)
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)
    return rows


if __name__ == "__main__":
    main()
