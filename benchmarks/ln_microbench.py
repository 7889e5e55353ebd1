#!/usr/bin/env python
"""Dropout + residual + LayerNorm kernels alone (csrc/layernorm.cu) against the HBM roofline.

    python benchmarks/ln_microbench.py [--rows 240000] [--C 768] [--p 0.1] [--iters 20] [--json out.json]

Times the C-ABI calls with the CUDA events the wrappers record (ops.ATTN_EVENTS); algorithmic bytes:
forward x, residual, y, z (2 B each) + mean / rstd; backward dy, z, dx, d_residual + mean / rstd.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from multimodaldiscussiontransformer_b200 import ops  # noqa: E402


def hbm_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        return json.load(open(path))["hbm_gbs"], "measured"
    return 6650.0, "fallback"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, nargs="+", default=[240000, 81000, 4128])
    ap.add_argument("--C", type=int, default=768)
    ap.add_argument("--p", type=float, default=0.1)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    peak, src = hbm_peak()
    out = []
    for rows in a.rows:
        for residual in (True, False):
            p = a.p if residual else 0.0
            x = torch.randn(rows, a.C, device="cuda").bfloat16().requires_grad_(True)
            r = torch.randn(rows, a.C, device="cuda").bfloat16().requires_grad_(True) if residual else None
            w = torch.ones(a.C, device="cuda", dtype=torch.bfloat16, requires_grad=True)
            b = torch.zeros(a.C, device="cuda", dtype=torch.bfloat16, requires_grad=True)
            gy = torch.randn(rows, a.C, device="cuda").bfloat16()
            for _ in range(3):
                y = ops.dropout_add_layer_norm(x, r, w, b, p, True)
                y.backward(gy)
            torch.cuda.synchronize()
            ops.ATTN_EVENTS = []
            for _ in range(a.iters):
                torch.cuda._sleep(200000)
                y = ops.dropout_add_layer_norm(x, r, w, b, p, True)
                torch.cuda._sleep(200000)
                y.backward(gy)
                torch.cuda.synchronize()
            ev, ops.ATTN_EVENTS = ops.ATTN_EVENTS, None
            f = [s.elapsed_time(e) for k, s, e, *_ in ev if k == "ln_fwd"]
            g = [s.elapsed_time(e) for k, s, e, *_ in ev if k == "ln_bwd"]
            f_ms, g_ms = sum(f) / len(f), sum(g) / len(g)
            fb = rows * a.C * 2.0 * (4 if residual else 2) + rows * 8.0
            gb = rows * a.C * 2.0 * (4 if residual else 3) + rows * 8.0
            row = dict(rows=rows, C=a.C, residual=residual, p=p, fwd_us=f_ms * 1e3, bwd_us=g_ms * 1e3,
                       fwd_gbs=fb / f_ms / 1e6, bwd_gbs=gb / g_ms / 1e6, fwd_frac=fb / f_ms / 1e6 / peak,
                       bwd_frac=gb / g_ms / 1e6 / peak, peak_gbs=peak, peak_source=src)
            out.append(row)
            print(json.dumps(row))
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
