#!/usr/bin/env python
"""Weight-gradient GEMM dW = dY^T X: this library's split-K tcgen05 kernel (csrc/wgrad.cu) against cuBLAS (``dy.t() @ x``) on
the shapes of a cfg3 training step.  python benchmarks/wgrad_microbench.py [--json out.json]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from multimodaldiscussiontransformer_b200 import ops  # noqa: E402


def timeit(fn, iters, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    ap.add_argument("--iters", type=int, default=10)
    a = ap.parse_args()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    rows = []
    for M, N, C, what in [(238806, 768, 768, "text out-proj / fc2 input side"), (238806, 2304, 768, "text fused q/k/v"),
                          (238806, 3072, 768, "text fc1"), (238806, 768, 3072, "text fc2"), (82410, 768, 768, "ViT out-proj"),
                          (82410, 2304, 768, "ViT q/k/v"), (82410, 3072, 768, "ViT fc1"), (4128, 768, 768, "graph layers")]:
        g = torch.Generator(device="cuda").manual_seed(M % 1000 + N)
        dy = torch.randn(M, N, device="cuda", generator=g).bfloat16()
        x = torch.randn(M, C, device="cuda", generator=g).bfloat16()
        ref = dy.t() @ x
        got = ops.wgrad_gemm(dy, x)
        exact = (dy.double().t() @ x.double()) if M <= 90000 and N * C <= 768 * 768 else None
        err = float((got - ref.float()).abs().max() / ref.float().abs().max())
        out = torch.zeros(N, C, dtype=torch.float32, device="cuda")
        t_own = timeit(lambda: ops.wgrad_gemm(dy, x, out), a.iters, flush)
        t_own_full = timeit(lambda: ops.wgrad_gemm(dy, x).to(torch.bfloat16), a.iters, flush)
        t_lib = timeit(lambda: dy.t() @ x, a.iters, flush)
        fl = 2.0 * M * N * C
        row = dict(M=M, N=N, C=C, what=what, mdt_us=t_own, mdt_with_zero_and_cast_us=t_own_full, cublas_us=t_lib,
                   mdt_tflops=fl / t_own / 1e6, cublas_tflops=fl / t_lib / 1e6, max_rel_diff_vs_cublas=err)
        if exact is not None:
            row["mdt_rel_err_vs_fp64"] = float((got.double() - exact).abs().max() / exact.abs().max())
            row["cublas_rel_err_vs_fp64"] = float((ref.double() - exact).abs().max() / exact.abs().max())
        rows.append(row)
        print(json.dumps(row))
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
