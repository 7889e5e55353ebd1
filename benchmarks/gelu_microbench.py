#!/usr/bin/env python
"""GELU kernels of csrc/gelu.cu against ATen at the fusion-layer sizes (238 806 text tokens x 3072; 82 410 ViT tokens x 3072):
CUDA events, L2 flushed, algorithmic bytes = 2 passes (forward) / 3 passes + column sums (backward) of rows x C bf16."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from multimodaldiscussiontransformer_b200 import ops  # noqa: E402

flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")


def timed(fn, iters=10):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.isfile(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
for rows, C in ((238806, 3072), (82410, 3072), (4128, 768)):
    u = torch.randn(rows, C, device="cuda").bfloat16().requires_grad_(True)
    go = torch.randn(rows, C, device="cuda").bfloat16()
    y = ops.gelu(u)
    y2 = torch.nn.functional.gelu(u)
    nb = rows * C * 2.0
    row = dict(rows=rows, C=C,
               mdt_fwd_us=timed(lambda: ops.gelu(u.detach())), aten_fwd_us=timed(lambda: torch.nn.functional.gelu(u.detach())),
               mdt_bwd_us=timed(lambda: torch.autograd.grad(y, u, go, retain_graph=True)),
               aten_bwd_us=timed(lambda: torch.autograd.grad(y2, u, go, retain_graph=True)),
               aten_colsum_us=timed(lambda: go.sum(0)))
    row["mdt_fwd_frac"] = 2 * nb / row["mdt_fwd_us"] / 1e3 / peak
    row["mdt_bwd_frac"] = 3 * nb / row["mdt_bwd_us"] / 1e3 / peak
    print(json.dumps(row))
