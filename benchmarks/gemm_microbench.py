#!/usr/bin/env python
"""csrc/gemm.cu against cuBLAS (torch F.linear) on the feed-forward shapes: fc1 of a fusion-layer BERT block
([238 806, 768] x [3072, 768]^T + bias -> GELU) and of a ViT block (82 410 rows).  CUDA events, L2 flushed."""
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


peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.isfile(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0}
for M, N, K in ((238806, 3072, 768), (238806, 2304, 768), (238806, 768, 768), (238806, 768, 3072), (82410, 3072, 768), (82410, 2304, 768),
                (82410, 768, 768), (82410, 768, 3072), (4128, 768, 768), (1000, 256, 64)):
    g = torch.Generator(device="cuda").manual_seed(M)
    x = (torch.randn(M, K, device="cuda", generator=g)).bfloat16()
    w = (torch.randn(N, K, device="cuda", generator=g) * K ** -0.5).bfloat16()
    b = torch.randn(N, device="cuda", generator=g).bfloat16()
    ref = torch.nn.functional.linear(x, w, b)
    out = ops.gemm_bias_act(x, w, b)
    o2, aux = ops.gemm_bias_act(x, w, b, gelu_act=True)
    refg = torch.nn.functional.gelu(ref.float()).bfloat16()
    flops = 2.0 * M * N * K
    row = dict(M=M, N=N, K=K, max_abs_diff_vs_cublas=float((out.float() - ref.float()).abs().max()),
               aux_diff=float((aux.float() - ref.float()).abs().max()), gelu_diff=float((o2.float() - refg.float()).abs().max()),
               mdt_bias_us=timed(lambda: ops.gemm_bias_act(x, w, b)), cublas_bias_us=timed(lambda: torch.nn.functional.linear(x, w, b)),
               mdt_bias_gelu_us=timed(lambda: ops.gemm_bias_act(x, w, b, gelu_act=True)),
               cublas_plus_mdt_gelu_us=timed(lambda: ops.gelu(torch.nn.functional.linear(x, w, b))))
    row["mdt_tflops"] = flops / row["mdt_bias_us"] / 1e6
    row["cublas_tflops"] = flops / row["cublas_bias_us"] / 1e6
    row["mdt_frac_burst"] = row["mdt_tflops"] / peaks["bf16_tflops"]
    print(json.dumps(row), flush=True)
