#!/usr/bin/env python
"""cuBLAS on the dense-layer shapes of a cfg3 step (isolated, L2 flushed): the headroom an own GEMM would have."""
import json
import torch

flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")


def timeit(fn, iters=8):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


for M in (238806, 82410):
    for N, K, what in [(768, 768, "o-proj fwd / dgrad"), (768, 3072, "fc2 fwd, fc1 dgrad"), (768, 2304, "qkv dgrad"), (3072, 768, "fc1 fwd, fc2 dgrad"),
                       (2304, 768, "qkv fwd")]:
        x = torch.randn(M, K, device="cuda").bfloat16()
        w = torch.randn(N, K, device="cuda").bfloat16()
        wt = w.t().contiguous()            # [K, N]: the NN (dgrad) form x @ wt
        b = torch.randn(N, device="cuda").bfloat16()
        acc = torch.randn(M, N, device="cuda").bfloat16()
        t_f = timeit(lambda: torch.nn.functional.linear(x, w, b))
        t_nn = timeit(lambda: x @ wt)
        t_add = timeit(lambda: acc.addmm_(x, wt))
        fl = 2.0 * M * N * K
        print(json.dumps(dict(M=M, N=N, K=K, what=what, linear_bias_us=t_f, linear_tflops=fl / t_f / 1e6, nn_us=t_nn, nn_tflops=fl / t_nn / 1e6,
                              addmm_us=t_add, addmm_tflops=fl / t_add / 1e6)))
