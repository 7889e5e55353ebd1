#!/usr/bin/env python
"""Probe for SURVEY 8f N1: which attention call serves the packed text sequences of a fusion layer best on a B200.

    python benchmarks/text_attn_probe.py [--n 4096] [--iters 10] [--json out.json]

Shape of one fusion-layer BERT attention at cfg3: n = 4096 comments, each ``4 bottleneck + U{8..100}`` tokens,
12 heads x 64.  Candidates (forward + backward timed separately with CUDA events, L2 flushed):
  * ``flash_varlen``  flash-attn 2.8 ``flash_attn_varlen_func`` on the packed tokens (what the step runs today)
  * ``cudnn_jagged``  torch SDPA, cuDNN backend, jagged nested tensors over the same packed buffer
  * ``cudnn_padded``  torch SDPA, cuDNN backend, padded ``[n, 104]`` with an additive key mask
  * ``mdt_padded``    this repo's tcgen05 kernel on the padded layout with the key-padding mask (no bias)
Each candidate that fails to run is reported with its error instead of a time.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(fn, iters, flush):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.fill_(1)
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
    ap.add_argument("--n", type=int, default=4096)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--json", default=None)
    ap.add_argument("--skip-padded", action="store_true")
    ap.add_argument("--jagged", action="store_true",
                    help="also try cuDNN SDPA on jagged nested tensors (on torch 2.11 this call went into an fx trace of the "
                         "nested-tensor ops and did not return within 5 minutes on the B200 box)")
    a = ap.parse_args()
    torch.manual_seed(0)
    H, d, nb = 12, 64, 4
    lens = torch.randint(8, 101, (a.n,)) + nb
    cu = torch.zeros(a.n + 1, dtype=torch.int32)
    cu[1:] = torch.cumsum(lens, 0)
    total, Tmax = int(cu[-1]), int(lens.max())
    cu_d = cu.cuda()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
    q, k, v = (torch.randn(total, H, d, device="cuda", dtype=torch.bfloat16).requires_grad_(True) for _ in range(3))
    go = torch.randn(total, H, d, device="cuda", dtype=torch.bfloat16)
    flops = 4.0 * H * d * float((lens.double() ** 2).sum())
    out = dict(n=a.n, total_tokens=total, max_len=Tmax, fwd_gflop=flops / 1e9, cases={})

    def run(name, fwd, grads, gout):
        try:
            o = fwd()
            f_us = timed(fwd, a.iters, flush)
            b_us = timed(lambda: torch.autograd.grad(o, grads, gout, retain_graph=True), a.iters, flush)
            out["cases"][name] = dict(fwd_us=f_us, bwd_us=b_us, fwd_tflops=flops / f_us / 1e6, bwd_tflops=2 * flops / b_us / 1e6)
            return o
        except Exception as ex:  # noqa: BLE001
            out["cases"][name] = dict(error=f"{type(ex).__name__}: {str(ex)[:300]}")
            return None
        finally:
            print(name, json.dumps(out["cases"][name]), flush=True)

    from flash_attn import flash_attn_varlen_func
    ref = run("flash_varlen", lambda: flash_attn_varlen_func(q, k, v, cu_d, cu_d, Tmax, Tmax), (q, k, v), go)

    # this repo's packed-sequence kernels (csrc/attn_varlen.cu): grouped tiles and one sequence per tile
    from multimodaldiscussiontransformer_b200 import ops as mops
    qkv = torch.stack([q.detach(), k.detach(), v.detach()], dim=1).contiguous().requires_grad_(True)
    bytes_fwd = 4.0 * total * H * d * 2            # q, k, v in, o out
    bytes_bwd = 8.0 * total * H * d * 2            # q, k, v, o, do in; dq, dk, dv out
    for group in (True, False):
        plan = mops.varlen_plan(cu_d, total, group=group)
        name = "mdt_varlen" if group else "mdt_varlen_ungrouped"
        o = run(name, lambda: mops.varlen_attention(qkv, plan), (qkv,), go)
        if o is not None:
            c = out["cases"][name]
            c["n_items"] = plan.n_items()
            c["fwd_hbm_gbs"] = bytes_fwd / c["fwd_us"] / 1e3
            c["bwd_hbm_gbs"] = bytes_bwd / c["bwd_us"] / 1e3
            if ref is not None:
                c["max_abs_diff_vs_flash"] = float((o.float() - ref.float()).abs().max())
            print(name, json.dumps(c), flush=True)

    from torch.nn.attention import SDPBackend, sdpa_kernel

    def jag(t):
        return torch.nested.nested_tensor_from_jagged(t, offsets=cu_d.long(), min_seqlen=int(lens.min()), max_seqlen=Tmax).transpose(1, 2)

    def cudnn_jagged():
        with sdpa_kernel(SDPBackend.CUDNN_ATTENTION):
            return F.scaled_dot_product_attention(jag(q), jag(k), jag(v)).transpose(1, 2).values()

    if a.jagged:
        o = run("cudnn_jagged", cudnn_jagged, (q, k, v), go)
        if o is not None and ref is not None:
            out["cases"]["cudnn_jagged"]["max_abs_diff_vs_flash"] = float((o.float() - ref.float()).abs().max())

    if a.skip_padded:
        print(json.dumps(out))
        if a.json:
            json.dump(out, open(a.json, "w"), indent=1)
        return
    # padded layouts
    T = Tmax
    idx = torch.arange(T, device="cuda")[None, :] < lens.cuda()[:, None]          # [n, T] valid
    rows = idx.reshape(-1).nonzero().squeeze(1)

    def pad(t):
        return t.new_zeros(a.n * T, H, d).index_copy(0, rows, t).view(a.n, T, H, d)

    qp, kp, vp = (pad(t.detach()).requires_grad_(True) for t in (q, k, v))
    gop = pad(go)
    mask = torch.zeros(a.n, 1, 1, T, device="cuda", dtype=torch.bfloat16).masked_fill(~idx[:, None, None, :], float("-inf"))

    def cudnn_padded():
        with sdpa_kernel(SDPBackend.CUDNN_ATTENTION):
            return F.scaled_dot_product_attention(qp.transpose(1, 2), kp.transpose(1, 2), vp.transpose(1, 2), attn_mask=mask).transpose(1, 2)

    run("cudnn_padded", cudnn_padded, (qp, kp, vp), gop)

    from multimodaldiscussiontransformer_b200 import ops
    key_pad = ~idx
    qs = (qp.detach() * d ** -0.5).requires_grad_(True)  # the kernel takes q already scaled (multihead_attention.py:137)

    def mdt_padded():
        return ops.biased_attention(qs, kp, vp, key_padding_mask=key_pad)

    try:
        import inspect
        out["mdt_signature"] = str(inspect.signature(ops.biased_attention))
    except Exception:  # noqa: BLE001
        pass
    o = run("mdt_padded", mdt_padded, (qs, kp, vp), gop)
    if o is not None and ref is not None:
        got = o.reshape(a.n * T, H, d).index_select(0, rows)
        out["cases"]["mdt_padded"]["max_abs_diff_vs_flash"] = float((got.float() - ref.float()).abs().max())
    print(json.dumps(out))
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
