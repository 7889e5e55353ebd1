#!/usr/bin/env python
"""Timeline of the persistent attention forward: per-CTA %globaltimer stamps (mdt_debug_attn_trace).

    python tools/attn_trace.py [T] [B]
"""
import os, sys
import torch
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
from multimodaldiscussiontransformer_b200 import _lib, ops  # noqa: E402
import attn_microbench as mb  # noqa: E402

T = int(sys.argv[1]) if len(sys.argv) > 1 else 129
B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
lib = _lib.load()
q, k, v, kw = mb.make_inputs(B, 12, T, "code")
for _ in range(3):
    ops.biased_attention(q, k, v, **kw)
torch.cuda.synchronize()
trace = torch.zeros(148, 64, dtype=torch.int64, device="cuda")
lib.mdt_debug_attn_trace(trace.data_ptr())
ops.biased_attention(q, k, v, **kw)
torch.cuda.synchronize()
lib.mdt_debug_attn_trace(None)
tr = trace.cpu()
t0 = int(tr[:, 0][tr[:, 0] > 0].min())
names = ["aux_full", "s_full", "pass1", "p_full", "o_done", "epi_end", "Q_issued", "V_issued", "mma_qfull", "PV_issued"]
for cta in (0, 73, 147):
    row = tr[cta]
    if row[0] == 0:
        continue
    print(f"CTA {cta}: start {int(row[0]) - t0} ns, prologue done {int(row[1]) - t0} ns")
    for n in range(4):
        ev = [(names[k], int(row[2 + 12 * n + k]) - t0) for k in range(10) if row[2 + 12 * n + k] > 0]
        if ev:
            print(f"  item {n}: " + "  ".join(f"{a}={b}" for a, b in sorted(ev, key=lambda x: x[1])))
