#!/usr/bin/env python
"""Condense attn_microbench JSON lines: T fwd_us fwd_TF bwd_us bwd_TF"""
import json, sys
for l in sys.stdin:
    try:
        d = json.loads(l)
        print(d["B"], d["T"], d["mode"], "fwd %.1f us %.1f TF | bwd %.1f us %.1f TF" % (d["fwd_us"], d["fwd_tflops"], d["bwd_us"], d["bwd_tflops"]))
    except Exception:
        print(l.rstrip()[:200])
