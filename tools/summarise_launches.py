#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list of bench.py: one timed training
step (delimited by the collate kernel), shares per kernel name, and the share of libmdt_b200's kernels.

    python tools/summarise_launches.py gpurun_out/launches.csv "<command line that was profiled>" [step_index]
"""
import collections
import csv
import re
import sys

path, cmd = sys.argv[1], sys.argv[2]
which = int(sys.argv[3]) if len(sys.argv) > 3 else 1
rows = []
with open(path) as f:
    lines = [l for l in f if not l.startswith("==")]
for row in csv.DictReader(lines):
    if row.get("Metric Name") == "gpu__time_duration.sum":
        rows.append((row["Kernel Name"], float(row["Metric Value"].replace(",", "")) / 1e3))
idx = [i for i, (n, v) in enumerate(rows) if "apsp_tile_kernel" in n]
lo, hi = idx[which] - 1, idx[which + 1] - 1  # edges_to_adj precedes the tile kernel
step = rows[lo:hi]
tot = sum(v for _, v in step)
print("ncu launch list (gpu__time_duration.sum, --clock-control none) of:", cmd)
print("per-launch times are cold-cache and serialised: compare SHARES, not absolutes")
print(f"one training step = launches [{lo},{hi}): {len(step)} launches, {tot / 1e3:.1f} ms summed device time\n")
agg = collections.defaultdict(lambda: [0, 0.0])


def short(n):
    n = re.sub(r"<.*", "", n)
    n = re.sub(r"\(.*", "", n)
    return n[:70]


for n, v in step:
    a = agg[short(n)]
    a[0] += 1
    a[1] += v
mine = ("mdt::", "attn_fwd_kernel", "attn_bwd_kernel", "apsp", "edges_to_adj", "node_feature", "move_rows", "attn_bias",
        "attn_delta", "attn_dq", "mdt_", "attn::", "ln::", "gelu::", "wgrad::", "gemm::", "gemm2::", "varlen_plan")
print(f"{'kernel':72s} {'count':>6s} {'total us':>11s} {'share':>7s} {'avg us':>9s}")
lib = sum(t for k, (c, t) in agg.items() if "nvjet" in k or "cublas" in k or "cutlass" in k or "cudnn" in k)
aten = sum(t for k, (c, t) in agg.items() if k.startswith("void at::") or "at_cuda_detail" in k or "elementwise_kernel_with_index" in k)
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print(f"{k:72s} {c:6d} {t:11.1f} {100 * t / tot:6.2f}% {t / c:9.1f}")
print("\nkernels of libmdt_b200.so in this step:")
mt = 0
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    if any(m in k for m in mine):
        print(f"{k:72s} {c:6d} {t:11.1f} {100 * t / tot:6.2f}% {t / c:9.1f}")
        mt += t
print(f"libmdt_b200 total: {mt:.1f} us = {100 * mt / tot:.2f}% of the step")
print(f"cuBLAS / cuDNN (nvjet_*, cutlass, cublasLt): {lib:.1f} us = {100 * lib / tot:.2f}%;  ATen element-wise / reduce / index / optimizer: "
      f"{aten:.1f} us = {100 * aten / tot:.2f}%")
