#!/usr/bin/env python
"""mDT training throughput on B200s (BASELINE.json: "mDT train discussion-graphs/sec").

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

One step = one pass of the hot path over one synthetic HatefulDiscussions-shaped batch per GPU
(BASELINE.json configs[2]: 32 discussion trees x 128 comments, bf16, the published model:
BERT-base + ViT-base trunks, 8 fusion layers in stacks of 2, 10 executed graph layers, 4 bottleneck
tokens, spatial_pos_max 5): on-device collate (Floyd-Warshall / spatial_pos / degrees / bias codes) ->
forward -> weighted node CE -> backward -> gradient all-reduce -> AdamW on fp32 master weights.
Prints ONE JSON line (rank 0).  `value` has the raw batch resident in HBM; `e2e` starts from pinned
host buffers and reads the loss back every step.  `--impl reference` times the reference's algorithm
on the host cores (oracle/mdt_oracle.py: the CPU restatement pinned to the unmodified reference).
"""

from __future__ import annotations

import argparse
import json
import math
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "mDT train discussion-graphs/sec"
UNIT = "graphs/s"
WORKLOAD = "configs[2]: HatefulDiscussions-shaped synthetic training batch, 32 graphs x 128 comments per GPU, bf16"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=8)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="discussion graphs per GPU per micro-batch (default 32; 16 for cfg5)")
    ap.add_argument("--nodes", type=int, default=128, help="comments per graph")
    ap.add_argument("--image-frac", type=float, default=0.10)
    ap.add_argument("--pool", type=int, default=2, help="distinct pre-packed host batches cycled through")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="cfg3", choices=["cfg3", "cfg5"],
                    help="cfg3 = BASELINE configs[2] (32 x 128 comments); cfg5 = configs[4] (trees of 16..512 comments, --batch 16)")
    ap.add_argument("--update-freq", type=int, default=1, help="micro-batches per optimizer update (published run: 3)")
    ap.add_argument("--grad-dtype", default="bf16", choices=["bf16", "fp32"], help="dtype of the all-reduced gradient buckets")
    ap.add_argument("--no-overlap", action="store_true", help="one all-reduce after the backward instead of overlapped buckets")
    ap.add_argument("--fp32-images", action="store_true", help="ship images as fp32 like the reference collator (default bf16)")
    ap.add_argument("--gaps", default=None, help="diagnostics: write the device idle gaps of one step (torch.profiler) to this JSON and exit")
    ap.add_argument("--fast", action="store_true", help="skip the T=1025 / preprocessing / CPU legs (development runs)")
    ap.add_argument("--prep", action="store_true",
                    help="configs[1] instead of the training step: Floyd-Warshall sweep (16-512 nodes, batch 256) and "
                         "GraphAttnBias GB/s, one JSON line per case, the CPU reference timed beside each APSP size")
    ap.add_argument("--prep-json", default=None)
    a = ap.parse_args()
    if a.batch is None:
        a.batch = 32 if a.workload == "cfg3" else 16
    return a


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        d = json.load(open(path))
        return dict(hbm_gbs=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sustained=d.get("bf16_tflops_sustained", d["bf16_tflops"]),
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm_gbs=6650.0, tf_burst=1590.0, tf_sustained=1400.0, source="fallback (B200_PROFILING.md)")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return None
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx.append(float(r[1]))
            except Exception:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return None
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the reference's algorithm on the host cores (oracle port), one graph per step
# ---------------------------------------------------------------------------------------------
def cpu_reference_arm(nodes: int, steps: int, warmup: int, image_frac: float, dropouts):
    import torch

    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import mdt_oracle as orc  # the one place bench.py executes oracle/: the timed CPU baseline

    from multimodaldiscussiontransformer_b200.config import PUBLISHED as P
    from multimodaldiscussiontransformer_b200.data import synthetic as syn

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    dp, adp, actdp = dropouts
    model = orc.OracleMDT(num_bottle_neck=P["num_bottleneck_tokens"], num_fusion_layers=P["num_fusion_layers"],
                          num_fusion_stack=P["num_fusion_stack"], num_graph_stack=P["num_graph_stack"],
                          embedding_dim=768, ffn_embedding_dim=768, num_attention_heads=12, dropout=dp,
                          attention_dropout=adp, activation_dropout=actdp)
    orc.seed_parameters(model, 0)
    for m in (model.text_model, model.vit_model):  # --freeze_initial_encoders (encoder.py:223-228)
        for p in m.parameters():
            p.requires_grad = False
    for m in (model.node_classifier, model.text_pooler, model.vit_pooler):
        for p in m.parameters():
            p.requires_grad = True
    model.train()
    opt = torch.optim.AdamW([p for p in model.parameters() if p.requires_grad], lr=P["lr"], betas=P["adam_betas"],
                            eps=P["adam_eps"], weight_decay=P["weight_decay"])
    pool = [syn.synthetic_items([nodes], 1000 + i, image_frac=image_frac) for i in range(2)]

    def step(i):
        batch = orc.batch_to_torch(orc.collator(pool[i % len(pool)], P["spatial_pos_max"]))
        text, bn, _ = model(batch)
        loss, n = orc.node_cross_entropy(model.head(text, bn), batch["y"], batch["y_mask"], P["positive_weight"],
                                         P["negative_weight"])
        opt.zero_grad(set_to_none=True)
        (loss / max(n, 1)).backward()
        opt.step()
        return float(loss)

    for i in range(warmup):
        step(i)
    t0 = time.perf_counter()
    for i in range(steps):
        step(warmup + i)
    dt = time.perf_counter() - t0
    return dict(value=steps / dt, unit=UNIT, cores=cores, kind="port",
                sample=f"{steps} steps of 1 graph x {nodes} comments (same model, fp32, torch CPU, {cores} threads): "
                       f"collate + fwd + bwd + AdamW"), dt / steps * 1e3


def run_prep(args):
    """BASELINE.json configs[1] + GraphAttnBias: benchmarks/prep_microbench.py with the CPU reference beside it."""
    sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import prep_microbench as pm

    ref = None
    try:  # the reference's own Cython floyd_warshall when oracle/_ref was built (it travels with the snapshot)
        import build_ref

        ref = build_ref.load_ref_algos(build=False)
    except Exception:
        ref = None
    import mdt_oracle as orc  # CPU-baseline leg (and the bit-exact check of the timed sample)

    from multimodaldiscussiontransformer_b200.data import algos
    import torch

    def cpu_fn(adj_np):
        n = adj_np.shape[1]
        k = max(1, min(adj_np.shape[0], int(1.5e9 / n ** 3)))  # bounded sample: <= ~1.5 G relaxations per size
        sample = adj_np[:k].astype("int64")
        t0 = time.perf_counter()
        if ref is not None:
            outs = [ref.floyd_warshall(sample[i]) for i in range(k)]
            kind = "reference"
        else:
            outs = [orc.floyd_warshall(sample[i]) for i in range(k)]
            kind = "port"
        dt = time.perf_counter() - t0
        M, path = algos.floyd_warshall_batched(torch.from_numpy(adj_np[:k]).cuda(), True)
        exact = all((M[i].cpu().numpy() == outs[i][0]).all() and (path[i].cpu().numpy() == outs[i][1]).all()
                    for i in range(k))
        return dict(graphs_per_s=k / dt, ms_per_graph=dt / k * 1e3, cores=1, kind=kind, sample=f"{k} graphs of {n} nodes",
                    bit_exact_vs_gpu=bool(exact))

    pm.main(cpu_fn=cpu_fn, argv=["--json", args.prep_json] if args.prep_json else [])


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    dropouts = DROPOUTS   # same as the GPU arm
    cb, ms = cpu_reference_arm(args.nodes, args.steps, args.warmup, args.image_frac, dropouts)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": cb["sample"]}, "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def instep_traffic():
    """DRAM bytes (read + write) of the attention kernels of one forward + one backward call inside a cfg3 training step,
    from the committed `ncu --set full` capture of bench.py (tools/ncu_traffic.py); None when the capture is absent."""
    for name in ("attn_instep_traffic_r02.json", "attn_instep_traffic_r01.json"):
        path = os.path.join(ROOT, "profiles", name)
        if not os.path.isfile(path):
            continue
        try:
            ks = json.load(open(path))["kernels"]
            per_name = {}
            for k in ks:  # one launch of each kernel kind
                per_name.setdefault(k["name"], k["dram_bytes"])
            return float(sum(per_name.values()))
        except Exception:
            continue
    return None


def attention_cfg4(peaks, T=1025, iters=10):
    """The attention kernels alone at SURVEY 8d's cfg4 size (B=32, H=12, T=1025, code + LUT bias, 25 % key padding), timed
    live after the training steps with the same events around the C-ABI call: the cfg3 step only has T=129, where two
    128-wide tiles hold 129 tokens and the kernels are latency-bound."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
    import attn_microbench as mb
    from multimodaldiscussiontransformer_b200 import ops
    B, H = 32, 12
    q, k, v, kw = mb.make_inputs(B, H, T, "code")
    o = ops.biased_attention(q, k, v, **kw)
    go = torch.randn_like(o)
    inputs = [q, k, v, kw["lut"]]

    def fwd():
        with torch.no_grad():
            ops.biased_attention(q, k, v, **kw)

    def bwd():
        torch.autograd.grad(o, inputs, go, retain_graph=True)

    f_ms, b_ms = mb.time_it(fwd, iters, None, "fwd"), mb.time_it(bwd, iters, None, "bwd")
    ff, fb = 4.0 * B * H * T * T * 64, 8.0 * B * H * T * T * 64
    ach = (ff + fb) / ((f_ms + b_ms) * 1e-3) / 1e12
    return {"B": B, "H": H, "T": T, "fwd_us": f_ms * 1e3, "bwd_us": b_ms * 1e3, "fwd_tflops": ff / f_ms / 1e9, "bwd_tflops": fb / b_ms / 1e9,
            "achieved": ach, "peak": peaks["tf_burst"], "frac": ach / peaks["tf_burst"], "unit": "TFLOP/s",
            "peak_kind": "burst bf16 (kernels timed alone)", "iters": iters}


def prep_leg(peaks, iters=5):
    """BASELINE.json configs[1] in the default line (<= ~5 s): the batched Floyd-Warshall kernel at 128 and 512 nodes
    (batch 256, int64 M out) with a bit-exactness check of a sample against the reference's own Cython build
    (oracle/_ref) or the C restatement, and GraphAttnBias forward / backward GB/s at N = 512."""
    import numpy as np
    import torch
    sys.path.insert(0, os.path.join(ROOT, "benchmarks"))
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import prep_microbench as pm
    from multimodaldiscussiontransformer_b200.data import algos

    ref, kind = None, "port"
    try:
        import build_ref
        ref = build_ref.load_ref_algos(build=False)
        kind = "reference"
    except Exception:
        ref = None
    import mdt_oracle as orc  # checker of the timed sample / CPU baseline

    timer = pm.Timer()
    mhz = 1965.0
    try:
        mhz = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("sm_max_mhz", mhz))
    except Exception:
        pass
    alu_peak = pm.SM_COUNT * pm.LANES * 2 * mhz * 1e6
    apsp = []
    for n in (128, 512):
        adj_np, depth = pm.tree_batch(n, 256)
        adj = torch.from_numpy(adj_np).cuda()
        M, _ = algos.floyd_warshall_batched(adj, False)
        pm.check_apsp(M, depth)
        ms = timer(lambda: algos.floyd_warshall_batched(adj, False), iters)
        k = 4 if n == 128 else 1   # bounded CPU sample: n^3 work per graph
        t0 = time.perf_counter()
        outs = [(ref.floyd_warshall(adj_np[i].astype("int64")) if ref is not None else orc.floyd_warshall(adj_np[i].astype("int64")))
                for i in range(k)]
        cpu_ms = (time.perf_counter() - t0) / k * 1e3
        exact = all(bool((M[i].cpu().numpy() == np.asarray(outs[i][0])).all()) for i in range(k))
        relax = 256.0 * n ** 3
        nbytes = 256.0 * n * n * 9
        apsp.append({"n": n, "batch": 256, "us": ms * 1e3, "graphs_per_s": 256 / ms * 1e3, "relaxations_per_s": relax / ms * 1e3,
                     "alu_frac": relax / (ms * 1e-3) / alu_peak, "hbm_gbs": nbytes / ms / 1e6,
                     "bit_exact": exact, "checked_against": kind, "checked_graphs": k, "cpu_ms_per_graph": cpu_ms, "cpu_cores": 1})
    rows = pm.attn_bias_sweep([(32, 512)], iters, timer)
    bias = next(r for r in rows if r["dtype"] == "bfloat16")
    return {
        "apsp": {"bound": "integer ALU: 148 SMs x 128 lanes x 2 relaxations (VIADDMNMX.U16x2) x %.0f MHz = %.1f T relaxations/s" % (mhz, alu_peak / 1e12),
                 "unit": "relaxations/s", "peak": alu_peak, "cases": apsp, "achieved": apsp[-1]["relaxations_per_s"],
                 "frac": apsp[-1]["alu_frac"]},
        "attn_bias": {"bound": "hbm", "unit": "GB/s", "peak": peaks["hbm_gbs"], "B": 32, "N": 512, "dtype": "bf16",
                      "fwd_us": bias["fwd_us"], "bwd_us": bias["bwd_us"], "fwd_gbs": bias["fwd_gbs"], "bwd_gbs": bias["bwd_gbs"],
                      "achieved": bias["fwd_gbs"], "frac": bias["fwd_frac"], "bwd_frac": bias["bwd_frac"]},
    }


# ---------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------
DROPOUTS = (0.4, 0.3, 0.3)  # dropout, attention_dropout, act_dropout: the published run (run_train.sh); attention dropout is in-kernel


def build_model(device):
    import torch

    from multimodaldiscussiontransformer_b200.config import published_args
    from multimodaldiscussiontransformer_b200.criterions import GraphPredictionNodeCrossEntropy
    from multimodaldiscussiontransformer_b200.models import GraphormerModel

    torch.manual_seed(0)
    args = published_args(dropout=DROPOUTS[0], attention_dropout=DROPOUTS[1], act_dropout=DROPOUTS[2])
    model = GraphormerModel.build_model(args, None).to(device=device, dtype=torch.bfloat16)
    model.encoder.graph_encoder.pack_text = os.environ.get("MDT_PACK_TEXT", "1") != "0"   # SURVEY 8f N1 (see config.text)
    crit = GraphPredictionNodeCrossEntropy(None, args.positive_weight, args.negative_weight).to(device)
    return args, model, crit


def workload_sizes(args, rank, world):
    """Comments per graph of this rank's batches (one list per pool entry) and the workload's name."""
    import numpy as np

    from multimodaldiscussiontransformer_b200.train import balanced_shards

    B = args.batch
    if args.workload == "cfg3":
        return [[args.nodes] * B for _ in range(args.pool)], WORKLOAD
    # BASELINE.json configs[4]: trees of up to 512 comments.  Every rank draws the same global list (B x world graphs per
    # pool entry), the shards are balanced by comment count (sum T^2 / sum T), this rank takes its own.
    out = []
    for i in range(args.pool):
        rng = np.random.default_rng(4242 + i)
        sizes = rng.integers(16, 513, size=B * world).tolist()
        out.append([sizes[j] for j in balanced_shards(sizes, world)[rank]])
    return out, ("configs[4]: synthetic trees of 16..512 comments (uniform), %d graphs per GPU per step, padded to the largest "
                 "tree of the batch, shards balanced by comment count, bf16" % B)


def run_b200(args):
    import torch
    import torch.distributed as dist

    from multimodaldiscussiontransformer_b200 import _lib, ops
    from multimodaldiscussiontransformer_b200.data import synthetic as syn
    from multimodaldiscussiontransformer_b200.data.collator import HostBatch, RawBatch, collate_device
    from multimodaldiscussiontransformer_b200.train import PolynomialDecayLR, Trainer

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a B200 (there is no CPU path behind libmdt_b200)"
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    lib = _lib.load()
    assert lib.mdt_device_ok() == 1, _lib.load().mdt_last_error_string()
    torch.backends.cuda.matmul.allow_tf32 = True

    margs, model, crit = build_model(device)
    sched = PolynomialDecayLR(margs.lr, margs.end_learning_rate, margs.warmup_updates, margs.total_num_update)
    grad_dtype = torch.bfloat16 if args.grad_dtype == "bf16" else torch.float32
    trainer = Trainer(model, crit, lr=margs.lr, betas=margs.adam_betas, eps=margs.adam_eps, weight_decay=margs.weight_decay,
                      update_freq=args.update_freq, lr_schedule=sched, grad_dtype=grad_dtype, overlap=not args.no_overlap)
    spmax = margs.spatial_pos_max
    B = args.batch
    uf = args.update_freq
    # graphs are sharded by rank: every rank packs its own pool of batches (weak scaling)
    sizes_pool, workload = workload_sizes(args, rank, world)
    img_dtype = torch.float32 if args.fp32_images else torch.bfloat16
    pool = [HostBatch(syn.synthetic_items(sz, 7919 * (rank + 1) + i, image_frac=args.image_frac), image_dtype=img_dtype)
            for i, sz in enumerate(sizes_pool)]
    resident = [RawBatch(hb, device) for hb in pool]
    h2d = max(hb.nbytes() for hb in pool) * uf
    copy_stream = torch.cuda.Stream(device=device)

    def step_resident(i):
        mbs = [{"net_input": {"batched_data": collate_device(resident[(i * uf + k) % len(resident)], spmax)}} for k in range(uf)]
        return trainer.train_step(mbs)

    inflight = {}

    def fetch(i):  # host -> HBM copies of step i's micro-batches on the copy stream
        return [RawBatch(pool[(i * uf + k) % len(pool)], device, stream=copy_stream) for k in range(uf)]

    readback = {"buf": [torch.empty(Trainer.N_STATS, dtype=torch.float32).pin_memory() for _ in range(2)], "pending": None,
                "last": None}

    def drain_stats():                          # host read of the previous step's loss + counters (its copy has landed)
        if readback["pending"] is not None:
            buf, ev = readback["pending"]
            ev.synchronize()
            readback["last"] = buf.tolist()
            readback["pending"] = None

    def step_e2e(i):
        raws = inflight.pop(i, None) or fetch(i)
        inflight[i + 1] = fetch(i + 1)          # the next step's batch crosses PCIe under this step's kernels
        mbs = [{"net_input": {"batched_data": collate_device(r.wait(), spmax)}} for r in raws]
        stats = trainer.train_step(mbs)
        # loss + counters back to the host every step: the copy is queued behind this step's kernels and read when the NEXT
        # step has been launched (fairseq logs the same way), so the host never stalls the device between steps
        drain_stats()
        buf = readback["buf"][i & 1]
        buf.copy_(stats, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        readback["pending"] = (buf, ev)
        return stats

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, record=False):
        barrier()
        if record:
            ops.ATTN_EVENTS = []
        comm = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            fn(i)
            if record and "comm_end" in trainer.timing:
                comm.append((trainer.timing["backward_end"], trainer.timing["comm_end"]))
        e1.record()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = max(e0.elapsed_time(e1), 0.0)
        ev, ops.ATTN_EVENTS = ops.ATTN_EVENTS, None
        exposed = sum(a_.elapsed_time(b_) for a_, b_ in comm) / max(len(comm), 1)
        t = torch.tensor([ms, wall, exposed], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), ev, float(t[2])

    for i in range(args.warmup):
        step_resident(i)
    if args.gaps:   # diagnostics: where the device idles inside one step (kernel timeline of torch.profiler), then exit
        from torch.profiler import ProfilerActivity, profile
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
            step_resident(args.warmup)
            torch.cuda.synchronize()
        ks = sorted(((e.time_range.start, e.time_range.end, e.name) for e in prof.events()
                     if str(e.device_type).endswith("CUDA") and e.time_range.end > e.time_range.start), key=lambda t: t[0])
        t0, t1 = ks[0][0], max(k[1] for k in ks)
        busy_until, idle, gaps = ks[0][0], 0.0, []
        for a_, b_, n_ in ks:
            if a_ > busy_until:
                idle += a_ - busy_until
                gaps.append((a_ - busy_until, (busy_until - t0) / (t1 - t0), n_[:70]))
            busy_until = max(busy_until, b_)
        nseg = 20
        seg_idle = [0.0] * nseg
        for g_, pos, _ in gaps:
            seg_idle[min(nseg - 1, int(pos * nseg))] += g_
        out = {"step_us": t1 - t0, "idle_us": idle, "idle_frac": idle / (t1 - t0), "kernels": len(ks),
               "idle_us_by_twentieth_of_step": [round(x) for x in seg_idle],
               "largest_gaps": [dict(us=round(g_, 1), at=round(pos, 3), next=n_) for g_, pos, n_ in sorted(gaps, reverse=True)[:40]]}
        with open(args.gaps, "w") as f:
            json.dump(out, f, indent=1)
        print(json.dumps({k: v for k, v in out.items() if k != "largest_gaps"}))
        return
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    calls0 = _lib.kernel_launches()
    ms_total, _, events, exposed_comm_ms = timed(step_resident, args.steps, record=True)
    launches = _lib.kernel_launches() - calls0
    clocks = sampler.stop() if sampler else None
    for i in range(2):
        step_e2e(i)
    inflight.clear()
    drain_stats()
    _, e2e_wall, _, _ = timed(lambda i: (step_e2e(i), drain_stats() if i == args.steps - 1 else None), args.steps)
    assert readback["last"] is not None and math.isfinite(readback["last"][0]), "e2e: the loss never reached the host"
    inflight.clear()
    peak_mem = torch.cuda.max_memory_allocated() / 2 ** 30

    graphs = B * uf * world * args.steps
    value = graphs / (ms_total / 1e3)
    e2e_value = graphs / (e2e_wall / 1e3)

    # ---- communication / copy costs measured alone (they explain the scaling numbers; not part of `value`)
    allreduce_ms = None
    if world > 1:
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record()
        for _ in range(3):
            dist.all_reduce(trainer.flat_grad)
        a1.record()
        barrier()
        allreduce_ms = a0.elapsed_time(a1) / 3
        trainer.flat_grad.zero_()
    c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    c0.record()
    RawBatch(pool[0], device)
    c1.record()
    torch.cuda.synchronize()
    h2d_ms = c0.elapsed_time(c1) * uf

    # ---- roofline of the north-star kernel of this library: biased graph attention (tensor pipe), in-step
    peaks = load_peaks()
    n_graph = max(sizes_pool[0])
    T = n_graph + 1
    fwd = [s.elapsed_time(e) for k, s, e, b_, h_, t_ in events if k == "fwd" and (b_, t_) == (B, T)]
    bwd = [s.elapsed_time(e) for k, s, e, b_, h_, t_ in events if k == "bwd" and (b_, t_) == (B, T)]
    flop_fwd = 4.0 * B * 12 * T * T * 64          # 4*B*H*T^2*d, true T, S-recompute not counted
    flop_bwd = 8.0 * B * 12 * T * T * 64
    roofline = {"bound": "tensor", "kernel": "attn_fwd3_kernel + attn_bwd2_kernel (biased graph attention, code+lut bias, B=%d T=%d)" % (B, T),
                "peak": peaks["tf_sustained"], "unit": "TFLOP/s", "peak_source": peaks["source"] + ", sustained bf16",
                "traffic": instep_traffic() if args.workload == "cfg3" else None}
    all_attn = sum(s.elapsed_time(e) for k, s, e, *_ in events if k in ("fwd", "bwd"))
    if fwd and bwd:
        fwd_ms, bwd_ms = statistics.mean(fwd), statistics.mean(bwd)
        achieved = (flop_fwd + flop_bwd) / ((fwd_ms + bwd_ms) * 1e-3) / 1e12
        roofline.update({"achieved": achieved, "frac": achieved / peaks["tf_sustained"], "fwd_us": fwd_ms * 1e3, "bwd_us": bwd_ms * 1e3,
                         "fwd_tflops": flop_fwd / (fwd_ms * 1e-3) / 1e12, "bwd_tflops": flop_bwd / (bwd_ms * 1e-3) / 1e12,
                         "launches_timed": len(fwd) + len(bwd), "attn_share_of_step": (sum(fwd) + sum(bwd)) / ms_total,
                         "all_mdt_biased_attn_share_of_step": all_attn / ms_total})

    # packed text attention of the fusion layers (HBM-bound: q, k, v in + o out; q, k, v, o, dO in + dq, dk, dv out)
    vl_f = [(s.elapsed_time(e), tot, h_) for k, s, e, tot, h_, nseq in events if k == "vl_fwd"]
    vl_b = [(s.elapsed_time(e), tot, h_) for k, s, e, tot, h_, nseq in events if k == "vl_bwd"]
    if vl_f and vl_b:
        bf = sum(4.0 * tot * h_ * 64 * 2 for _, tot, h_ in vl_f)
        bb = sum(8.0 * tot * h_ * 64 * 2 for _, tot, h_ in vl_b)
        tf_, tb_ = sum(x[0] for x in vl_f), sum(x[0] for x in vl_b)
        roofline["text_attention"] = {"bound": "hbm", "kernel": "attn_varlen_fwd_kernel + attn_varlen_bwd_kernel (packed text sequences)",
                                      "achieved": (bf + bb) / ((tf_ + tb_) * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                      "frac": (bf + bb) / ((tf_ + tb_) * 1e-3) / 1e9 / peaks["hbm_gbs"],
                                      "fwd_us": tf_ / len(vl_f) * 1e3, "bwd_us": tb_ / len(vl_b) * 1e3,
                                      "share_of_step": (tf_ + tb_) / ms_total}

    # the LayerNorm kernels (HBM-bound)
    ln_b = ln_ms = 0.0
    for ev in events:
        k = ev[0]
        if k == "ln_fwd":
            _, s_, e_, rows_, C_, flags = ev
            ln_b += rows_ * C_ * 2.0 * (2 + (flags >> 1) + (flags & 1)) + rows_ * 8.0   # x, y (+ residual) (+ z) + mean / rstd
        elif k == "ln_bwd":
            _, s_, e_, rows_, C_, flags = ev
            ln_b += rows_ * C_ * 2.0 * (2 + flags) + rows_ * 8.0                         # dy, z, gradients written
        else:
            continue
        ln_ms += s_.elapsed_time(e_)
    if ln_ms > 0:
        roofline["layernorm"] = {"bound": "hbm", "kernel": "ln_fwd_kernel + ln_bwd_kernel (dropout + residual + LayerNorm)",
                                 "achieved": ln_b / (ln_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                                 "frac": ln_b / (ln_ms * 1e-3) / 1e9 / peaks["hbm_gbs"], "share_of_step": ln_ms / ms_total}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    if world == 1 and not args.fast:
        roofline["attn_T1025"] = attention_cfg4(peaks)
        try:
            roofline.update(prep_leg(peaks))
        except Exception as ex:  # the prep leg must never cost the headline line
            roofline["prep_error"] = "%s: %s" % (type(ex).__name__, str(ex)[:200])
    cpu = None
    if not args.no_cpu_baseline and world == 1 and not args.fast:
        cpu, _ = cpu_reference_arm(args.nodes, 3, 1, args.image_frac, DROPOUTS)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": workload, "graphs_per_gpu": B, "update_freq": uf,
                   "comments_per_graph": args.nodes if args.workload == "cfg3" else "16..512 (padded to %d)" % n_graph,
                   "image_frac": args.image_frac,
                   "model": "published run_train.sh 8 4 5 2 2: 8 fusion layers (stacks of 2), 10 executed graph layers, nb=4, "
                            "spatial_pos_max=5, freeze_initial_encoders, AdamW fp32 master weights",
                   "dropout": {"dropout": DROPOUTS[0], "attention_dropout": DROPOUTS[1], "act_dropout": DROPOUTS[2]},
                   "text": "fusion-layer BERT blocks on packed (unpadded) tokens, attention on this library's packed-sequence kernels "
                           "(MDT_TEXT_ATTN=flash / MDT_PACK_TEXT=0 restore the library call / the padded path for A/B runs)",
                   "images": "shipped %s over PCIe (the model computes in bf16)" % ("fp32" if args.fp32_images else "bf16"),
                   "comm": "gradient buckets (%s) all-reduced on a side stream while the backward runs" % args.grad_dtype
                           if not args.no_overlap else "one all-reduce after the backward",
                   "l2": "working set per step (activations, weights) >> 126 MB L2; batches cycle through a pool",
                   "parallelism": f"dp{world}", "peak_mem_gib": round(peak_mem, 1)},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": Trainer.N_STATS * 4,
                "note": "step i+1's batch is copied from pinned memory on a copy stream while step i computes; step i's loss + counters are copied to pinned host memory behind its kernels and read on the host once step i+1 is launched (every step, the last one before the timer stops)"},
        "gpu_launches": launches, "roofline": roofline, "clocks": clocks,
        "h2d_ms": h2d_ms, "allreduce_ms": allreduce_ms, "exposed_comm_ms": exposed_comm_ms if world > 1 else 0.0,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    elif a.prep:
        run_prep(a)
    else:
        run_b200(a)
