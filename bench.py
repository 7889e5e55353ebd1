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
    ap.add_argument("--batch", type=int, default=32, help="discussion graphs per GPU per step")
    ap.add_argument("--nodes", type=int, default=128, help="comments per graph")
    ap.add_argument("--image-frac", type=float, default=0.10)
    ap.add_argument("--pool", type=int, default=2, help="distinct pre-packed host batches cycled through")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--prep", action="store_true",
                    help="configs[1] instead of the training step: Floyd-Warshall sweep (16-512 nodes, batch 256) and "
                         "GraphAttnBias GB/s, one JSON line per case, the CPU reference timed beside each APSP size")
    ap.add_argument("--prep-json", default=None)
    return ap.parse_args()


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
    dropouts = (0.4, 0.0, 0.3)
    cb, ms = cpu_reference_arm(args.nodes, args.steps, args.warmup, args.image_frac, dropouts)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "sample": cb["sample"]}, "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def instep_traffic():
    """DRAM bytes (read + write) of the attention kernels of one forward + one backward call inside a cfg3 training step,
    from the committed `ncu --set full` capture of bench.py (profiles/attn_instep_traffic_r01.json, tools/ncu_traffic.py);
    None when the capture is absent."""
    path = os.path.join(ROOT, "profiles", "attn_instep_traffic_r01.json")
    if not os.path.isfile(path):
        return None
    try:
        ks = json.load(open(path))["kernels"]
        per_name = {}
        for k in ks:  # one launch of each kernel kind
            per_name.setdefault(k["name"], k["dram_bytes"])
        return float(sum(per_name.values()))
    except Exception:
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
    trainer = Trainer(model, crit, lr=margs.lr, betas=margs.adam_betas, eps=margs.adam_eps, weight_decay=margs.weight_decay,
                      update_freq=1, lr_schedule=sched)
    spmax = margs.spatial_pos_max
    B, n = args.batch, args.nodes
    # graphs are sharded by rank: every rank packs its own pool of batches (weak scaling)
    pool = [HostBatch(syn.synthetic_items([n] * B, 7919 * (rank + 1) + i, image_frac=args.image_frac)) for i in range(args.pool)]
    resident = [RawBatch(hb, device) for hb in pool]
    h2d = max(hb.nbytes() for hb in pool)

    def step_resident(i):
        bd = collate_device(resident[i % len(resident)], spmax)
        return trainer.train_step([{"net_input": {"batched_data": bd}}])

    def step_e2e(i):
        bd = collate_device(RawBatch(pool[i % len(pool)], device), spmax)   # pinned host -> HBM every step
        stats = trainer.train_step([{"net_input": {"batched_data": bd}}])
        return stats.cpu()                                                   # loss + counters back to the host

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, record=False):
        barrier()
        if record:
            ops.ATTN_EVENTS = []
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for i in range(steps):
            fn(i)
        e1.record()
        barrier()
        wall = (time.perf_counter() - t0) * 1e3
        ms = max(e0.elapsed_time(e1), 0.0)
        ev, ops.ATTN_EVENTS = ops.ATTN_EVENTS, None
        t = torch.tensor([ms, wall], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0]), float(t[1]), ev

    for i in range(args.warmup):
        step_resident(i)
    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    calls0 = _lib.kernel_launches()
    ms_total, _, events = timed(step_resident, args.steps, record=True)
    launches = _lib.kernel_launches() - calls0
    clocks = sampler.stop() if sampler else None
    for i in range(2):
        step_e2e(i)
    _, e2e_wall, _ = timed(step_e2e, args.steps)
    peak_mem = torch.cuda.max_memory_allocated() / 2 ** 30

    graphs = B * world * args.steps
    value = graphs / (ms_total / 1e3)
    e2e_value = graphs / (e2e_wall / 1e3)

    # ---- roofline of the dominant kernel of this library: biased attention (tensor pipe)
    peaks = load_peaks()
    fwd = [s.elapsed_time(e) for k, s, e, *_ in events if k == "fwd"]
    bwd = [s.elapsed_time(e) for k, s, e, *_ in events if k == "bwd"]
    T = n + 1
    flop_fwd = 4.0 * B * 12 * T * T * 64          # 4*B*H*T^2*d, true T, S-recompute not counted
    flop_bwd = 8.0 * B * 12 * T * T * 64
    fwd_ms, bwd_ms = statistics.mean(fwd), statistics.mean(bwd)
    achieved = (flop_fwd + flop_bwd) / ((fwd_ms + bwd_ms) * 1e-3) / 1e12
    roofline = {"bound": "tensor", "kernel": "attn_fwd_kernel + attn_bwd_kernel (biased attention, code+lut bias)",
                "achieved": achieved, "peak": peaks["tf_sustained"], "unit": "TFLOP/s", "frac": achieved / peaks["tf_sustained"],
                "traffic": instep_traffic(), "peak_source": peaks["source"] + ", sustained bf16", "fwd_us": fwd_ms * 1e3, "bwd_us": bwd_ms * 1e3,
                "fwd_tflops": flop_fwd / (fwd_ms * 1e-3) / 1e12, "bwd_tflops": flop_bwd / (bwd_ms * 1e-3) / 1e12,
                "launches_timed": len(fwd) + len(bwd), "attn_share_of_step": (sum(fwd) + sum(bwd)) / ms_total}

    # the LayerNorm kernels (HBM-bound) are the largest share of the step among this library's kernels
    ln_b = ln_ms = 0.0
    for k, s_, e_, rows_, C_, flags in events:
        if k == "ln_fwd":
            ln_b += rows_ * C_ * 2.0 * (2 + (flags >> 1) + (flags & 1)) + rows_ * 8.0   # x, y (+ residual) (+ z) + mean / rstd
        elif k == "ln_bwd":
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
    if world == 1:
        roofline["attn_T1025"] = attention_cfg4(peaks)
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        cpu, _ = cpu_reference_arm(n, 2, 1, args.image_frac, DROPOUTS)
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "bf16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "graphs_per_gpu": B, "comments_per_graph": n, "image_frac": args.image_frac,
                   "model": "published run_train.sh 8 4 5 2 2: 8 fusion layers (stacks of 2), 10 executed graph layers, nb=4, "
                            "spatial_pos_max=5, freeze_initial_encoders, AdamW fp32 master weights, update_freq 1",
                   "dropout": {"dropout": DROPOUTS[0], "attention_dropout": DROPOUTS[1], "act_dropout": DROPOUTS[2]},
                   "text": "fusion-layer BERT blocks on packed (unpadded) tokens: padding tokens (46 % of the batch) are not computed; "
                           "logits / loss / gradients equal the padded path (tests/test_gpu_modules.py), MDT_PACK_TEXT=0 restores it",
                   "note": "graph-layer attention dropout 0.3 runs inside the attention kernels (counter-hash mask, regenerated in the backward)",
                   "l2": "working set per step (activations, weights) >> 126 MB L2; batches cycle through a pool",
                   "parallelism": f"dp{world}", "peak_mem_gib": round(peak_mem, 1)},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": Trainer.N_STATS * 4},
        "gpu_launches": launches, "roofline": roofline, "clocks": clocks,
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
