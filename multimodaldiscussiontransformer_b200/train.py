"""Data-parallel training driver for mDT on B200s: one process per GPU, discussion graphs sharded by
rank, bf16 model with fp32 master weights, gradients reduced over NCCL (NVLink 5 / NVSwitch).

What it stands in for: the fairseq trainer the reference is launched through
(``fairseq-train --fp16 --update-freq 3 --optimizer adam ... --lr-scheduler polynomial_decay``,
mDT/experiments/hateful_discussions/run_train.sh:28-65).  Semantics kept:
  * the criterion returns a *summed* loss and ``sample_size`` = labelled comments (hatespeech_loss.py:97,113-118);
    gradients are divided by the global ``sample_size`` after the all-reduce;
  * ``update_freq`` micro-batches accumulate before one reduce + optimizer step;
  * parameters that never receive a gradient (the graph stack the forward skips, ``bert_projection`` ...,
    SURVEY.md 0.4) are zero-filled in the flat gradient buffer, so every rank reduces the same bytes;
  * logging counters travel in the same flat buffer (one collective per update, not two).
"""

from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional

import torch
import torch.distributed as dist


class PolynomialDecayLR:
    """``--lr-scheduler polynomial_decay --power p --warmup-updates W --total-num-update N`` (run_train.sh:38-40)."""

    def __init__(self, lr: float, end_lr: float, warmup_updates: int, total_updates: int, power: float = 1.0):
        self.lr, self.end_lr, self.warmup, self.total, self.power = lr, end_lr, warmup_updates, total_updates, power

    def __call__(self, num_updates: int) -> float:
        if self.warmup > 0 and num_updates <= self.warmup:
            return self.lr * num_updates / float(self.warmup)
        if num_updates >= self.total:
            return self.end_lr
        pct = 1.0 - (num_updates - self.warmup) / float(self.total - self.warmup)
        return (self.lr - self.end_lr) * pct ** self.power + self.end_lr


def shard_indices(n_items: int, rank: int, world: int, epoch: int, seed: int = 1) -> List[int]:
    """Epoch-seeded permutation (EpochShuffleDataset, dataset.py:217-233) split round-robin across ranks;
    the tail is padded by wrapping so every rank runs the same number of steps."""
    g = torch.Generator().manual_seed(seed + epoch)
    perm = torch.randperm(n_items, generator=g).tolist()
    per = (n_items + world - 1) // world
    perm = perm + perm[: per * world - n_items]
    return perm[rank::world]


def balanced_shards(sizes: List[int], world: int) -> List[List[int]]:
    """Deal graphs to ranks so that every rank gets the same number of graphs and a similar sum of T^2 (attention cost)
    and of T (token count): indices sorted by size, dealt in a serpentine (0..w-1, w-1..0, ...).  SURVEY.md 8e: the
    straggler control for batches whose trees range from 16 to 512 comments (BASELINE configs[4])."""
    order = sorted(range(len(sizes)), key=lambda i: -sizes[i])
    shards: List[List[int]] = [[] for _ in range(world)]
    for pos, i in enumerate(order):
        lap, k = divmod(pos, world)
        shards[k if lap % 2 == 0 else world - 1 - k].append(i)
    return shards


class Trainer:
    """Flat master weights / gradients over the trainable parameters + fused AdamW; gradients of different ranks are
    summed bucket by bucket WHILE the backward is still running.

    Layout: the flat buffers hold the parameters in REVERSE registration order, so the parameters whose gradients are
    produced first (the head, the last fusion / graph stacks) come first and consecutive parameters form buckets of about
    ``bucket_mb``.  A ``post_accumulate_grad`` hook counts a bucket's gradients down; when the last one lands, the bucket is
    packed into the flat buffer (``grad_dtype``: bf16 halves the bytes on NVLink) and its NCCL all-reduce is launched on a
    side stream.  Parameters that never receive a gradient (the graph stack the forward skips, ``bert_projection`` ...,
    SURVEY.md 0.4) are found in the first step, which runs without overlap; afterwards they are neither waited for nor
    copied -- their slots stay zero, so every rank still reduces identical bytes (fairseq ``legacy_ddp`` semantics).
    The logging counters travel in their own fp32 tensor (exact whatever ``grad_dtype`` is).
    """

    N_STATS = 8  # loss, sample_size, ncorrect, num_positive_correct, total_positive, num_pred_positive, ntokens, nbatches

    def __init__(self, model: torch.nn.Module, criterion, lr: float = 3e-5, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, update_freq: int = 1, clip_norm: float = 0.0,
                 lr_schedule: Optional[PolynomialDecayLR] = None, process_group=None, grad_dtype=torch.float32,
                 bucket_mb: float = 64.0, overlap: bool = True):
        self.model, self.criterion = model, criterion
        self.update_freq, self.clip_norm, self.lr_schedule = update_freq, clip_norm, lr_schedule
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.params = [p for p in model.parameters() if p.requires_grad][::-1]   # gradient-ready order (approximately)
        dev = self.params[0].device
        n = sum(p.numel() for p in self.params)
        self.numel = n
        self.flat_grad = torch.zeros(n, dtype=grad_dtype, device=dev)
        self.stats_buf = torch.zeros(self.N_STATS, dtype=torch.float32, device=dev)
        self.master = torch.empty(n, dtype=torch.float32, device=dev)
        self.master_grad = self.flat_grad if grad_dtype == torch.float32 else torch.zeros(n, dtype=torch.float32, device=dev)
        o = 0
        self.offsets = []
        for p in self.params:
            self.master[o:o + p.numel()].copy_(p.detach().reshape(-1))
            self.offsets.append(o)
            o += p.numel()
        self.master.requires_grad_(False)
        # per-parameter views of the flat buffers, built once: re-slicing ~400 parameters per bucket pack and per write-back was
        # ~5 ms of host time per update during which the device idled (profiles/step_gaps_r02.json)
        self._grad_views = [self.flat_grad[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, self.offsets)]
        self._master_views = [self.master[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, self.offsets)]
        self.master_param = torch.nn.Parameter(self.master, requires_grad=True)
        self.optimizer = torch.optim.AdamW([self.master_param], lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                           fused=dev.type == "cuda")
        self.num_updates = 0
        # ---- buckets over consecutive parameters
        limit = int(bucket_mb * 2 ** 20 / self.flat_grad.element_size())
        self.buckets: List[List[int]] = [[]]
        size = 0
        for i, p in enumerate(self.params):
            if self.buckets[-1] and size + p.numel() > limit:
                self.buckets.append([])
                size = 0
            self.buckets[-1].append(i)
            size += p.numel()
        self.bucket_of = [b for b, idxs in enumerate(self.buckets) for _ in idxs]
        self.overlap = bool(overlap) and self.world > 1 and dev.type == "cuda"
        self._dead: Optional[set] = None       # parameters without gradient, known after the first step
        self._armed = False                    # hooks act only during the last micro-batch of an update
        self._pending: List[int] = []
        self._launched: List[bool] = []
        self._works: list = []
        self._comm_stream = torch.cuda.Stream(device=dev) if self.overlap else None
        self.timing: Dict[str, Any] = {}       # CUDA events of the last update: backward end / communication end
        if self.overlap:
            for i, p in enumerate(self.params):
                p.register_post_accumulate_grad_hook(self._make_hook(i))

    # ------------------------------------------------------------------------------------------
    def _bucket_slice(self, b: int) -> torch.Tensor:
        idxs = self.buckets[b]
        lo = self.offsets[idxs[0]]
        hi = self.offsets[idxs[-1]] + self.params[idxs[-1]].numel()
        return self.flat_grad[lo:hi]

    def _pack_bucket(self, b: int) -> None:
        dsts, srcs = [], []
        for i in self.buckets[b]:
            p = self.params[i]
            dst = self._grad_views[i]
            if p.grad is None:
                if self._dead is None or i not in self._dead:
                    dst.zero_()      # dead parameters: reduce zeros like fairseq's legacy DDP
            else:
                dsts.append(dst)
                srcs.append(p.grad)
        if dsts:
            torch._foreach_copy_(dsts, srcs)

    def _launch_bucket(self, b: int) -> None:
        """Pack bucket b on the compute stream and start its all-reduce on the side stream."""
        self._pack_bucket(b)
        self._launched[b] = True
        ready = torch.cuda.Event()
        ready.record()
        with torch.cuda.stream(self._comm_stream):
            self._comm_stream.wait_event(ready)
            self._works.append(dist.all_reduce(self._bucket_slice(b), op=dist.ReduceOp.SUM, group=self.group, async_op=True))

    def _make_hook(self, i: int):
        def hook(_param):
            if not self._armed:
                return
            b = self.bucket_of[i]
            self._pending[b] -= 1
            if self._pending[b] == 0 and not self._launched[b]:
                self._launch_bucket(b)
        return hook

    def _write_back(self):
        torch._foreach_copy_([p.data for p in self.params], self._master_views)

    def train_step(self, samples: Iterable[Dict[str, Any]]) -> torch.Tensor:
        """One optimizer update over ``update_freq`` micro-batches.  Returns the reduced stats tensor
        (device, fp32, ``N_STATS`` entries, summed over ranks) without synchronising."""
        if not self.model.training:      # walking ~1000 modules per update is 1 ms of host time
            self.model.train()
        samples = list(samples)
        stats = torch.zeros(self.N_STATS, dtype=torch.float32, device=self.flat_grad.device)
        overlap_now = self.overlap and self._dead is not None
        for k, sample in enumerate(samples):
            loss, sample_size, log = self.criterion(self.model, sample)
            if overlap_now and k == len(samples) - 1:   # gradients are final after this backward: reduce them as they land
                self._pending = [sum(1 for i in idxs if i not in self._dead) for idxs in self.buckets]
                self._launched = [False] * len(self.buckets)
                self._works = []
                self._armed = True
            loss.backward()
            self._armed = False
            vals = [log["loss"], sample_size, log["ncorrect"], log["num_positive_correct"], log["total_positive"],
                    log["num_pred_positive"], log["ntokens"], 1]
            stats += torch.stack([torch.as_tensor(v, dtype=torch.float32, device=stats.device).reshape(()) for v in vals])
        cuda = self.flat_grad.is_cuda
        if cuda:
            self.timing["backward_end"] = torch.cuda.Event(enable_timing=True)
            self.timing["backward_end"].record()
        self.stats_buf.copy_(stats)
        if self.world > 1:
            if overlap_now:
                for b in range(len(self.buckets)):      # buckets a hook did not complete (nothing should be left but all-dead ones)
                    if not self._launched[b]:
                        self._launch_bucket(b)
                self._works.append(dist.all_reduce(self.stats_buf, op=dist.ReduceOp.SUM, group=self.group, async_op=True))
                for w in self._works:
                    w.wait()                            # the compute stream waits; the host does not
                self._works = []
            else:
                for b in range(len(self.buckets)):
                    self._pack_bucket(b)
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
                dist.all_reduce(self.stats_buf, op=dist.ReduceOp.SUM, group=self.group)
        else:
            for b in range(len(self.buckets)):
                self._pack_bucket(b)
        if cuda:
            self.timing["comm_end"] = torch.cuda.Event(enable_timing=True)
            self.timing["comm_end"].record()
        if self._dead is None:   # first update: remember which parameters never get a gradient; their slots are zero for good
            self._dead = {i for i, p in enumerate(self.params) if p.grad is None}
        stats = self.stats_buf.clone()
        if self.master_grad.data_ptr() != self.flat_grad.data_ptr():
            self.master_grad.copy_(self.flat_grad)
        # gradient denominator = global number of labelled comments (fairseq: sample_size)
        self.master_grad.div_(stats[1].clamp_min(1.0))
        if self.clip_norm > 0:
            norm = torch.linalg.vector_norm(self.master_grad)
            self.master_grad.mul_((self.clip_norm / (norm + 1e-6)).clamp_max(1.0))
        self.num_updates += 1
        if self.lr_schedule is not None:
            for g in self.optimizer.param_groups:
                g["lr"] = self.lr_schedule(self.num_updates)
        self.master_param.grad = self.master_grad
        self.optimizer.step()
        self._write_back()
        for p in self.params:
            p.grad = None
        return stats

    @staticmethod
    def stats_to_metrics(stats: torch.Tensor) -> Dict[str, float]:
        s = stats.tolist()
        from .criterions.hatespeech_loss import GraphPredictionNodeCrossEntropy as C

        return C.aggregate([dict(loss=s[0], sample_size=s[1], ncorrect=s[2], num_positive_correct=s[3], total_positive=s[4],
                                 num_pred_positive=s[5])])
