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


class Trainer:
    """Flat fp32 master weights / gradients over the trainable parameters + fused AdamW."""

    N_STATS = 8  # loss, sample_size, ncorrect, num_positive_correct, total_positive, num_pred_positive, ntokens, nbatches

    def __init__(self, model: torch.nn.Module, criterion, lr: float = 3e-5, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.01, update_freq: int = 1, clip_norm: float = 0.0,
                 lr_schedule: Optional[PolynomialDecayLR] = None, process_group=None, grad_dtype=torch.float32):
        self.model, self.criterion = model, criterion
        self.update_freq, self.clip_norm, self.lr_schedule = update_freq, clip_norm, lr_schedule
        self.group = process_group
        self.world = dist.get_world_size(process_group) if dist.is_available() and dist.is_initialized() else 1
        self.params = [p for p in model.parameters() if p.requires_grad]
        dev = self.params[0].device
        n = sum(p.numel() for p in self.params)
        self.numel = n
        # [grads | stats]: one buffer, one collective
        self.flat_grad = torch.zeros(n + self.N_STATS, dtype=grad_dtype, device=dev)
        self.master = torch.empty(n, dtype=torch.float32, device=dev)
        self.master_grad = self.flat_grad[:n] if grad_dtype == torch.float32 else torch.zeros(n, dtype=torch.float32, device=dev)
        o = 0
        self.offsets = []
        for p in self.params:
            self.master[o:o + p.numel()].copy_(p.detach().reshape(-1))
            self.offsets.append(o)
            o += p.numel()
        self.master.requires_grad_(False)
        self.master_param = torch.nn.Parameter(self.master, requires_grad=True)
        self.optimizer = torch.optim.AdamW([self.master_param], lr=lr, betas=betas, eps=eps, weight_decay=weight_decay,
                                           fused=dev.type == "cuda")
        self.num_updates = 0

    # ------------------------------------------------------------------------------------------
    def _flatten_grads(self):
        fg = self.flat_grad
        views = []
        for p, o in zip(self.params, self.offsets):
            dst = fg[o:o + p.numel()].view(p.shape)
            if p.grad is None:
                dst.zero_()      # dead parameters: reduce zeros like fairseq's legacy DDP
            else:
                views.append((dst, p.grad))
        if views:
            torch._foreach_copy_([d for d, _ in views], [g for _, g in views])

    def _write_back(self):
        srcs = [self.master[o:o + p.numel()].view(p.shape) for p, o in zip(self.params, self.offsets)]
        torch._foreach_copy_([p.data for p in self.params], srcs)

    def train_step(self, samples: Iterable[Dict[str, Any]]) -> torch.Tensor:
        """One optimizer update over ``update_freq`` micro-batches.  Returns the reduced stats tensor
        (device, fp32, ``N_STATS`` entries, summed over ranks) without synchronising."""
        self.model.train()
        stats = torch.zeros(self.N_STATS, dtype=torch.float32, device=self.flat_grad.device)
        for sample in samples:
            loss, sample_size, log = self.criterion(self.model, sample)
            loss.backward()
            vals = [log["loss"], sample_size, log["ncorrect"], log["num_positive_correct"], log["total_positive"],
                    log["num_pred_positive"], log["ntokens"], 1]
            stats += torch.stack([torch.as_tensor(v, dtype=torch.float32, device=stats.device).reshape(()) for v in vals])
        self._flatten_grads()
        self.flat_grad[self.numel:].copy_(stats)
        if self.world > 1:
            dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
        stats = self.flat_grad[self.numel:].float().clone()
        if self.master_grad.data_ptr() != self.flat_grad.data_ptr():
            self.master_grad.copy_(self.flat_grad[: self.numel])
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
