"""``node_cross_entropy`` criterion (mDT/src/criterions/hatespeech_loss.py:23-182): class-weighted
cross entropy, summed over the labelled comments, plus the counters behind accuracy / precision /
recall / F1.  The reference's device->host->device round trip of the logits (:95) is gone, and so is
the boolean-mask gather: the forward does not synchronise with the host."""

from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List

import torch
import torch.nn.functional as F

from .._registry import FairseqCriterion, register_criterion


@dataclass
class GraphPredictionNodeCrossEntropyConfig:
    positive_weight: float = field(default=1.0, metadata={"help": "Weight to associate to the positive class"})
    negative_weight: float = field(default=1.0, metadata={"help": "Weight to associate to the negative class"})


@register_criterion("node_cross_entropy", dataclass=GraphPredictionNodeCrossEntropyConfig)
class GraphPredictionNodeCrossEntropy(FairseqCriterion):
    def __init__(self, task=None, positive_weight: float = 1.0, negative_weight: float = 1.0) -> None:
        super().__init__(task)
        self.register_buffer("weight", torch.tensor([negative_weight, positive_weight], dtype=torch.float32),
                             persistent=False)

    def forward(self, model, sample: Dict[str, Any], reduce=True):
        """-> ``(loss, sample_size, logging_output)``; sample_size = number of labelled comments (:97)."""
        bd = sample["net_input"]["batched_data"]
        num_comments = bd["x"].shape[1]
        comment_logits, _ = model(**sample["net_input"])
        # The reference indexes ``comment_logits[y_mask]`` (:95-99), a boolean-mask gather that costs a device->host
        # sync.  Same numbers without it: every comment gets the label of its rank among the labelled ones, the
        # per-comment losses / counters are masked and summed.  ``sample_size`` = number of labels = ``len(logits)`` (:97).
        mask = bd["y_mask"].reshape(-1).bool()
        targets = torch.flatten(bd["y"]).long()
        sample_size = int(targets.numel())
        logits = comment_logits.float()
        if sample_size == 0:
            tgt_all = torch.zeros(mask.shape[0], dtype=torch.long, device=mask.device)
        else:
            rank = (torch.cumsum(mask.to(torch.int32), 0) - 1).clamp_(0, sample_size - 1)
            tgt_all = targets[rank.long()]
        weight = self.weight.to(logits.device)
        with torch.no_grad():
            pred = torch.argmax(logits, dim=-1)
            hit = (pred == tgt_all) & mask
            ncorrect = hit.sum()
            num_positive_correct = (hit & (pred == 1)).sum()
            total_positive = ((tgt_all == 1) & mask).sum()
            num_pred_positive = ((pred == 1) & mask).sum()
        per_comment = F.cross_entropy(logits, tgt_all, reduction="none", weight=weight) * mask.to(logits.dtype)
        loss = per_comment.sum() if reduce else per_comment
        logging_output = {"loss": loss.data, "sample_size": sample_size, "nsentences": sample_size, "ntokens": num_comments,
                          "ncorrect": ncorrect, "num_positive_correct": num_positive_correct,
                          "total_positive": total_positive, "num_pred_positive": num_pred_positive}
        return loss, sample_size, logging_output

    @staticmethod
    def aggregate(logging_outputs: List[Dict[str, Any]]) -> Dict[str, float]:
        """The numbers ``reduce_metrics`` logs (:134-173), returned as a dict."""
        tot = lambda k: float(sum(float(log.get(k, 0)) for log in logging_outputs))
        n = tot("sample_size")
        out = {"loss": tot("loss") / max(n, 1.0), "sample_size": n}
        if logging_outputs and "ncorrect" in logging_outputs[0]:
            pc, tp, pp = tot("num_positive_correct"), tot("total_positive"), tot("num_pred_positive")
            recall = pc / tp if tp else 0.0
            precision = pc / pp if pp else 0.0
            f1 = 0.0 if precision == 0 and recall == 0 else 2 * precision * recall / (precision + recall)
            out.update(accuracy=tot("ncorrect") / max(n, 1.0), recall=recall, precision=precision, f1=f1)
        return out

    @staticmethod
    def reduce_metrics(logging_outputs: List[Dict[str, Any]]) -> None:
        agg = GraphPredictionNodeCrossEntropy.aggregate(logging_outputs)
        try:
            from fairseq import metrics
        except Exception:
            return
        for k in ("loss", "accuracy", "recall", "precision", "f1"):
            if k in agg:
                metrics.log_scalar(k, agg[k], agg["sample_size"], round=3)

    @staticmethod
    def logging_outputs_can_be_summed() -> bool:
        return True
