"""Host-side logic of the data-parallel driver on CPU: world_size 2 over gloo.

The model here is a tiny CPU stand-in with the criterion's interface; the point is the flat
gradient buffer, the zero-fill of dead parameters, the sample_size denominator and that two ranks
with different shards end up with identical weights equal to a single-process run over both shards."""

from __future__ import annotations

import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


class Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        torch.manual_seed(3)
        self.lin = torch.nn.Linear(6, 2)
        self.dead = torch.nn.Linear(4, 4)  # never used: grad stays None

    def forward(self, batched_data):
        return self.lin(batched_data["feat"]), None


def make_sample(seed, n):
    g = torch.Generator().manual_seed(seed)
    feat = torch.randn(n, 6, generator=g)
    y = torch.randint(0, 2, (n,), generator=g).float()
    mask = torch.ones(n, dtype=torch.bool)
    return {"net_input": {"batched_data": {"feat": feat, "x": torch.zeros(1, n, 1), "y": y, "y_mask": mask}}}


def run_trainer(samples_per_step, steps, update_freq=1):
    from multimodaldiscussiontransformer_b200.criterions import GraphPredictionNodeCrossEntropy
    from multimodaldiscussiontransformer_b200.train import Trainer

    model = Toy()
    crit = GraphPredictionNodeCrossEntropy(None, 1.5, 1.0)
    tr = Trainer(model, crit, lr=1e-2, weight_decay=0.0, update_freq=update_freq)
    stats = None
    for s in range(steps):
        stats = tr.train_step(samples_per_step(s))
    return model, stats


def worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # rank r sees shard r: different sizes so the sample_size denominator matters
    model, stats = run_trainer(lambda s: [make_sample(100 * s + rank, 3 + 2 * rank)], steps=3)
    torch.save({"w": model.lin.weight.detach(), "dead": model.dead.weight.detach(), "stats": stats}, f"{out}.{rank}")
    dist.destroy_process_group()


def test_two_ranks_match_single_process(tmp_path):
    port = 29500 + os.getpid() % 2000
    out = str(tmp_path / "res")
    mp.spawn(worker, args=(2, port, out), nprocs=2, join=True)
    r0, r1 = torch.load(out + ".0"), torch.load(out + ".1")
    assert torch.equal(r0["w"], r1["w"]) and torch.equal(r0["dead"], r1["dead"])
    assert torch.equal(r0["stats"], r1["stats"])
    assert float(r0["stats"][1]) == 3 + 5  # global sample_size of the last step
    # one process accumulating both shards per step (update_freq 2) is the same optimisation problem
    single, _ = run_trainer(lambda s: [make_sample(100 * s + 0, 3), make_sample(100 * s + 1, 5)], steps=3, update_freq=2)
    assert torch.allclose(single.lin.weight, r0["w"], atol=1e-6)


def test_shard_indices_cover_everything():
    from multimodaldiscussiontransformer_b200.train import shard_indices

    shards = [shard_indices(103, r, 4, epoch=2) for r in range(4)]
    assert len({len(s) for s in shards}) == 1
    assert set().union(*map(set, shards)) == set(range(103))
    assert shard_indices(103, 1, 4, epoch=2) == shards[1] and shard_indices(103, 1, 4, epoch=3) != shards[1]


def test_polynomial_decay():
    from multimodaldiscussiontransformer_b200.train import PolynomialDecayLR

    s = PolynomialDecayLR(3e-5, 3e-7, 3246, 10820)
    assert s(0) == 0.0 and abs(s(3246) - 3e-5) < 1e-12 and s(10820) == 3e-7 and s(20000) == 3e-7
    assert 3e-7 < s(7000) < 3e-5
