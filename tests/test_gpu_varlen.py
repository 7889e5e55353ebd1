"""GPU parity of the packed-sequence attention kernels (the BERT side of the fusion layers,
multi_graphormer_fusion_layer.py:105-149) against the fp32 formula ``softmax(q k^T * scale) v`` evaluated per
sequence on the same bf16 inputs.  Tolerances as for the graph attention: forward <= 2e-2 absolute on N(0,1)
inputs, gradients <= 3e-2 relative Frobenius."""

from __future__ import annotations

import math

import pytest
import torch

pytestmark = pytest.mark.gpu

FWD_ATOL = 2e-2
GRAD_RTOL = 3e-2
_M32 = 0xFFFFFFFF


@pytest.fixture(scope="module")
def ops():
    from multimodaldiscussiontransformer_b200 import ops as _ops

    return _ops


def make(lens, H, seed):
    g = torch.Generator().manual_seed(seed)
    total = sum(lens)
    qkv = torch.randn(total, 3, H, 64, generator=g).bfloat16().cuda()
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32).cuda()
    return qkv, cu


def reference(qkv, lens, scale, keep=None, p=0.0):
    """fp32, sequence by sequence; ``keep`` [H, total, 128] (tile-local columns handled by the caller) or None."""
    q, k, v = (qkv[:, i].float() for i in range(3))
    outs, o = [], 0
    for n in lens:
        s = torch.einsum("thd,shd->hts", q[o:o + n], k[o:o + n]) * scale
        pr = torch.softmax(s, dim=-1)
        if keep is not None:
            pr = pr * keep[o:o + n].float() / (1.0 - p)
        outs.append(torch.einsum("hts,shd->thd", pr, v[o:o + n]))
        o += n
    return torch.cat(outs)


def rel_err(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-3 * math.sqrt(b.numel())))


LENS = {
    "text_like": None,                       # 4 + U{8..100}, 60 sequences
    "one_row": [1],
    "full_tile": [128],
    "mixed_edges": [128, 128, 1, 127, 2, 126, 64, 64, 33],
    "many_small": [5] * 61 + [3, 1, 1, 7],
    "single_long_tail": [100, 28, 1],
}


def lens_of(name):
    if LENS[name] is None:
        g = torch.Generator().manual_seed(7)
        return (4 + torch.randint(8, 101, (60,), generator=g)).tolist()
    return LENS[name]


@pytest.mark.parametrize("group", [True, False])
@pytest.mark.parametrize("H", [1, 12])
@pytest.mark.parametrize("name", list(LENS))
def test_varlen_forward_backward(ops, name, H, group):
    lens = lens_of(name)
    qkv, cu = make(lens, H, seed=len(lens) + H)
    plan = ops.varlen_plan(cu, sum(lens), group=group)
    n_items = plan.n_items()
    assert not plan.error()
    firsts = plan.item_first_seq()
    assert firsts[0] == 0 and firsts[-1] == len(lens) and all(a < b for a, b in zip(firsts, firsts[1:]))
    cul = cu.tolist()
    assert all(cul[b] - cul[a] <= 128 for a, b in zip(firsts, firsts[1:]))
    assert plan.item_rows() == [cul[s] for s in firsts]
    if not group:
        assert n_items == len(lens)
    elif name == "many_small":
        assert n_items <= 4                                   # 319 rows -> 3 tiles (greedy)
    scale = 64 ** -0.5
    x = qkv.detach().clone().requires_grad_(True)
    o = ops.varlen_attention(x, plan)
    xr = qkv.detach().float().requires_grad_(True)
    o_ref = reference(xr, lens, scale)
    assert o.shape == o_ref.shape and o.dtype == torch.bfloat16
    assert float((o.float() - o_ref).abs().max()) <= FWD_ATOL
    go = torch.randn(o.shape, generator=torch.Generator().manual_seed(3)).bfloat16().cuda()
    o.backward(go)
    o_ref.backward(go.float())
    for i, nm in enumerate(("dq", "dk", "dv")):
        assert rel_err(x.grad[:, i], xr.grad[:, i]) <= GRAD_RTOL, (nm, rel_err(x.grad[:, i], xr.grad[:, i]))
    assert bool(torch.isfinite(x.grad.float()).all())


def test_varlen_strided_rows_and_scale(ops):
    """q / k / v as column slices of a wider row (row stride 3 * H * 64 + padding), non-default softmax scale."""
    lens = [17, 90, 104, 12, 55]
    H = 4
    g = torch.Generator().manual_seed(0)
    total = sum(lens)
    wide = torch.randn(total, 3 * H * 64 + 64, generator=g).bfloat16().cuda()
    qkv = wide[:, :3 * H * 64].view(total, 3, H, 64)          # non-contiguous rows
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32).cuda()
    plan = ops.varlen_plan(cu, total)
    o = ops.varlen_attention(qkv, plan, softmax_scale=0.2)
    o_ref = reference(qkv, lens, 0.2)
    assert float((o.float() - o_ref).abs().max()) <= FWD_ATOL


def keep_mask_from_plan(plan, cu, H, seed, p):
    """bool [total, H, 128]: the kernels' counter-hash mask for (packed row, head, tile-local key column)."""
    thr = min(65535, int(p * 65536.0 + 0.5)) << 16
    lo, hi = seed & _M32, (seed >> 32) & _M32
    mul = 0x9E3779B1

    def fmix(h):
        h = h ^ (h >> 16)
        h = (h * 0x85EBCA6B) & _M32
        h = h ^ (h >> 13)
        h = (h * 0xC2B2AE35) & _M32
        return h ^ (h >> 16)

    n_items = plan.n_items()
    firsts = plan.item_first_seq()
    cul = cu.tolist()
    total = cul[-1]
    keep = torch.zeros(total, H, 128, dtype=torch.bool)
    c = torch.arange(128, dtype=torch.int64).view(1, 1, 128)
    for item in range(n_items):
        r0, r1 = cul[firsts[item]], cul[firsts[item + 1]]
        w = item * H + torch.arange(H, dtype=torch.int64)
        salt = ((fmix((lo ^ ((w * mul) & _M32)) & _M32) + hi) & _M32).view(1, H, 1)
        r = torch.arange(r1 - r0, dtype=torch.int64).view(-1, 1, 1)
        chunk = fmix(salt ^ (((r * 4 + (c >> 5)) * mul) & _M32))           # drop_chunk(salt, r, c >> 5)
        lane = fmix((((c & 31) + 1) * mul) & _M32)                          # drop_lane(c & 31)
        x = ((chunk ^ lane) * mul) & _M32
        keep[r0:r1] = x >= thr
    return keep


@pytest.mark.parametrize("group", [True, False])
def test_varlen_dropout_forward_backward(ops, group):
    """dropout_p = 0.3 against the fp32 formula with the very same mask (regenerated from seed + plan)."""
    lens = lens_of("text_like")[:24] + [128, 3]
    H, p, seed = 4, 0.3, 0x0BAD_5EED_1234
    qkv, cu = make(lens, H, seed=99)
    plan = ops.varlen_plan(cu, sum(lens), group=group)
    keep_tile = keep_mask_from_plan(plan, cu.cpu(), H, seed, p)       # columns are tile-local
    # tile-local column of key j of sequence s = (cu[s] - row0 of its item) + j  ->  per-sequence [H, n, n] masks
    n_items = plan.n_items()
    firsts = plan.item_first_seq()
    cul = cu.tolist()
    keep_rows = torch.zeros(sum(lens), H, 128, dtype=torch.bool)
    seq_keep = []
    for item in range(n_items):
        row0 = cul[firsts[item]]
        for s in range(firsts[item], firsts[item + 1]):
            a, b = cul[s], cul[s + 1]
            seq_keep.append(keep_tile[a:b, :, a - row0:b - row0].permute(1, 0, 2))   # [H, n, n]
    assert abs(float(torch.cat([k.reshape(-1) for k in seq_keep]).float().mean()) - (1 - p)) < 2e-2

    def ref(x):
        q, k, v = (x[:, i].float() for i in range(3))
        outs = []
        for s, n in enumerate(lens):
            o = cul[s]
            sc = torch.einsum("thd,shd->hts", q[o:o + n], k[o:o + n]) * (64 ** -0.5)
            pr = torch.softmax(sc, dim=-1) * seq_keep[s].cuda().float() / (1.0 - p)
            outs.append(torch.einsum("hts,shd->thd", pr, v[o:o + n]))
        return torch.cat(outs)

    x = qkv.detach().clone().requires_grad_(True)
    o = ops.varlen_attention(x, plan, dropout_p=p, seed=seed)
    xr = qkv.detach().float().requires_grad_(True)
    o_ref = ref(xr)
    assert float((o.float() - o_ref).abs().max()) <= 3e-2
    go = torch.randn(o.shape, generator=torch.Generator().manual_seed(5)).bfloat16().cuda()
    o.backward(go)
    o_ref.backward(go.float())
    for i, nm in enumerate(("dq", "dk", "dv")):
        assert rel_err(x.grad[:, i], xr.grad[:, i]) <= GRAD_RTOL, nm


def test_varlen_plan_flags_long_sequences(ops):
    cu = torch.tensor([0, 10, 150, 160], dtype=torch.int32).cuda()
    plan = ops.varlen_plan(cu, 160)
    assert plan.error()


def test_varlen_large_batch_matches_reference(ops):
    """cfg3-sized packed text batch (4096 comments): every tile of a full-chip launch against the formula, forward only."""
    g = torch.Generator().manual_seed(11)
    lens = (4 + torch.randint(8, 101, (4096,), generator=g)).tolist()
    H = 12
    qkv, cu = make(lens, H, seed=5)
    plan = ops.varlen_plan(cu, sum(lens))
    o = ops.varlen_attention(qkv, plan)
    # reference on a sample of sequences (first, last, and every 97th)
    cul = cu.tolist()
    for s in list(range(0, 4096, 97)) + [4095]:
        a, b = cul[s], cul[s + 1]
        ref = reference(qkv[a:b], [b - a], 64 ** -0.5)
        assert float((o[a:b].float() - ref).abs().max()) <= FWD_ATOL, s


def _greedy_plan(lens, group):
    """Python restatement of the grouping: consecutive sequences share an item while they fit 128 rows."""
    items, rows, start, first, pos = [], [], 0, 0, 0
    for s, n in enumerate(lens):
        if s > first and (not group or pos + n - start > 128):
            items.append((start, first))
            start, first = pos, s
        rows += [(pos - start, pos - start + n)] * n
        pos += n
    items.append((start, first))
    return items, rows


@pytest.mark.parametrize("n_seq,group", [(1, True), (7, True), (4096, True), (4096, False), (11264, True), (11300, True)])
def test_varlen_plan_matches_greedy_restatement(ops, n_seq, group):
    """The plan words (items, per-row key windows) bit for bit against the sequential greedy grouping: the pointer-doubling
    kernels up to 11264 sequences, the single-thread kernel beyond."""
    g = torch.Generator().manual_seed(n_seq)
    lens = torch.randint(1, 129 if n_seq < 100 else 60, (n_seq,), generator=g).tolist()
    lens[0] = 128
    total = sum(lens)
    cu = torch.tensor([0] + list(torch.tensor(lens).cumsum(0)), dtype=torch.int32).cuda()
    plan = ops.varlen_plan(cu, total, group=group)
    items, rows = _greedy_plan(lens, group)
    assert not plan.error()
    assert plan.n_items() == len(items)
    w = plan.data.cpu().tolist()
    got_items = [(w[4 + 2 * i], w[5 + 2 * i]) for i in range(len(items) + 1)]
    assert got_items[:-1] == items and got_items[-1] == (total, n_seq)
    off = 4 + 2 * (n_seq + 1) + n_seq
    got_rows = [(x & 0xFFFF, x >> 16) for x in w[off:off + total]]
    assert got_rows == rows
