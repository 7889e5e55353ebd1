"""Parity of the integer / HBM-bound CUDA kernels against the oracle and the golden fixtures.

Every call goes through the C ABI (ctypes) of libmdt_b200.so.  Integer outputs are compared
bit for bit; float32 outputs of the gather kernels to 1e-6 (pure adds), bf16 to one ulp.
"""

import json
import os

import numpy as np
import pytest
import torch

import cases
import mdt_oracle as orc

pytestmark = pytest.mark.gpu

G = cases.GOLDEN_DIR


def load(name):
    return np.load(os.path.join(G, name), allow_pickle=False)


@pytest.fixture(scope="module")
def algos():
    from multimodaldiscussiontransformer_b200.data import algos as a

    return a


@pytest.fixture(scope="module")
def ops():
    from multimodaldiscussiontransformer_b200 import ops as o

    return o


# ---------------------------------------------------------------- A1 floyd_warshall
@pytest.mark.parametrize("name", list(cases.apsp_cases().keys()))
def test_floyd_warshall_golden(algos, name):
    gold = load("apsp_small.npz")
    adj = cases.apsp_cases()[name]
    M, path = algos.floyd_warshall(adj)                      # Cython-signature wrapper (path kernel)
    assert M.dtype == np.int64 and path.dtype == np.int64
    np.testing.assert_array_equal(M, gold[name + ".M"])
    np.testing.assert_array_equal(path, gold[name + ".path"])
    t = torch.from_numpy(np.ascontiguousarray(adj.astype(np.int64))).cuda().unsqueeze(0)
    M2, p2 = algos.floyd_warshall_batched(t)                 # register-tile kernel, M only
    assert p2 is None
    np.testing.assert_array_equal(M2[0].cpu().numpy(), gold[name + ".M"])


@pytest.mark.parametrize("name", list(cases.apsp_digest_cases().keys()))
def test_floyd_warshall_large_digest(algos, name):
    dig = json.load(open(os.path.join(G, "apsp_digests.json")))
    adj = cases.apsp_digest_cases()[name]
    M, path = algos.floyd_warshall(adj)
    assert cases.digest(M) == dig[name]["M"]
    assert cases.digest(path) == dig[name]["path"]
    t = torch.from_numpy(adj.astype(np.uint8)).cuda().unsqueeze(0)
    M2, _ = algos.floyd_warshall_batched(t)
    assert cases.digest(M2[0].cpu().numpy()) == dig[name]["M"]


@pytest.mark.parametrize("n", [16, 32, 64, 100, 128, 200, 256, 384, 512])
@pytest.mark.parametrize("dtype", [torch.uint8, torch.int32, torch.int64])
def test_floyd_warshall_batched_vs_oracle(algos, n, dtype):
    """BASELINE config 2 sweep (reduced batch so the CPU oracle finishes in seconds)."""
    rng = np.random.default_rng(n)
    B = 6 if n <= 256 else 3
    adj = np.stack([cases.syn.tree_adjacency(cases.syn.random_tree(n, rng)) for _ in range(B - 1)]
                   + [cases.syn.forest_adjacency(n, 3, rng)])
    M_ref, path_ref = orc.floyd_warshall_batch(adj.astype(np.int64))
    t = torch.from_numpy(adj).cuda().to(dtype)
    M, _ = algos.floyd_warshall_batched(t)
    np.testing.assert_array_equal(M.cpu().numpy(), M_ref)
    if dtype == torch.uint8:
        M, path = algos.floyd_warshall_batched(t, need_path=True)
        np.testing.assert_array_equal(M.cpu().numpy(), M_ref)
        np.testing.assert_array_equal(path.cpu().numpy(), path_ref)


def test_floyd_warshall_weighted_and_cyclic(algos):
    rng = np.random.default_rng(7)
    for n in (12, 37, 90):
        w = rng.integers(0, 6, size=(4, n, n)).astype(np.int64)
        w[rng.random((4, n, n)) < 0.7] = 0
        M_ref, path_ref = orc.floyd_warshall_batch(w)
        M, path = algos.floyd_warshall_batched(torch.from_numpy(w).cuda(), need_path=True)
        np.testing.assert_array_equal(M.cpu().numpy(), M_ref)
        np.testing.assert_array_equal(path.cpu().numpy(), path_ref)
        M2, _ = algos.floyd_warshall_batched(torch.from_numpy(w).cuda())
        np.testing.assert_array_equal(M2.cpu().numpy(), M_ref)


def test_floyd_warshall_full_size_properties(algos):
    """Config 2 at full size (batch 256, n = 512): size-independent properties + sampled oracle rows."""
    rng = np.random.default_rng(99)
    B, n = 256, 512
    parents = [cases.syn.random_tree(n, rng) for _ in range(B)]
    adj = torch.from_numpy(np.stack([cases.syn.tree_adjacency(p) for p in parents])).cuda()
    M, _ = algos.floyd_warshall_batched(adj)
    assert M.shape == (B, n, n)
    assert bool((M == M.transpose(1, 2)).all())
    assert bool((M.diagonal(dim1=1, dim2=2) == 0).all())
    assert int(M.max()) <= 510
    # tree identity: M == up + down, checked on 4 graphs
    for b in (0, 17, 128, 255):
        hops = cases.syn.hop_pairs(parents[b]).sum(-1)
        np.testing.assert_array_equal(M[b].cpu().numpy(), np.minimum(hops, 510))
    # idempotence: APSP of the weighted closure is the closure itself
    Mc = M[:8].clone()
    M2, _ = algos.floyd_warshall_batched(Mc)
    assert bool((M2 == Mc).all())


def test_floyd_warshall_empty_and_errors(algos):
    M, p = algos.floyd_warshall_batched(torch.zeros((0, 4, 4), dtype=torch.uint8, device="cuda"))
    assert M.shape == (0, 4, 4)
    with pytest.raises(AssertionError):
        algos.floyd_warshall(np.zeros((3, 4), dtype=np.int64))
    with pytest.raises(RuntimeError):
        algos.floyd_warshall_batched(torch.zeros((1, 513, 513), dtype=torch.uint8, device="cuda"))


@pytest.mark.parametrize("name", ["tree16", "cyclic20"])
def test_gen_edge_input(algos, name):
    gold = load("apsp_small.npz")
    adj = cases.apsp_cases()[name]
    M, path = algos.floyd_warshall(adj)
    max_dist = int(M[M < 510].max())
    out = algos.gen_edge_input(max_dist, path, gold[name + ".edge_feat"])
    np.testing.assert_array_equal(out, gold[name + ".edge_input"])
    with pytest.raises(IndexError):
        algos.gen_edge_input(max_dist - 1, path, gold[name + ".edge_feat"])


# ---------------------------------------------------------------- A2 + A3 preprocess + collate
@pytest.mark.parametrize("name", list(cases.COLLATE_CASES.keys()))
def test_collator_golden(name):
    from multimodaldiscussiontransformer_b200.data.collator import collator

    gold = load("collate.npz")
    items, spmax = cases.collate_items(name)
    batch = collator(items, spmax)
    for k in ("idx", "attn_bias", "spatial_pos", "in_degree", "out_degree", "x_token_mask", "x", "x_token_type_ids",
              "x_attention_mask", "x_image_indexes", "y", "y_mask"):
        g = gold[f"{name}.{k}"]
        v = batch[k].cpu().numpy()
        assert v.dtype == g.dtype, (k, v.dtype, g.dtype)
        assert v.shape == g.shape, (k, v.shape, g.shape)
        np.testing.assert_array_equal(v, g, err_msg=k)
    shp = tuple(gold[f"{name}.x_images.shape"])
    if batch["x_images"] is None:
        assert shp == (0,)
    else:
        assert tuple(batch["x_images"].shape) == shp


@pytest.mark.parametrize("sizes,spmax", [([128] * 8, 5), ([16, 300, 77, 512, 1, 2], 5), ([257, 255], 3), ([40, 33], 100)])
def test_collate_graph_tensors_vs_oracle(sizes, spmax):
    from multimodaldiscussiontransformer_b200.data.collator import collate_graph_tensors

    items = cases.syn.synthetic_items(sizes, 5, with_images=False, as_torch=False)
    ref = orc.collator(items, spmax)
    out = collate_graph_tensors([it.edge_index for it in items], sizes, spmax, torch.device("cuda"))
    np.testing.assert_array_equal(out["spatial_pos"].cpu().numpy(), ref["spatial_pos"])
    np.testing.assert_array_equal(out["attn_bias"].cpu().numpy(), ref["attn_bias"])
    np.testing.assert_array_equal(out["in_degree"].cpu().numpy(), ref["in_degree"])
    # bias codes: 255 where attn_bias is -inf, 254 on the graph-token row/column, spatial index elsewhere
    code = out["bias_code"].cpu().numpy()
    T = max(sizes) + 1
    assert (code[:, :, T:] == 255).all()
    code = code[:, :, :T]
    ab = ref["attn_bias"]
    assert ((code == 255) == np.isneginf(ab)).all()
    live = ~np.isneginf(ab)
    assert (code[:, 0, :][live[:, 0, :]] == 254).all() and (code[:, :, 0][live[:, :, 0]] == 254).all()
    inner = live[:, 1:, 1:]
    np.testing.assert_array_equal(code[:, 1:, 1:][inner], ref["spatial_pos"][inner].astype(np.uint8))


# ---------------------------------------------------------------- A5 GraphAttnBias
def _bias_inputs(B, N, H, seed, dtype, device="cuda"):
    g = torch.Generator().manual_seed(seed)
    T = N + 1
    sp = torch.randint(0, 22, (B, N, N), generator=g).int()
    ab = torch.zeros(B, T, T)
    ab[:, 1:, 1:][torch.rand(B, N, N, generator=g) < 0.3] = float("-inf")
    ab[:, :, T - max(1, N // 5):] = float("-inf")
    table = torch.randn(512, H, generator=g) * 0.5
    virt = torch.randn(1, H, generator=g)
    return ab.to(dtype).to(device), sp.to(device), table.to(device), virt.to(device)


@pytest.mark.parametrize("B,N,H", [(3, 8, 4), (2, 128, 12), (1, 512, 12), (5, 33, 32), (2, 1, 12)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_attn_bias_fwd_bwd(ops, B, N, H, dtype):
    ab, sp, table, virt = _bias_inputs(B, N, H, N, dtype)
    table.requires_grad_(True)
    virt.requires_grad_(True)
    out = ops.attn_bias(ab, sp, table, virt)
    # oracle in the same dtype (the reference module runs in the model dtype)
    t_ref = table.detach().cpu().to(dtype).requires_grad_(True)
    v_ref = virt.detach().cpu().to(dtype).requires_grad_(True)
    ref = orc.graph_attn_bias(ab.cpu(), sp.cpu(), t_ref, v_ref)
    assert out.dtype == dtype and out.shape == ref.shape
    o, r = out.float().cpu(), ref.float()
    assert torch.equal(torch.isneginf(o), torch.isneginf(r))
    fin = ~torch.isneginf(r)
    assert torch.equal(o[fin], r[fin]), float((o[fin] - r[fin]).abs().max())
    g = torch.Generator().manual_seed(3)
    go = torch.randn(out.shape, generator=g).to(dtype)
    out.backward(go.cuda())
    # fp64 reference of the reduction (the reference accumulates in the model dtype)
    t64 = table.detach().cpu().double().requires_grad_(True)
    v64 = virt.detach().cpu().double().requires_grad_(True)
    orc.graph_attn_bias(torch.zeros_like(ab.cpu(), dtype=torch.float64), sp.cpu(), t64, v64).backward(go.double())
    gt = t64.grad.clone()
    gt[0] = 0  # padding_idx=0: no gradient (graphormer_layers.py:80)
    scale = max(1.0, float(gt.abs().max()))
    tol = 1e-5 * scale if dtype == torch.float32 else 2e-3 * scale
    np.testing.assert_allclose(table.grad.cpu().double().numpy(), gt.numpy(), atol=tol, rtol=1e-4)
    np.testing.assert_allclose(virt.grad.cpu().double().numpy(), v64.grad.numpy(), atol=tol * 4, rtol=1e-4)


@pytest.mark.parametrize("N,H,hi", [(33, 12, 60), (130, 16, 40), (64, 5, 22)])
def test_attn_bias_bwd_bf16_wide_indices_and_unaligned_gradient(ops, N, H, hi):
    """The tensor-core histogram (bf16, H <= 16): spatial indices >= 32 take its atomic side path, and a gradient that is a
    2-byte-aligned view (planes start at any alignment) is read through aligned vectors; against an fp64 scatter-add."""
    from multimodaldiscussiontransformer_b200 import _lib

    B, T = 3, N + 1
    g = torch.Generator().manual_seed(N)
    sp = torch.randint(0, hi, (B, N, N), generator=g).int().cuda()
    buf = torch.randn(B * H * T * T + 3, generator=g).bfloat16().cuda()
    go = buf[3:].view(B, H, T, T)                                   # 6 bytes off the allocation's alignment
    S = 64
    g_table = torch.zeros((S, H), dtype=torch.float32, device="cuda")
    g_virt = torch.zeros((H,), dtype=torch.float32, device="cuda")
    lib = _lib.load()
    _lib.check(lib.mdt_attn_bias_bwd(_lib.ptr(go), _lib.ptr(sp), B, H, N, S, _lib.dtype_code(go.dtype), _lib.ptr(g_table),
                                     _lib.ptr(g_virt), _lib.stream()), "mdt_attn_bias_bwd")
    g64 = go.double()
    want = torch.zeros(S, H, dtype=torch.float64, device="cuda")
    want.index_add_(0, sp.long().reshape(-1), g64[:, :, 1:, 1:].permute(0, 2, 3, 1).reshape(-1, H))
    want[0] = 0
    want_v = g64[:, :, 0, :].sum((0, 2)) + g64[:, :, 1:, 0].sum((0, 2))
    scale = max(1.0, float(want.abs().max()))
    np.testing.assert_allclose(g_table.double().cpu().numpy(), want.cpu().numpy(), atol=1e-4 * scale, rtol=1e-5)
    np.testing.assert_allclose(g_virt.double().cpu().numpy(), want_v.cpu().numpy(), atol=1e-4 * scale, rtol=1e-5)


def test_attn_bias_golden(ops):
    gold = load("graph_modules.npz")
    inp = cases.graph_module_inputs()
    keys = [k for k in gold.files if k.startswith("gab.")]
    assert "gab.out" in keys, keys
    cb = torch.nn.Module()
    H = 4
    cb.edge_encoder = torch.nn.Embedding(17, H, padding_idx=0)
    cb.edge_dis_encoder = torch.nn.Embedding(8 * H * H, 1)
    cb.spatial_pos_encoder = torch.nn.Embedding(32, H, padding_idx=0)
    cb.graph_token_virtual_distance = torch.nn.Embedding(1, H)
    orc.seed_parameters(cb, 32)
    out = ops.attn_bias(inp["attn_bias"].cuda(), inp["spatial_pos"].cuda(), cb.spatial_pos_encoder.weight.detach().cuda(),
                        cb.graph_token_virtual_distance.weight.detach().cuda())
    g = gold["gab.out"]
    o = out.cpu().numpy()
    assert (np.isneginf(o) == np.isneginf(g)).all()
    np.testing.assert_allclose(o[~np.isneginf(g)], g[~np.isneginf(g)], atol=1e-6)


def test_attn_bias_general_values(ops):
    """attn_bias holding arbitrary finite values takes the element-wise path."""
    ab, sp, table, virt = _bias_inputs(2, 19, 6, 1, torch.float32)
    ab = torch.where(torch.isneginf(ab), ab, torch.randn_like(ab))
    out = ops.attn_bias(ab, sp, table, virt)
    ref = orc.graph_attn_bias(ab.cpu(), sp.cpu(), table.cpu(), virt.cpu())
    fin = ~torch.isneginf(ref)
    assert torch.equal(torch.isneginf(out.cpu()), ~fin)
    np.testing.assert_allclose(out.cpu()[fin].numpy(), ref[fin].numpy(), atol=1e-6)


# ---------------------------------------------------------------- A4 GraphNodeFeature
@pytest.mark.parametrize("B,N,C", [(3, 8, 64), (4, 128, 768), (2, 1, 768)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_node_feature_fwd_bwd(ops, B, N, C, dtype):
    g = torch.Generator().manual_seed(B * 100 + N)
    x = torch.randn(B, N, C, generator=g).to(dtype)
    deg = torch.randint(0, 7, (B, N), generator=g)
    w_in = (torch.randn(512, C, generator=g) * 0.1).to(dtype)
    w_out = (torch.randn(512, C, generator=g) * 0.1).to(dtype)
    tok = torch.randn(1, C, generator=g).to(dtype)
    xs = [t.cuda().requires_grad_(True) for t in (x, w_in, w_out, tok)]
    out = ops.node_feature(xs[0], deg.cuda(), deg.cuda(), xs[1], xs[2], xs[3])
    ref = orc.graph_node_feature(x, deg, deg, w_in, w_out, tok)
    assert torch.equal(out.cpu(), ref), float((out.cpu().float() - ref.float()).abs().max())
    go = torch.randn(out.shape, generator=g).to(dtype)
    out.backward(go.cuda())
    x64, wi64, wo64, tk64 = (t.double().requires_grad_(True) for t in (x, w_in, w_out, tok))
    orc.graph_node_feature(x64, deg, deg, wi64, wo64, tk64).backward(go.double())
    assert torch.equal(xs[0].grad.cpu(), go[:, 1:, :])
    gi = wi64.grad.clone()
    gi[0] = 0
    go_ = wo64.grad.clone()
    go_[0] = 0
    tol = 1e-4 if dtype == torch.float32 else 0.05 * float(gi.abs().max() + 1)
    np.testing.assert_allclose(xs[1].grad.cpu().double().numpy(), gi.numpy(), atol=tol)
    np.testing.assert_allclose(xs[2].grad.cpu().double().numpy(), go_.numpy(), atol=tol)
    np.testing.assert_allclose(xs[3].grad.cpu().double().numpy(), tk64.grad.numpy(), atol=tol)


# ---------------------------------------------------------------- K12 gather / scatter rows
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gather_scatter_rows(ops, dtype):
    g = torch.Generator().manual_seed(0)
    rows, C = 1000, 768
    src = torch.randn(rows, C, generator=g).to(dtype).cuda().requires_grad_(True)
    index = torch.randperm(rows, generator=g)[:377].int().cuda()
    out = ops.gather_rows(src, index)
    assert torch.equal(out, src.detach()[index.long()])
    go = torch.randn(out.shape, generator=g).to(dtype).cuda()
    out.backward(go)
    ref = torch.zeros_like(src)
    ref[index.long()] = go
    assert torch.equal(src.grad, ref)

    base = torch.randn(rows, C, generator=g).to(dtype).cuda().requires_grad_(True)
    upd = torch.randn(377, C, generator=g).to(dtype).cuda().requires_grad_(True)
    res = ops.scatter_rows(base, upd, index)
    ref = base.detach().clone()
    ref[index.long()] = upd.detach()
    assert torch.equal(res, ref)
    g2 = torch.randn(res.shape, generator=g).to(dtype).cuda()
    res.backward(g2)
    assert torch.equal(upd.grad, g2[index.long()])
    gb = g2.clone()
    gb[index.long()] = 0
    assert torch.equal(base.grad, gb)
    # strided views (a [n, nb, C] bottleneck tensor's token 0)
    bn = torch.randn(50, 4, C, generator=g).to(dtype).cuda()
    idx = torch.arange(50, dtype=torch.int32, device="cuda") * 3
    dst = torch.zeros(200, C, dtype=dtype, device="cuda")
    from multimodaldiscussiontransformer_b200.ops import _move

    _move(True, bn[:, 0, :], idx, dst)
    assert torch.equal(dst[idx.long()], bn[:, 0, :])


# ---------------------------------------------------------------- tcgen05 building blocks
def test_umma_descriptor_selftest():
    from multimodaldiscussiontransformer_b200 import _lib

    lib = _lib.load()
    g = torch.Generator().manual_seed(0)
    mk = lambda *s: torch.randn(*s, generator=g).bfloat16().cuda()
    a, b, p, v, at, bk = mk(128, 64), mk(128, 64), mk(128, 128), mk(128, 64), mk(128, 128), mk(64, 128)
    d1 = torch.zeros(128, 128, device="cuda")
    d2 = torch.zeros(128, 64, device="cuda")
    d3 = torch.zeros(128, 64, device="cuda")
    d4 = torch.zeros(128, 64, device="cuda")
    _lib.check(lib.mdt_umma_selftest(*[_lib.ptr(t) for t in (a, b, p, v, at, bk, d1, d2, d3, d4)], _lib.stream()),
               "mdt_umma_selftest")
    torch.cuda.synchronize()
    r1 = a.float() @ b.float().T
    r2 = p.float() @ v.float()
    r3 = at.float().T @ bk.float().T
    for name, d, r in (("K-major x K-major", d1, r1), ("K-major x MN-major", d2, r2), ("MN-major x K-major", d3, r3),
                       ("A from tensor memory x MN-major", d4, r2)):
        err = float((d - r).abs().max())
        assert err < 1e-2, (name, err)
