"""The committed fixtures ARE the unmodified reference's outputs: regenerate each of them in memory from
``/root/reference`` (through ``tests/golden/ref_harness.py`` + ``make_golden.py``) and compare with the files
under ``tests/golden/``.  CPU only; skipped where ``/root/reference`` does not exist (the GPU box).

Integer / index fixtures must match exactly; float fixtures to 2e-5 absolute (the reference's CPU
kernels may split reductions differently with another thread count)."""

from __future__ import annotations

import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
G = os.path.join(HERE, "golden")
for p in (G, os.path.join(HERE, "..", "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)
import ref_harness as rh  # noqa: E402

pytestmark = pytest.mark.skipif(not rh.reference_available(), reason="/root/reference is absent on this box")


def _regenerate(fn, *args, **kw):
    import make_golden as mg

    mg.CAPTURE = {}
    try:
        fn(mg, *args, **kw)
        return mg.CAPTURE
    finally:
        mg.CAPTURE = None


def _compare_npz(name, live, only_prefix=None):
    committed = np.load(os.path.join(G, name))
    keys = [k for k in live if only_prefix is None or k.startswith(only_prefix)]
    assert keys, name
    for k in keys:
        assert k in committed.files, (name, k)
        a, b = live[k], committed[k]
        assert a.shape == b.shape and a.dtype == b.dtype, (name, k, a.shape, b.shape, a.dtype, b.dtype)
        if np.issubdtype(a.dtype, np.floating):
            both_inf = np.isinf(a) & np.isinf(b) & (np.sign(a) == np.sign(b))
            np.testing.assert_allclose(np.where(both_inf, 0, a), np.where(both_inf, 0, b), atol=2e-5, rtol=1e-4, err_msg=f"{name}:{k}")
        else:
            assert np.array_equal(a, b), (name, k)
    if only_prefix is None:
        assert set(committed.files) == set(live), (name, set(committed.files) ^ set(live))


def test_live_reference_apsp():
    cap = _regenerate(lambda mg: mg.gen_apsp())
    _compare_npz("apsp_small.npz", cap["apsp_small.npz"])
    committed = json.load(open(os.path.join(G, "apsp_digests.json")))
    for name, d in cap["apsp_digests.json"].items():
        assert committed[name]["M"] == d["M"] and committed[name]["path"] == d["path"] and committed[name]["n"] == d["n"], name


def test_live_reference_hop_pairs():
    cap = _regenerate(lambda mg: mg.gen_hops())
    _compare_npz("hop_pairs.npz", cap["hop_pairs.npz"])


def test_live_reference_collate():
    cap = _regenerate(lambda mg: mg.gen_collate())
    _compare_npz("collate.npz", cap["collate.npz"])


def test_live_reference_graph_modules():
    cap = _regenerate(lambda mg: mg.gen_modules())
    _compare_npz("graph_modules.npz", cap["graph_modules.npz"])


def _check_meta(committed, live):
    for case, m in live.items():
        c = committed[case]
        assert c["n_params"] == m["n_params"] and c["state_keys"] == m["state_keys"], case
        for k, v in m["grad_norms"].items():
            w = c["grad_norms"][k]
            assert (v is None) == (w is None), (case, k)
            if v is not None:
                assert abs(v - w) <= 1e-3 * max(abs(w), 1e-6) + 1e-7, (case, k, v, w)


def test_live_reference_encoder_small():
    cap = _regenerate(lambda mg: mg.gen_encoder(skip_cfg1=True))
    live = cap["encoder.npz"]
    for case in ("tiny", "tiny_stack2", "tiny_noimg"):
        _compare_npz("encoder.npz", live, only_prefix=case + ".")
    _check_meta(json.load(open(os.path.join(G, "encoder_meta.json"))), cap["encoder_meta.json"])
    assert cap["encoder_state_keys.json"] == json.load(open(os.path.join(G, "encoder_state_keys.json")))


@pytest.mark.slow
def test_live_reference_encoder_cfg1():
    """BASELINE.json configs[0] through the verbatim reference (about 20 s of CPU)."""
    cap = _regenerate(lambda mg: mg.gen_encoder(only=("cfg1",)))
    _compare_npz("encoder.npz", cap["encoder.npz"], only_prefix="cfg1.")
    _check_meta(json.load(open(os.path.join(G, "encoder_meta.json"))), cap["encoder_meta.json"])


@pytest.mark.slow
def test_live_reference_encoder_big():
    cap = _regenerate(lambda mg: mg.gen_encoder_big())
    _compare_npz("encoder_big.npz", cap["encoder_big.npz"])
    _check_meta(json.load(open(os.path.join(G, "encoder_big_meta.json"))), cap["encoder_big_meta.json"])
