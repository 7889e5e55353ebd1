"""The C-ABI library builds, loads and exports every symbol ``include/mdt_b200.h`` declares
(CPU box: no compute calls)."""

import ctypes
import os
import re

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
HEADER = os.path.join(ROOT, "include", "mdt_b200.h")


def header_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mdt_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib_path():
    from multimodaldiscussiontransformer_b200.csrc import build

    return build.build()


def test_header_declares_the_hot_path():
    syms = header_symbols()
    for need in ("mdt_apsp_batched", "mdt_collate_graphs", "mdt_attn_bias_fwd", "mdt_attn_bias_bwd",
                 "mdt_biased_attn_fwd", "mdt_biased_attn_bwd", "mdt_node_feature_fwd", "mdt_gather_rows"):
        assert need in syms


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    missing = [s for s in header_symbols() if not hasattr(lib, s)]
    assert not missing, f"declared in mdt_b200.h but not exported: {missing}"
    lib.mdt_abi_version.restype = ctypes.c_int
    assert lib.mdt_abi_version() == 1


def test_ctypes_signatures_cover_the_header(lib_path):
    from multimodaldiscussiontransformer_b200 import _lib

    assert sorted(_lib.SIGNATURES) == header_symbols()
    _lib.load()


def test_no_cpu_fallback_without_device(lib_path):
    """On a box without a GPU the library refuses instead of computing on the host."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from multimodaldiscussiontransformer_b200 import _lib

    lib = _lib.load()
    assert lib.mdt_device_ok() == 0
    rc = lib.mdt_apsp_batched(ctypes.c_void_p(8), 1, 1, 4, ctypes.c_void_p(8), None, None, None)
    assert rc != 0
    assert b"CUDA" in lib.mdt_last_error_string() or b"sm_100" in lib.mdt_last_error_string()


def test_product_does_not_import_the_oracle():
    pkg = os.path.join(ROOT, "multimodaldiscussiontransformer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert "mdt_oracle" not in src and "import oracle" not in src and "from oracle" not in src, f
