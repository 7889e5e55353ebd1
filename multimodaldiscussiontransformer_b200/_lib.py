"""ctypes binding of ``libmdt_b200.so`` (the C ABI declared in ``include/mdt_b200.h``).

There is no fallback: if the shared object is missing or a call fails, a ``RuntimeError``
is raised.  PyTorch is used only for device memory and streams; every pointer handed to
the library is ``tensor.data_ptr()`` and every call runs on ``torch.cuda.current_stream()``.
"""

from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmdt_b200.so")

F32, BF16 = 0, 1
SPATIAL_CANTOR, SPATIAL_SPD = 0, 1
CODE_VIRTUAL, CODE_MASKED = 254, 255
CODE_SPATIAL_MAX, CODE_GRAD_MAX = 61, 22  # include/mdt_b200.h: spatial codes the fused kernels index / bin gradients for

_lib = None


class AttnDims(Structure):
    _fields_ = [("B", c_int32), ("H", c_int32), ("T", c_int32)] + [
        (f"{t}_stride_{a}", c_int64) for t in ("q", "k", "v", "o") for a in ("b", "t", "h")
    ]


# name -> (restype, argtypes); must list every symbol of include/mdt_b200.h (tests check this)
SIGNATURES = {
    "mdt_abi_version": (c_int, []),
    "mdt_last_error_string": (c_char_p, []),
    "mdt_device_ok": (c_int, []),
    "mdt_apsp_workspace_bytes": (c_int64, [c_int, c_int]),
    "mdt_apsp_batched": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "mdt_gen_edge_input": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "mdt_collate_graphs": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                   c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "mdt_collate_graphs_static": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p,
                                          c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "mdt_attn_bias_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int,
                                  c_void_p, c_void_p]),
    "mdt_attn_bias_bwd": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "mdt_biased_attn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, POINTER(AttnDims), c_void_p, c_int, c_void_p, c_int,
                                    c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_uint64, c_void_p]),
    "mdt_debug_attn_trace": (c_int, [c_void_p]),
    "mdt_biased_attn_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, POINTER(AttnDims),
                                    c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                    POINTER(AttnDims), c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_uint64, c_void_p]),
    "mdt_varlen_attn_plan_words": (c_int64, [c_int, c_int]),
    "mdt_varlen_attn_plan": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "mdt_varlen_attn_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_int, c_int, c_int,
                                    c_void_p, c_void_p, c_int64, c_void_p, c_float, c_float, c_uint64, c_void_p]),
    "mdt_varlen_attn_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int64, c_int64, c_void_p, c_int64, c_void_p,
                                    c_int64, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                    c_int64, c_int64, c_int64, c_float, c_float, c_uint64, c_void_p]),
    "mdt_node_feature_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                     c_int, c_int, c_int, c_void_p, c_void_p]),
    "mdt_node_feature_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p]),
    "mdt_gather_rows": (c_int, [c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "mdt_scatter_rows": (c_int, [c_void_p, c_int64, c_void_p, c_int, c_int, c_int, c_void_p, c_int64, c_void_p]),
    "mdt_dropout_add_ln_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64,
                                       c_int, c_float, c_uint64, c_float, c_void_p]),
    "mdt_dropout_add_ln_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                       c_int64, c_int, c_float, c_uint64, c_void_p]),
    "mdt_dropout_add_ln_bwd_colsum": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                              c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_uint64, c_void_p]),
    "mdt_dropout_add_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_uint64, c_void_p]),
    "mdt_dropout_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_uint64, c_void_p]),
    "mdt_gelu_fwd": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "mdt_gelu_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p]),
    "mdt_gemm_bias_act": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_void_p]),
    "mdt_gemm_dgrad_gelu": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "mdt_wgrad_gemm": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_void_p]),
    "mdt_umma_selftest": (c_int, [c_void_p] * 11),
}


def load():
    """Load the shared library once; raises if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m multimodaldiscussiontransformer_b200.csrc.build` "
                "(there is no CPU or PyTorch fallback for the mDT hot path)")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        if lib.mdt_abi_version() != 1:
            raise RuntimeError("libmdt_b200.so ABI version mismatch")
        _lib = lib
    return _lib


# C-ABI calls made so far (name -> count) and the kernels each one launches: bench.py reports
# the number of this library's kernel launches inside its timed region from these.
CALLS: dict = {}
KERNELS_PER_CALL = {"mdt_apsp_batched": 1, "mdt_gen_edge_input": 1, "mdt_collate_graphs": 2, "mdt_collate_graphs_static": 2, "mdt_attn_bias_fwd": 1,
                    "mdt_attn_bias_bwd": 1, "mdt_biased_attn_fwd": 1, "mdt_biased_attn_bwd": 2, "mdt_varlen_attn_plan": 2,
                    "mdt_varlen_attn_fwd": 1, "mdt_varlen_attn_bwd": 1, "mdt_node_feature_fwd": 1,
                    "mdt_node_feature_bwd": 1, "mdt_gather_rows": 1, "mdt_scatter_rows": 1, "mdt_umma_selftest": 1, "mdt_dropout_add_ln_fwd": 1,
                    "mdt_dropout_add_ln_bwd": 1}


def kernel_launches() -> int:
    return sum(KERNELS_PER_CALL.get(k, 1) * n for k, n in CALLS.items())


def check(rc: int, what: str) -> None:
    CALLS[what] = CALLS.get(what, 0) + 1
    if rc != 0:
        msg = load().mdt_last_error_string()
        raise RuntimeError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    return None if t is None else c_void_p(t.data_ptr())


def stream():
    import torch

    return c_void_p(torch.cuda.current_stream().cuda_stream)


def dtype_code(dt) -> int:
    import torch

    if dt == torch.float32:
        return F32
    if dt == torch.bfloat16:
        return BF16
    raise TypeError(f"libmdt_b200 handles float32 and bfloat16 buffers, not {dt}")
