"""Build ``libmdt_b200.so`` in-tree with nvcc for sm_100a (no torch headers: plain C ABI).

    python -m multimodaldiscussiontransformer_b200.csrc.build [--force] [--verbose]

The shared object lands next to the package (``multimodaldiscussiontransformer_b200/libmdt_b200.so``),
is git-ignored and travels to the GPU box with the gpurun snapshot.
"""

from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.dirname(HERE)
ROOT = os.path.dirname(PKG)
LIB = os.path.join(PKG, "libmdt_b200.so")
OBJ_DIR = os.path.join(HERE, "_obj")
SOURCES = ["runtime.cu", "tensor_map.cu", "apsp.cu", "attn_bias.cu", "rows.cu", "layernorm.cu", "gelu.cu", "gemm.cu", "gemm2.cu", "wgrad.cu", "attn_fwd.cu", "attn_fwd3.cu", "attn_tail.cu", "attn_bwd.cu", "attn_bwd2.cu", "attn_varlen.cu", "umma_selftest.cu"]
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
FLAGS = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def nvcc() -> str:
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.isfile(exe):
        raise RuntimeError("nvcc not found: libmdt_b200 cannot be built")
    return exe


def _digest(paths) -> str:
    h = hashlib.sha256()
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def sources():
    return [os.path.join(HERE, s) for s in SOURCES if os.path.isfile(os.path.join(HERE, s))]


def build(force: bool = False, verbose: bool = False) -> str:
    srcs = sources()
    deps = srcs + [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".cuh")] + [
        os.path.join(ROOT, "include", "mdt_b200.h")]
    stamp = os.path.join(OBJ_DIR, "stamp")
    want = _digest(deps)
    if not force and os.path.isfile(LIB) and os.path.isfile(stamp) and open(stamp).read() == want:
        return LIB
    os.makedirs(OBJ_DIR, exist_ok=True)
    exe = nvcc()

    def compile_one(src):
        obj = os.path.join(OBJ_DIR, os.path.basename(src)[:-3] + ".o")
        cmd = [exe, *ARCH, *FLAGS, "-I", os.path.join(ROOT, "include"), "-c", src, "-o", obj]
        r = subprocess.run(cmd, capture_output=True, text=True)
        log = r.stdout + r.stderr
        with open(obj + ".log", "w") as f:
            f.write(log)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed for %s:\n%s" % (src, log))
        if verbose:
            print(log)
        return obj

    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(compile_one, srcs))
    cmd = [exe, *ARCH, "-shared", "-o", LIB, *objs, "-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
    with open(stamp, "w") as f:
        f.write(want)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
