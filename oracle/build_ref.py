"""Build / load the reference's own Cython file as ``oracle/_ref/algos*.so``.

TEST INFRASTRUCTURE.  ``/root/reference/mDT/src/data/algos.pyx`` is cythonized from
where it lies (``language_level=2``: the file uses the Python-2 ``long`` type,
algos.pyx:13) and compiled with gcc.  Only the generated C file and the shared object
are written, both under ``oracle/_ref/`` (git-ignored, shipped to the GPU box by
gpurun).  No reference source is copied into the repository.
"""

from __future__ import annotations

import glob
import importlib.util
import os
import subprocess
import sys
import sysconfig

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")
REF_PYX = os.path.join(os.environ.get("MDT_REFERENCE_ROOT", "/root/reference"), "mDT", "src", "data", "algos.pyx")


def _so_path():
    hits = glob.glob(os.path.join(REF_DIR, "algos*.so"))
    return hits[0] if hits else None


def build_ref_algos(verbose: bool = False) -> str:
    import numpy as np

    if not os.path.isfile(REF_PYX):
        raise FileNotFoundError(REF_PYX)
    os.makedirs(REF_DIR, exist_ok=True)
    c_file = os.path.join(REF_DIR, "algos.c")
    subprocess.check_call(
        [sys.executable, "-m", "cython", "-2", REF_PYX, "-o", c_file],
        stdout=None if verbose else subprocess.DEVNULL,
        stderr=None if verbose else subprocess.DEVNULL,
    )
    ext = sysconfig.get_config_var("EXT_SUFFIX")
    so = os.path.join(REF_DIR, "algos" + ext)
    cmd = [
        "gcc", "-O2", "-fPIC", "-shared", "-w",
        "-DNPY_NO_DEPRECATED_API=NPY_1_7_API_VERSION",
        "-I" + sysconfig.get_paths()["include"], "-I" + np.get_include(),
        c_file, "-o", so,
    ]
    subprocess.check_call(cmd)
    return so


def load_ref_algos(build: bool = False):
    """Import ``oracle/_ref/algos*.so`` (optionally building it first). None if unavailable."""
    so = _so_path()
    if so is None and build and os.path.isfile(REF_PYX):
        so = build_ref_algos()
    if so is None:
        return None
    spec = importlib.util.spec_from_file_location("algos", so)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


if __name__ == "__main__":
    print(build_ref_algos(verbose=True))
