"""Build recipe for liblmi_b200.so: one nvcc invocation, sm_100a only, in-tree output.

``python -m lmi_b200.build`` (or ``__graft_entry__.build()``) cross-compiles without a GPU.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "liblmi_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-shared", "-Xcompiler", "-fPIC",
]


def sources():
    out = []
    for root, _, files in os.walk(CSRC):
        out += [os.path.join(root, f) for f in files if f.endswith((".cu", ".cuh", ".inl", ".h"))]
    out.append(os.path.join(os.path.dirname(PKG_DIR), "include", "lmi_b200.h"))
    return out


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(s) > t for s in sources() if os.path.exists(s))


# development-only twin with the kernels' experiment switches compiled in (-DLMI_EXPERIMENTS: option "dbg_flags",
# results invalid by construction).  Used by scripts/dev only; the package never loads it.
EXP_LIB_PATH = os.path.join(PKG_DIR, "liblmi_b200_exp.so")


def _stale(path) -> bool:
    if not os.path.exists(path):
        return True
    t = os.path.getmtime(path)
    return any(os.path.getmtime(s) > t for s in sources() if os.path.exists(s))


def _nvcc(out_path, extra, verbose):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: cannot build " + os.path.basename(out_path))
    cmd = [nvcc] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", out_path, os.path.join(CSRC, "lmi_api.cu")]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return out_path


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    return _nvcc(LIB_PATH, [], verbose)


def build_experiments(force: bool = False, verbose: bool = False) -> str:
    if not force and not _stale(EXP_LIB_PATH):
        return EXP_LIB_PATH
    return _nvcc(EXP_LIB_PATH, ["-DLMI_EXPERIMENTS"], verbose)


if __name__ == "__main__":
    if "--experiments" in sys.argv:
        print(build_experiments(force="--force" in sys.argv, verbose="-v" in sys.argv))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
