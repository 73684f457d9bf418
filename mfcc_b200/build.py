"""In-tree build of libmfcc_b200.so with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "csrc")
SOURCES = ["mfcc_kernels.cu", "mfcc_post.cu", "mfcc_librosa.cu"]
DEPS = ["mfcc_kernels.cu", "mfcc_post.cu", "mfcc_librosa.cu", "mfcc_dev.cuh", "mfcc_f32.cuh", os.path.join("..", "..", "include", "mfcc_b200.h")]
OUT = os.path.join(_HERE, "libmfcc_b200.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC"]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(_SRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return OUT
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", OUT] + [
        os.path.join(_SRC, s) for s in SOURCES]
    env = dict(os.environ)
    # the image's CC/CXX may point at a wrapper nvcc does not understand
    env.pop("CC", None)
    env.pop("CXX", None)
    subprocess.check_call(cmd, cwd=_SRC, env=env)
    return OUT
