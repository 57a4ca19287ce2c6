"""Builds fvm_b200/libfvm_b200.so (the C-ABI library, include/fvm_b200.h) in-tree with
nvcc for sm_100a.  The .so is git-ignored but travels to the GPU box with the snapshot."""
from __future__ import annotations

import concurrent.futures
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
CSRC = os.path.join(PKG, "csrc")
BUILD = os.path.join(PKG, "_build")
LIB = os.path.join(PKG, "libfvm_b200.so")
SOURCES = ["api.cu", "assemble.cu", "thomas.cu", "fused.cu", "fused_cta.cu", "spike.cu", "diagnostics.cu", "explicit.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: cannot build libfvm_b200.so")
    return exe


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(ROOT, "include", "fvm_b200.h"))
    return hdrs


def _stale(target, srcs):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in srcs)


def _compile(src, build_dir=None):
    obj = os.path.join(build_dir or BUILD, src.replace(".cu", ".o"))
    path = os.path.join(CSRC, src)
    if not _stale(obj, [path] + _deps()):
        return obj, ""
    cmd = [_nvcc()] + NVCC_FLAGS + ["-c", path, "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, res.stdout, res.stderr))
    with open(obj + ".ptxas.log", "w") as fh:
        fh.write(res.stderr)
    return obj, res.stderr


def build_library(force=False, verbose=False, build_dir=None, lib_path=None):
    """Compile (if stale) and return the path of libfvm_b200.so.  ``build_dir`` / ``lib_path``: build somewhere
    else than in the tree (tests/test_clean_build.py: a build from nothing but the sources)."""
    BUILD, LIB = build_dir or globals()["BUILD"], lib_path or globals()["LIB"]
    os.makedirs(BUILD, exist_ok=True)
    if force:
        for f in os.listdir(BUILD):
            os.remove(os.path.join(BUILD, f))
        if os.path.exists(LIB):
            os.remove(LIB)
    with concurrent.futures.ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        results = list(ex.map(lambda src: _compile(src, BUILD), SOURCES))
    objs = [r[0] for r in results]
    if verbose:
        for _, log in results:
            if log:
                print(log)
    if _stale(LIB, objs):
        cmd = [_nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (res.stdout, res.stderr))
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
