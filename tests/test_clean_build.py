"""The library builds from nothing but the sources: every .cu compiled for sm_100a into an empty directory,
linked, and the result exports what include/fvm_b200.h declares.  (The GPU box runs the .so that travels with
the snapshot; this is the check that a clean checkout can make it.  ~1.5 minutes of nvcc.)"""
import ctypes
import os
import re

from fvm_b200 import build as B

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_clean_build_exports_the_header(tmp_path):
    lib = B.build_library(force=True, build_dir=str(tmp_path / "obj"), lib_path=str(tmp_path / "libfvm_b200.so"))
    assert os.path.exists(lib) and lib.startswith(str(tmp_path))
    objs = sorted(f for f in os.listdir(tmp_path / "obj") if f.endswith(".o"))
    assert objs == sorted(s.replace(".cu", ".o") for s in B.SOURCES)
    handle = ctypes.CDLL(lib)
    header = open(os.path.join(ROOT, "include", "fvm_b200.h")).read()
    names = set(re.findall(r"\b(fvm_[a-z0-9_]+)\s*\(", header)) - {"fvm_allgather_fn"}
    assert len(names) > 25
    for n in names:
        assert hasattr(handle, n), n
    assert handle.fvm_version() >= 200
