"""Host-side logic and the C-ABI surface, without a GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

import fvm_b200
from fvm_b200 import _lib, workloads as W
from fvm_b200.batched import _alpha_ends, _bc_kinds

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    """Every function include/fvm_b200.h declares is exported by the built library and
    bound with a signature in fvm_b200/_lib.py."""
    hdr = open(os.path.join(ROOT, "include", "fvm_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(fvm_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    assert declared == set(_lib.exported_symbols())
    handle = _lib.lib()                       # raises if the .so is missing
    for name in declared:
        assert hasattr(handle, name)
    assert handle.fvm_version() >= 100
    assert handle.fvm_fused_max_cells() >= 1024
    assert b"invalid" in handle.fvm_status_string(_lib.FVM_ERR_INVALID_ARG)


def test_argument_validation_needs_no_gpu():
    """Bad arguments are rejected before anything is launched."""
    handle = _lib.lib()
    pr = _lib.FvmProblem()
    assert handle.fvm_assemble(ctypes.byref(pr), None, None, None, None, None, 1, 1, None, None, None) \
        == _lib.FVM_ERR_INVALID_ARG
    assert handle.fvm_theta_run_fused(ctypes.byref(pr), None, 0.5, 1, 0, None, None, None) \
        == _lib.FVM_ERR_INVALID_ARG
    assert handle.fvm_theta_step_thomas(2, 1, None, None, None, 1, 1, 0.5, None, None, None, 1, 1, 1, None) \
        == _lib.FVM_ERR_INVALID_ARG


def test_struct_layout_matches_header():
    # 2 + 1 + 3*3 int64/pointer fields, 5 int32 (+4 pad), 2 pointers
    assert ctypes.sizeof(_lib.FvmProblem) == 8 * 12 + 4 * 5 + 4 + 8 * 2


def test_struct_layouts_against_the_c_compiler(tmp_path):
    """include/fvm_b200.h compiles as plain C, and the ctypes mirrors of its two structs have the sizes and field
    offsets the C compiler gives them."""
    import shutil
    import subprocess
    cc = shutil.which("gcc") or shutil.which("cc")
    if cc is None:
        pytest.skip("no C compiler")
    fields = {"fvm_problem": [f[0] for f in _lib.FvmProblem._fields_],
              "fvm_spike_plan": [f[0] for f in _lib.FvmSpikePlan._fields_]}
    src = ['#include <stdio.h>', '#include <stddef.h>', '#include "fvm_b200.h"', 'int main(void) {']
    for st, names in fields.items():
        src.append('  printf("%s %%zu\\n", sizeof(%s));' % (st, st))
        for n in names:
            src.append('  printf("%s.%s %%zu\\n", offsetof(%s, %s));' % (st, n, st, n))
    src += ['  return 0;', '}']
    c = tmp_path / "layout.c"
    c.write_text("\n".join(src))
    exe = tmp_path / "layout"
    subprocess.run([cc, "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)],
                   check=True)
    out = dict(ln.split() for ln in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.splitlines())
    for st, cls in (("fvm_problem", _lib.FvmProblem), ("fvm_spike_plan", _lib.FvmSpikePlan)):
        assert int(out[st]) == ctypes.sizeof(cls), st
        for n in fields[st]:
            assert int(out["%s.%s" % (st, n)]) == getattr(cls, n).offset, (st, n)


def test_mesh_and_cellvariable():
    f = np.array([0.0, 0.1, 0.4, 0.5, 1.0])
    m = fvm_b200.Mesh(f)
    assert m.J == 4 and np.allclose(m.cells, [0.05, 0.25, 0.45, 0.75])
    assert np.allclose(m.cell_widths, [0.1, 0.3, 0.1, 0.5])
    assert m.h(1) == pytest.approx(0.3) and m.hm(1) == pytest.approx(0.2) and m.hp(1) == pytest.approx(0.2)
    with pytest.raises(ValueError):
        m.hm(0)                                               # fvm_classes.py:49-50
    with pytest.raises(ValueError):
        m.hp(3)                                               # :55-56
    with pytest.raises(ValueError):
        fvm_b200.Mesh([0, 0.5, 0.5, 1])                       # :36-37
    fvm_b200.Mesh([0, 0.6, 0.3, 1])                           # unsorted faces are accepted, like the reference
    v = fvm_b200.CellVariable(2.0, mesh=m)
    assert v.shape == (4,) and np.all(v == 2.0) and (v * 3).mesh is m
    u = fvm_b200.CellVariable(np.array([1.0, 2.0, 4.0, 8.0]), mesh=m)
    i = np.array([1, 2])
    # p(i) and m(i+1) are the same face
    assert np.allclose(u.p(i), u.m(i + 1))
    assert u.m(1) == pytest.approx(0.3 / 0.4 * 1.0 + 0.1 / 0.4 * 2.0)


def test_bc_resolution_and_alpha_quirk():
    assert _bc_kinds(None, None, 1.0, 0.0) == (_lib.BC_DIRICHLET, _lib.BC_DIRICHLET)
    assert _bc_kinds(0.0, 0.0, None, None) == (_lib.BC_ROBIN, _lib.BC_ROBIN)
    assert _bc_kinds(0.0, 0.0, 1.0, None) == (_lib.BC_DIRICHLET, _lib.BC_ROBIN)   # value beats flux
    with pytest.raises(UnboundLocalError):
        _bc_kinds(None, None, None, None)
    # both ends follow left_flux (fvm_classes.py:237-238)
    assert _alpha_ends(None, _lib.BC_DIRICHLET, _lib.BC_ROBIN, False) == (0, 0)
    assert _alpha_ends(0.0, _lib.BC_ROBIN, _lib.BC_DIRICHLET, False) == (1, 1)
    assert _alpha_ends(0.0, _lib.BC_ROBIN, _lib.BC_DIRICHLET, True) == (1, 0)


def test_product_fails_loudly_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fvm_b200.BatchedAdvectionDiffusionModel(np.linspace(0, 1, 9), 1.0, 1.0, 0.1)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        fvm_b200.AdvectionDiffusionModel(np.linspace(0, 1, 9), 1.0, 1.0, 0.1)


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "fvm_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, fn)).read()
                assert "oracle" not in text or fn == "workloads.py", fn


def test_workloads_are_shard_invariant():
    full = W.batched_sweep(64, 32)
    part = W.batched_sweep(64, 32, first=16, count=8)
    for key in ("a", "d", "k", "w0"):
        assert np.array_equal(full[key][16:24], part[key])
    f = W.graded_faces(1024)
    h = np.diff(f)
    assert f[0] == 0 and abs(f[-1] - 1) < 1e-15 and abs(h[0] / h[-1] - 100) < 1e-9


def test_bench_contract_pieces():
    """bench.py without a GPU: argument defaults, the workload description, and the committed ncu
    capture that `roofline.traffic` is read from (within 5 % of the algorithmic bytes of a launch)."""
    import importlib.util
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_under_test", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    argv, sys.argv = sys.argv, ["bench.py"]
    try:
        spec.loader.exec_module(bench)
        args = bench.parse()
    finally:
        sys.argv = argv
    assert (args.gpus, args.impl, args.workload, args.problems, args.cells) == (1, "b200", "cfg3", 65536, 1024)
    assert args.warmup >= 3 and args.steps >= 1 and args.checkpoint_every == 1000 and args.inner is None
    args.inner = 10000
    cfg = bench.config(args)
    assert "workload" in cfg and "model" not in cfg and cfg["theta_steps_per_bench_step"] == 10000
    assert bench.METRIC == "fp64 cell-updates/sec (theta-step)" and bench.UNIT == "cell-updates/s"
    traffic = bench.ncu_traffic(bench.FUSED_SUMMARY)
    algorithmic = (16.0 + 8.0 * 10) * 65536 * 1024
    assert traffic is not None and abs(traffic / algorithmic - 1.0) < 0.05
    assert bench.ncu_traffic("no_such_capture.txt") is None
    peaks, src = bench.measured_peaks()
    assert peaks["hbm_gbs"] > 1000 and "sm_max_mhz" in peaks and src


def test_bench_slab_rows_checker_matches_the_oracle():
    """bench.py's torch restatement of the rows of A and B (the full-size residual check of the
    cfg 5 leg) against the oracle, for every kind of slab: first, middle, last, whole mesh."""
    import importlib.util
    import torch
    from oracle import fvm_oracle as O
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("bench_under_test2", os.path.join(root, "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    J, a, d, dt, theta = 96, 1.0, 3e-3, 2e-3, 0.5
    faces = O.graded_faces(J, 30.0)
    asm = O.assemble(faces, a, d, "exponential", left_flux=0.0, right_flux=0.0, exponential="fitted")
    A, B = O.theta_operators(asm, dt, theta)
    for j0, j1 in ((0, 32), (32, 64), (64, 96), (0, 96)):
        f_lo, f_hi = max(j0 - 1, 0), min(j1 + 1, J)
        f = torch.as_tensor(faces[f_lo:f_hi + 1])
        (Al, Ad, Au), (Bl, Bd, Bu) = bench._slab_rows(f, a, d, dt, theta, J, j0, j1, f_lo)
        for got, want in ((Al, A["lo"]), (Ad, A["di"]), (Au, A["up"]), (Bl, B["lo"]), (Bd, B["di"]), (Bu, B["up"])):
            assert np.allclose(got.numpy(), want[j0:j1], rtol=1e-12, atol=1e-13), (j0, j1)
