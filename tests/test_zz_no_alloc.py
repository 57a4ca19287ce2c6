"""fvm_spike_local_static: the variant of fvm_spike_local for callers that cannot allocate.

Some users of this library run the time loop where device memory cannot be requested: the host
application took all of the GPU's memory for its own pool when it started, or the step runs in
a section where cudaMalloc's device-wide synchronisation is not allowed.  For them the one work
array fvm_spike_local needs from its caller (`rs`, two doubles per tile) sits in a fixed 512 KiB device array inside
the library.  This test calls the entry point through the C-ABI, twice in a row on different
states, and requires the same bits as fvm_spike_local produces with a caller-owned `rs`; it
also checks that a request beyond the built-in capacity is refused without launching anything.

The file is named so that pytest collects it last: a run that stops at the first failure has
then already executed every other test.
"""

import numpy as np
import pytest
import torch

from fvm_b200 import PartitionedAdvectionDiffusionModel, _lib, workloads as W

pytestmark = pytest.mark.gpu


def _model(P=3, J=3000):
    f = W.graded_faces(J, 20.0)
    h = np.diff(f)
    rng = np.random.default_rng(42)
    a = rng.uniform(0.5, 2.0, P)
    d = a * h.max() / rng.uniform(0.3, 1.5, P)
    k = rng.uniform(1.0, 6.0, P) * h.min()
    c = 0.5 * (f[1:] + f[:-1])
    w0 = 0.2 + np.exp(-((c[None, :] - rng.uniform(0.3, 0.7, (P, 1))) / 0.1) ** 2)
    m = PartitionedAdvectionDiffusionModel(f, a, d, k, "exponential", P=P, config=1)
    m.set_boundary_conditions(left_flux=0.25, right_flux=-0.1)
    m.prepare(0.5)
    m.set_state(w0)
    return m


def _call(m, static, w_in):
    """One fvm_spike_local / fvm_spike_local_static call on state buffer w_in; returns copies of
    everything the call writes."""
    lib = _lib.lib()
    dev = m.device
    w_out = torch.full_like(w_in, float("nan"))
    rr = torch.full((m.P, m.nt), float("nan"), dtype=torch.float64, device=dev)
    gf = torch.full((m.P,), float("nan"), dtype=torch.float64, device=dev)
    wptr = _lib.ptr
    common = (m.nt, m.P, m.config, _lib.ptr(m.p), _lib.ptr(m.q), _lib.ptr(m.b), _lib.ptr(m.sep),
              _lib.ptr(m.bends), 1, m._last_local, m._theta, wptr(w_in), wptr(w_out), m.w_stride)
    if static:
        st = lib.fvm_spike_local_static(*common, _lib.ptr(rr), m.nt, _lib.ptr(gf), _lib.stream_ptr())
    else:
        rs = torch.full((m.P, m.nt, 2), float("nan"), dtype=torch.float64, device=dev)
        st = lib.fvm_spike_local(*common, _lib.ptr(rs), _lib.ptr(rr), m.nt, _lib.ptr(gf), _lib.stream_ptr())
    assert st == 0, lib.fvm_status_string(st)
    torch.cuda.synchronize()
    return w_out.clone(), rr.clone(), gf.clone()


def test_static_scratch_matches_caller_scratch_twice_in_a_row():
    m = _model()
    lib = _lib.lib()
    assert lib.fvm_spike_static_max_tiles() >= m.P * m.nt
    first_state = m._wbuf[m._cur].clone()
    m.step(7)                                   # a different state for the second call
    second_state = m._wbuf[m._cur].clone()
    for state in (first_state, second_state):   # two calls in a row share the built-in scratch
        want = _call(m, False, state)
        got = _call(m, True, state)
        for g, w in zip(got, want):
            assert torch.isfinite(g).all()
            assert torch.equal(g, w)            # same kernels, same inputs: identical bits
    # and once more on the first state: what the previous call left in the scratch does not leak
    for g, w in zip(_call(m, True, first_state), _call(m, False, first_state)):
        assert torch.equal(g, w)


def test_static_scratch_refuses_sizes_beyond_its_capacity():
    m = _model(P=1, J=600)
    lib = _lib.lib()
    cap = int(lib.fvm_spike_static_max_tiles())
    buf = m._wbuf[0]
    out = torch.zeros_like(buf)
    rr = torch.zeros((1, m.nt), dtype=torch.float64, device=m.device)
    gf = torch.zeros((1,), dtype=torch.float64, device=m.device)
    wptr = _lib.ptr
    before = out.clone()
    st = lib.fvm_spike_local_static(cap + 1, 1, m.config, _lib.ptr(m.p), _lib.ptr(m.q), _lib.ptr(m.b),
                                    _lib.ptr(m.sep), _lib.ptr(m.bends), 1, -1, 0.5, wptr(buf), wptr(out),
                                    m.w_stride, _lib.ptr(rr), m.nt, _lib.ptr(gf), _lib.stream_ptr())
    assert st == _lib.FVM_ERR_SCRATCH_LIMIT
    assert b"built-in scratch" in lib.fvm_status_string(st)
    torch.cuda.synchronize()
    assert torch.equal(out, before)             # nothing was launched
