"""ctypes binding of the C-ABI in include/fvm_b200.h.

There is no CPU fallback: if the library or a CUDA device is missing, the first
call raises.  torch is used for device memory and streams only.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libfvm_b200.so")

FVM_OK = 0
FVM_ERR_INVALID_ARG = -1
FVM_ERR_UNSUPPORTED = -2
FVM_ERR_SCRATCH_LIMIT = -3
FVM_ERR_COLLECTIVE = -4
CENTRAL, UPWIND, EXPONENTIAL = 0, 1, 2
SCHEMES = {"central": CENTRAL, "upwind": UPWIND, "exponential": EXPONENTIAL}
BC_DIRICHLET, BC_ROBIN = 0, 1

c_i64 = ctypes.c_int64
c_i32 = ctypes.c_int32
c_dbl = ctypes.c_double
c_ptr = ctypes.c_void_p


class FvmProblem(ctypes.Structure):
    """struct fvm_problem (include/fvm_b200.h)."""
    _fields_ = [
        ("J", c_i64), ("P", c_i64),
        ("faces", c_ptr),
        ("a", c_ptr), ("a_ps", c_i64), ("a_cs", c_i64),
        ("d", c_ptr), ("d_ps", c_i64), ("d_cs", c_i64),
        ("reaction", c_ptr), ("r_ps", c_i64), ("r_cs", c_i64),
        ("scheme", c_i32), ("left_kind", c_i32), ("right_kind", c_i32),
        ("alpha_left", c_i32), ("alpha_right", c_i32),
        ("left_flux", c_ptr), ("right_flux", c_ptr),
    ]


class FvmSpikePlan(ctypes.Structure):
    """struct fvm_spike_plan (include/fvm_b200.h)."""
    _fields_ = [
        ("n_tiles", c_i64), ("P", c_i64),
        ("config", c_i32), ("lazy", c_i32), ("has_first_cell", c_i32), ("world", c_i32), ("rank", c_i32),
        ("extents", c_i32 * 2),
        ("last_cell_local", c_i64), ("B", c_i64), ("m", c_i64),
        ("theta", c_dbl), ("w_stride", c_i64),
        ("p", c_ptr), ("q", c_ptr), ("b", c_ptr), ("Vt", c_ptr), ("Wt", c_ptr), ("sep", c_ptr), ("bends", c_ptr),
        ("rp", c_ptr), ("rq", c_ptr), ("rinv", c_ptr),
        ("Vr", c_ptr), ("Wr", c_ptr),
        ("tp", c_ptr), ("tq", c_ptr), ("tinv", c_ptr), ("cst", c_ptr),
        ("rs", c_ptr), ("rr", c_ptr), ("g_first", c_ptr),
        ("gr", c_ptr), ("packet", c_ptr), ("allp", c_ptr),
        ("xs", c_ptr),
    ]


class FvmError(RuntimeError):
    def __init__(self, status, where):
        self.status = status
        msg = lib().fvm_status_string(status).decode()
        super().__init__("%s failed with status %d: %s" % (where, status, msg))


_SIGNATURES = {
    "fvm_version": (ctypes.c_int, []),
    "fvm_status_string": (ctypes.c_char_p, [ctypes.c_int]),
    "fvm_assemble": (ctypes.c_int, [ctypes.POINTER(FvmProblem), c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                    c_i64, c_i64, c_ptr, c_ptr, c_ptr]),
    "fvm_theta_factor": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_dbl,
                                        c_i32, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64,
                                        c_ptr, c_ptr, c_ptr]),
    "fvm_theta_step_thomas": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_dbl,
                                             c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr]),
    "fvm_theta_step_thomas_tv": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_dbl,
                                                c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_i64, c_ptr]),
    "fvm_state_compare": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_i64, c_i64,
                                         c_ptr, c_ptr]),
    "fvm_explicit_step": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_i32, c_i32,
                                         c_ptr, c_ptr, c_ptr, c_ptr, c_i64, c_i64, c_ptr]),
    "fvm_mesh_check": (ctypes.c_int, [c_ptr, c_i64, c_ptr, c_ptr]),
    "fvm_diagnostics": (ctypes.c_int, [ctypes.POINTER(FvmProblem), c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_ptr]),
    "fvm_fused_max_cells": (c_i64, []),
    "fvm_theta_run_fused": (ctypes.c_int, [ctypes.POINTER(FvmProblem), c_ptr, c_dbl, c_i64, c_i64,
                                           c_ptr, c_ptr, c_ptr]),
    "fvm_fused_ex_max_cells": (c_i64, []),
    "fvm_theta_run_fused_ex": (ctypes.c_int, [ctypes.POINTER(FvmProblem), c_ptr, c_dbl, c_i64, c_i64, c_ptr, c_ptr,
                                              c_ptr, c_ptr, c_i64, c_i64, c_ptr, c_i64, c_i64, c_ptr, c_i64, c_ptr]),
    "fvm_spike_tile_cells": (c_i64, [c_i32]),
    "fvm_spike_tile_threads": (c_i64, [c_i32]),
    "fvm_spike_setup": (ctypes.c_int, [ctypes.POINTER(FvmProblem), c_ptr, c_dbl, c_i64, c_i64, c_i32,
                                       c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_dbl, c_ptr]),
    "fvm_spike_reduced_factor": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "fvm_spike_local": (ctypes.c_int, [c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i32, c_i64,
                                       c_dbl, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_ptr, c_ptr]),
    "fvm_spike_local_lazy": (ctypes.c_int, [c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i32, c_i64,
                                            c_dbl, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_i64, c_ptr, c_ptr, c_ptr,
                                            c_ptr, c_i64, c_i64, ctypes.POINTER(c_i32), c_ptr]),
    "fvm_spike_rank_pack": (ctypes.c_int, [c_i64, c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr]),
    "fvm_spike_rank_top": (ctypes.c_int, [c_i64, c_i64, c_i64, c_i32, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr,
                                          c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "fvm_spike_static_max_tiles": (c_i64, []),
    "fvm_spike_graph_create": (ctypes.c_int, [ctypes.POINTER(FvmSpikePlan), c_ptr, c_ptr, c_ptr, ctypes.POINTER(c_ptr)]),
    "fvm_spike_graph_run": (ctypes.c_int, [c_ptr, c_i64, c_ptr]),
    "fvm_spike_graph_destroy": (None, [c_ptr]),
    "fvm_spike_steps": (ctypes.c_int, [ctypes.POINTER(FvmSpikePlan), c_ptr, c_ptr, c_i32, c_i32, c_i64, c_ptr, c_ptr,
                                       c_ptr]),
    "fvm_spike_local_static": (ctypes.c_int, [c_i64, c_i64, c_i32, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_i32, c_i64,
                                              c_dbl, c_ptr, c_ptr, c_i64, c_ptr, c_i64, c_ptr, c_ptr]),
    "fvm_spike_reduce_solve": (ctypes.c_int, [c_i64, c_i64, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr, c_ptr]),
    "fvm_spike_finalize": (ctypes.c_int, [c_i64, c_i64, c_i32, ctypes.POINTER(c_i32), c_i64, c_i64, c_ptr, c_ptr,
                                          c_ptr, c_ptr, c_i64, c_ptr]),
}

_lib = None


def lib():
    """The loaded shared library (raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                "%s is missing: build it with `python -m fvm_b200.build` (needs nvcc). "
                "fvm_b200 has no CPU fallback." % LIB_PATH)
        handle = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(handle, name)          # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


def exported_symbols():
    return sorted(_SIGNATURES)


def require_cuda():
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("fvm_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    return torch


def ptr(t):
    """Device pointer of a torch tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda or t.dtype.is_floating_point and t.element_size() != 8:
        raise TypeError("expected a CUDA float64 tensor")
    return ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


CALLS = {}      # entry point name -> number of calls that went through check() (tests count launches with it)


def check(status, where):
    CALLS[where] = CALLS.get(where, 0) + 1
    if status != FVM_OK:
        raise FvmError(status, where)


def bind_to_gpu_numa_node(device_index):
    """Pin the calling process to the CPUs of the NUMA node the GPU hangs off, so that host buffers
    allocated (and pinned) afterwards land in the memory next to that GPU.  With one process per GPU
    on a two-socket box, half of the ranks otherwise stream their checkpoints across the socket
    interconnect.  Linux only; returns the node number, or None when it cannot be determined (then
    nothing is changed)."""
    try:
        torch = require_cuda()
        props = torch.cuda.get_device_properties(int(device_index))
        bus = "%04x:%02x:%02x.0" % (props.pci_domain_id, props.pci_bus_id, props.pci_device_id)
        with open("/sys/bus/pci/devices/%s/numa_node" % bus) as fh:
            node = int(fh.read().strip())
        if node < 0:
            return None
        with open("/sys/devices/system/node/node%d/cpulist" % node) as fh:
            cpus = set()
            for part in fh.read().strip().split(","):
                lo, _, hi = part.partition("-")
                cpus.update(range(int(lo), int(hi or lo) + 1))
        allowed = cpus & set(os.sched_getaffinity(0))
        if not allowed:
            return None
        os.sched_setaffinity(0, allowed)
        return node
    except (OSError, ValueError, AttributeError, RuntimeError):
        return None
