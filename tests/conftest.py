import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def bc_from_array(arr):
    """Inverse of make_golden.bc_arrays: NaN means None."""
    keys = ("left_flux", "right_flux", "left_value", "right_value")
    return {k: (None if np.isnan(v) else float(v)) for k, v in zip(keys, arr)}


def step_error(got, want, w0_scale):
    """Per-snapshot error of SURVEY.md 8(d): max_i |dw_i| / max(||w_n||_inf, 1e-6 ||w_0||_inf),
    maximised over the leading (snapshot) axis."""
    got = np.asarray(got, dtype=float)
    want = np.asarray(want, dtype=float)
    scale = np.maximum(np.max(np.abs(want), axis=-1), 1e-6 * w0_scale)
    return float(np.max(np.max(np.abs(got - want), axis=-1) / scale))


@pytest.fixture(scope="session")
def golden_assembly():
    return load_golden("assembly_cases")
