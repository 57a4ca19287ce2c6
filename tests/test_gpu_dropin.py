"""Script-level drop-in: a script in the reference's own idiom -- ``from
fvm_classes import *``, ``A = I - k*theta*alpha*M``, ``w = linalg.spsolve(A, d)`` in a Python loop
[fvm_classes.py:367-377,408-411] and the older ``Model`` / ``A_matrix`` / ``spsolve(A1.tocsc(), M1*w1 + b1)``
form [fvm_examples.py:29-45,84-87] -- runs unchanged (only the movie writer is left out), every
``spsolve`` of the loop is a launch of the CUDA path, and every step matches the UNMODIFIED
reference's output (tests/golden/demo_J149.npz, cfg1_examples.npz) to 1e-10."""
import os
import sys

import numpy as np
import pytest

from conftest import load_golden, step_error
from oracle import fvm_oracle as O

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

DEMO_SCRIPT = """
from fvm_classes import *
faces = np.concatenate([np.linspace(0, 0.99, 50), np.logspace(np.log10(0.991), np.log10(1.0), 100)])
mesh = Mesh(faces)
a = CellVariable(1, mesh=mesh)
d = CellVariable(1e-3, mesh=mesh)
k = 0.01
theta = 1.0
left_value = 1.0
w_init = np.sin(np.pi*mesh.cells)**100
w_init[0] = left_value
model = AdvectionDiffusionModel(faces, a, d, k, discretisation="exponential")
model.set_boundary_conditions(left_value=1., right_value=0.)
M = model.coefficient_matrix()
alpha = model.alpha_matrix()
beta  = model.beta_vector()
I = sparse.identity(model.mesh.J)
A = I - k*theta*alpha*M
d = (I + k*(1-theta)*alpha*M)*w_init + beta
w = w_init
states = []
for i in range(201):
    d = (I + k*(1-theta)*alpha*M)*w + beta
    w = linalg.spsolve(A, d)
    area = np.sum(w * mesh.cell_widths)
    states.append(w)
"""

EXAMPLES_SCRIPT = """
from fvm_classes import *
a = 1
d = 1e-3
k = 0.01
central = Model(faces, a, d, k, discretisation="central")
central.set_boundary_conditions(left_value=1., right_value=0.)
A1 = central.A_matrix()
M1 = central.M_matrix()
b1 = central.b_vector()
upwind = Model(faces, a, d, k, discretisation="upwind")
upwind.set_boundary_conditions(left_value=1., right_value=0.)
A2 = upwind.A_matrix()
M2 = upwind.M_matrix()
b2 = upwind.b_vector()
exponential = Model(faces, a, d, k, discretisation="exponential")
exponential.set_boundary_conditions(left_value=1., right_value=0.)
A3 = exponential.A_matrix()
M3 = exponential.M_matrix()
b3 = exponential.b_vector()
mesh = central.mesh
w_init = np.sin(np.pi*mesh.cells)**100
w1 = w_init
w2 = w_init
w3 = w_init
out = {"central": [], "upwind": [], "exponential": []}
iters = 251
for i in range(iters):
    w1 = linalg.spsolve(A1.tocsc(), M1 * w1 + b1)
    w2 = linalg.spsolve(A2.tocsc(), M2 * w2 + b2)
    w3 = linalg.spsolve(A3.tocsc(), M3 * w3 + b3)
    out["central"].append(w1); out["upwind"].append(w2); out["exponential"].append(w3)
"""


def _run(script, **names):
    import warnings
    from fvm_b200 import _lib
    sys.path.insert(0, ROOT)                       # the only change a user makes: where fvm_classes comes from
    try:
        before = _lib.CALLS.get("fvm_theta_run_fused", 0)
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")        # the reference's own Peclet / CFL warnings
            exec(compile(script, "<script>", "exec"), names)
        return names, _lib.CALLS.get("fvm_theta_run_fused", 0) - before
    finally:
        sys.path.remove(ROOT)


def test_shipped_demo_runs_unchanged_on_the_gpu():
    g = load_golden("demo_J149")
    ns, launches = _run(DEMO_SCRIPT)
    assert launches == 201                         # every spsolve of the loop was one fused-kernel launch
    assert np.array_equal(ns["faces"], g["faces"])
    assert step_error(np.array(ns["states"]), g["states"], 1.0) < 1e-10
    assert abs(ns["area"] - 0.9989593187683591) < 1e-10        # the value the demo prints last
    assert isinstance(ns["A"], ns["dia_matrix"])   # still the scipy type the reference returns


@pytest.mark.parametrize("kind", ["uniform", "random", "geometric"])
def test_example_driver_runs_unchanged_on_the_gpu(kind):
    g = load_golden("cfg1_examples")
    ns, launches = _run(EXAMPLES_SCRIPT, faces=O.example_faces(kind))
    assert launches == 3 * 251
    # the two wild meshes: 2e-10, as in tests/test_gpu_theta.py (SuperLU itself is noisy there)
    tol = 1e-10 if kind == "uniform" else 2e-10
    for scheme in O.SCHEMES:
        err = step_error(np.array(ns["out"][scheme]), g["%s_%s_theta0.5" % (kind, scheme)], np.max(np.abs(g[kind + "_w0"])))
        assert err < tol, (kind, scheme, err)


def test_import_star_surface_and_seed():
    sys.path.insert(0, ROOT)
    try:
        import importlib
        mod = importlib.import_module("fvm_classes")
        importlib.reload(mod)                                     # the import seeds numpy, like the reference's (:7)
        first = np.random.uniform(0, 1, 3)
        assert np.allclose(first, np.random.RandomState(1).uniform(0, 1, 3))
        for name in ("Mesh", "CellVariable", "AdvectionDiffusionModel", "Model", "np", "sparse", "linalg", "dia_matrix",
                     "gaussian", "H", "TH", "check_index_within_bounds"):
            assert hasattr(mod, name), name
    finally:
        sys.path.remove(ROOT)
