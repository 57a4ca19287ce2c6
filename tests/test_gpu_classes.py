"""The reference's class surface on top of the CUDA path: what a script that does
``from fvm_classes import *`` sees (fvm_classes.py:367-377,408-411; fvm_examples.py:29-45,84-87)."""
import warnings

import numpy as np
import pytest

from conftest import load_golden, step_error
from oracle import fvm_oracle as O

pytestmark = pytest.mark.gpu


def test_demo_script_idiom_with_gpu_assembled_matrices():
    """The shipped demo's own lines, on matrices assembled by the GPU and handed back as
    scipy objects: A = I - k*theta*alpha*M ; d = (I + k*(1-theta)*alpha*M)*w + beta ; spsolve."""
    from fvm_b200 import AdvectionDiffusionModel, sparse, linalg, dia_matrix
    g = load_golden("demo_J149")
    faces, k, theta = g["faces"], 0.01, 1.0
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        model = AdvectionDiffusionModel(faces, 1, 1e-3, k, discretisation="exponential")
        model.set_boundary_conditions(left_value=1., right_value=0.)
        M = model.coefficient_matrix()
        alpha = model.alpha_matrix()
        beta = model.beta_vector()
        assert isinstance(M, dia_matrix) and M.shape == (149, 149)
        I = sparse.identity(model.mesh.J)
        A = I - k * theta * alpha * M
        w = g["w0"].copy()
        for i in range(201):
            d = (I + k * (1 - theta) * alpha * M) * w + beta
            w = linalg.spsolve(A, d)
    assert step_error(w, g["states"][-1], 1.0) < 1e-10
    assert np.allclose(model.kappa, g["kappa"], rtol=0, atol=1e-12)
    assert np.allclose(np.asarray(model.d), g["D"], rtol=1e-13)
    assert np.max(np.abs(M.diagonal() - g["di"])) < 1e-12 * np.max(np.abs(g["di"]))
    # peclet_number() after construction reports the modified Peclet number (<= 2)
    assert np.max(model.peclet_number()) < 2.0
    # and the accelerated loop gives the same answer
    wf = model.theta_run(g["w0"], theta, 201)
    assert step_error(wf, g["states"][-1], 1.0) < 1e-10


def test_old_api_example_flow():
    """fvm_examples.py:29-45,84-87 through the compat class Model (A_matrix/M_matrix/b_vector)."""
    from fvm_b200 import Model, linalg
    g = load_golden("cfg1_examples")
    f = g["uniform_faces"]
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for scheme in O.SCHEMES:
            m = Model(f, 1, 1e-3, 0.01, discretisation=scheme)
            m.set_boundary_conditions(left_value=1., right_value=0.)
            A1, M1, b1 = m.A_matrix(), m.M_matrix(), m.b_vector()
            w1 = np.sin(np.pi * m.mesh.cells) ** 100
            for i in range(251):
                w1 = linalg.spsolve(A1.tocsc(), M1 * w1 + b1)
            want = g["uniform_%s_theta0.5" % scheme][-1]
            assert step_error(w1, want, 1.0) < 1e-10


def test_warnings_and_errors_match_the_reference():
    from fvm_b200 import AdvectionDiffusionModel
    f = np.linspace(0, 1, 50)
    with pytest.warns(UserWarning, match="Peclet number"):
        AdvectionDiffusionModel(f, 1, 1e-3, 1e-4, discretisation="central")        # Pe ~ 20 > 2
    with pytest.warns(UserWarning, match="CFL condition"):
        AdvectionDiffusionModel(f, 1, 1.0, 0.015, discretisation="central")        # CFL ~ 0.73
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        m = AdvectionDiffusionModel(f, 1, 1.0, 1e-4)
        with pytest.raises(AttributeError):
            m.coefficient_matrix()                                                 # BCs never set
        m.set_boundary_conditions()
        with pytest.raises(UnboundLocalError):
            m.beta_vector()                                                        # all None, :260
        with pytest.raises(AttributeError):
            AdvectionDiffusionModel(f, 1, 1.0, 1e-4, discretisation="quick")       # :128,131
        with pytest.raises(ValueError):
            AdvectionDiffusionModel([0, 0.5, 0.5, 1], 1, 1.0, 1e-4)                # duplicate faces
        with pytest.raises(IndexError):
            AdvectionDiffusionModel([0, 0.5, 1], 1, 1.0, 1e-4)                     # J = 2
        m.set_boundary_conditions(left_flux=0.5, right_flux=0.25)
        b = m.beta_vector()
        h = np.diff(f)
        assert b[0] == 0.5 / h[0] and b[-1] == -0.25 / h[-1] and np.all(b[1:-1] == 0)
        assert np.all(m.alpha_matrix().diagonal() == 1)


def test_device_diagnostics_match_numpy():
    """Peclet / CFL extrema (fvm_classes.py:103-114,379-380) and the printed "area" (:420-421)."""
    from fvm_b200 import BatchedAdvectionDiffusionModel, workloads as W
    rng = np.random.default_rng(9)
    J, P = 777, 5
    f = W.graded_faces(J, 30.0)
    h = np.diff(f)
    a = rng.normal(size=(P, J))
    d = 10 ** rng.uniform(-3, 0, size=(P, 1)) * np.ones((P, J))
    k = rng.uniform(1e-4, 1e-2, P)
    w = rng.random((P, J))
    m = BatchedAdvectionDiffusionModel(f, a, d, k, "upwind")
    out = {key: val.cpu().numpy() for key, val in m.diagnostics(w).items()}
    mu = a * h / d
    cfl = a * k[:, None] / h
    assert np.array_equal(out["peclet_min"], mu.min(1)) and np.array_equal(out["peclet_max"], mu.max(1))
    assert np.array_equal(out["cfl_min"], cfl.min(1)) and np.array_equal(out["cfl_max"], cfl.max(1))
    assert np.allclose(out["area"], (w * h).sum(1), rtol=1e-13)


def test_checkpoint_interval_longer_than_the_run_and_divergence_guard():
    """theta_run with checkpoint_every > nsteps returns an empty checkpoint array (not a TypeError);
    run_until_steady raises instead of stepping a NaN state to max_steps."""
    import warnings
    from fvm_b200 import AdvectionDiffusionModel, BatchedAdvectionDiffusionModel
    f = np.linspace(0, 1, 41)
    m = AdvectionDiffusionModel(f, 1.0, 0.05, 1e-3, discretisation="central")
    m.set_boundary_conditions(left_value=1.0, right_value=0.0)
    w0 = np.zeros(40); w0[0] = 1.0
    w, ck = m.theta_run(w0, 0.5, 5, checkpoint_every=10)
    assert ck.shape == (0, 40) and np.all(np.isfinite(w))
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        b = BatchedAdvectionDiffusionModel(f, 1.0, 0.05, 1e-3, "central", P=1)
    b.set_boundary_conditions(left_value=1.0, right_value=0.0)
    bad = np.full((1, 40), np.nan)
    with pytest.raises(FloatingPointError):
        b.run_until_steady(bad, 1.0, check_every=5, max_steps=1000)
