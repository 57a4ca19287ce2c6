"""SURVEY.md 8(f) extensions around the path: time-varying boundary fluxes, distances between
states (steady-state detection, analytic comparison), .npy snapshots.  None of these exist in the
reference; the oracle is the reference's own step (fvm_classes.py:408-411) with beta rebuilt
per step."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from conftest import step_error
from oracle import fvm_oracle as O
from fvm_b200 import AdvectionDiffusionModel, BatchedAdvectionDiffusionModel, workloads as W

pytestmark = pytest.mark.gpu
TOL = 1e-10


def oracle_time_varying(faces, a, d, k, scheme, bc_of_step, theta, w0, nsteps):
    """w <- spsolve(A, B w + beta_n), beta_n assembled from the fluxes of step n."""
    w = np.array(w0, dtype=float)
    out = []
    for n in range(nsteps):
        asm = O.assemble(faces, a, d, scheme, **bc_of_step(n))
        A, B = O.theta_operators(asm, k, theta)
        J = len(w)
        mat = lambda D: sp.diags([D["lo"][1:], D["di"], D["up"][:-1]], [-1, 0, 1], shape=(J, J), format="csc")
        w = spla.spsolve(mat(A), mat(B) @ w + asm["beta"])
        out.append(w.copy())
    return np.array(out)


@pytest.mark.parametrize("variant,J", [("fused", 300), ("fused", 1000), ("fused", 61), ("thomas", 300)])
@pytest.mark.parametrize("theta", [0.5, 1.0, 0.7])
def test_time_varying_fluxes_match_per_step_oracle(theta, variant, J):
    P, n = 3, 24
    f = W.graded_faces(J)
    rng = np.random.default_rng(11)
    a, d, k = rng.uniform(0.5, 1.5, P), rng.uniform(0.01, 0.05, P), rng.uniform(1e-3, 4e-3, P)
    cells = 0.5 * (f[1:] + f[:-1])
    w0 = np.stack([np.exp(-((cells - 0.5) / 0.1) ** 2)] * P)
    lf = 0.3 * np.sin(np.arange(n)[:, None] * 0.4 + np.arange(P)[None, :])      # (n, P)
    rf = 0.2 * np.cos(np.arange(n) * 0.25)                                       # (n,)
    m = BatchedAdvectionDiffusionModel(f, a, d, k, "upwind")
    m.set_boundary_conditions(left_flux=0.7, right_flux=-0.1)
    # both ends vary
    w, ck = m.run_time_varying(w0, theta, left_flux_t=lf, right_flux_t=rf, checkpoint_every=1, variant=variant)
    assert ck.shape == (n, P, J)
    for p in range(P):
        want = oracle_time_varying(f, a[p], d[p], k[p], "upwind",
                                   lambda s: dict(left_flux=lf[s, p], right_flux=rf[s]), theta, w0[p], n)
        assert step_error(ck[:, p].cpu().numpy(), want, 1.0) < TOL, p
    assert np.array_equal(w.cpu().numpy(), ck[-1].cpu().numpy())
    # only the right end varies: the left keeps the flux of set_boundary_conditions
    w2, _ = m.run_time_varying(w0, theta, right_flux_t=rf, variant=variant)
    want = oracle_time_varying(f, a[0], d[0], k[0], "upwind",
                               lambda s: dict(left_flux=0.7, right_flux=rf[s]), theta, w0[0], n)
    assert step_error(w2[0].cpu().numpy()[None], want[-1][None], 1.0) < TOL
    # a constant history reproduces run()
    w3, _ = m.run_time_varying(w0, theta, left_flux_t=np.full(n, 0.7), right_flux_t=np.full(n, -0.1), variant=variant)
    w4, _ = m.run(w0, theta, n, variant=variant)
    assert step_error(w3.cpu().numpy(), w4.cpu().numpy(), 1.0) < 1e-12      # unit flux x 0.7 against flux 0.7: a rounding apart


def test_time_varying_with_a_dirichlet_end_and_bad_arguments():
    J, n = 120, 10
    f = np.linspace(0, 1, J + 1)
    m = BatchedAdvectionDiffusionModel(f, 1.0, 0.02, 2e-3, "central", P=1)
    m.set_boundary_conditions(left_value=1.0, right_flux=0.0)
    w0 = np.zeros((1, J)); w0[0, 0] = 1.0
    rf = np.linspace(0.0, 0.5, n)
    w, _ = m.run_time_varying(w0, 0.5, right_flux_t=rf)
    want = oracle_time_varying(f, 1.0, 0.02, 2e-3, "central",
                               lambda s: dict(left_value=1.0, right_flux=rf[s]), 0.5, w0[0], n)
    assert step_error(w.cpu().numpy(), want[-1][None], 1.0) < TOL
    with pytest.raises(ValueError):
        m.run_time_varying(w0, 0.5, left_flux_t=rf)            # the left end is Dirichlet
    with pytest.raises(ValueError):
        m.run_time_varying(w0, 0.5)                            # nothing varies
    with pytest.raises(ValueError):
        m.run_time_varying(w0, 0.0, right_flux_t=rf)           # theta = 0 is not covered


def test_state_compare_matches_numpy_and_carries_nan():
    J, P = 1000, 5
    f = W.graded_faces(J)
    rng = np.random.default_rng(3)
    a_, b_ = rng.normal(size=(P, J)), rng.normal(size=(P, J))
    m = BatchedAdvectionDiffusionModel(f, 1.0, 1e-2, 1e-3, "central", P=P)
    c = {k: v.cpu().numpy() for k, v in m.compare(a_, b_).items()}
    h = np.diff(f)
    assert np.array_equal(c["max_abs_diff"], np.max(np.abs(a_ - b_), axis=1))
    assert np.array_equal(c["max_abs_ref"], np.max(np.abs(b_), axis=1))
    assert np.allclose(c["l1_diff"], np.sum(np.abs(a_ - b_) * h, axis=1), rtol=1e-13)
    assert np.allclose(c["area_ref"], np.sum(b_ * h, axis=1), rtol=1e-12, atol=1e-15)
    a_[2, 777] = np.nan
    c = m.compare(a_, b_)["max_abs_diff"].cpu().numpy()
    assert np.isnan(c[2]) and not np.isnan(c[[0, 1, 3, 4]]).any()


def test_run_until_steady_reaches_the_analytic_steady_state(tmp_path):
    """Advection-diffusion between Dirichlet ends has the steady profile
    (exp(Pe x) - 1) / (exp(Pe) - 1) with Pe = a/d (the long-time limit of the analytic curve the
    example driver plots, fvm_examples.py:71); the discrete steady state is second-order close
    to it on a uniform mesh."""
    J, a, d = 400, 1.0, 0.1
    f = np.linspace(0, 1, J + 1)
    model = AdvectionDiffusionModel(f, a, d, 5e-3, discretisation="central")
    model.set_boundary_conditions(left_value=0.0, right_value=1.0)
    w0 = np.zeros(J); w0[-1] = 1.0
    w, steps = model.theta_run_until_steady(w0, 1.0, tol=1e-12, check_every=200, max_steps=100000)
    assert 200 <= steps < 100000
    # the Dirichlet cells hold their flag values at the cell CENTRES: the profile lives
    # on [x_0, x_{J-1}]
    x = model.mesh.cells
    L = x[-1] - x[0]
    exact = np.expm1(a / d * (x - x[0])) / np.expm1(a / d * L)
    cmp = model._gpu.compare(w[None], exact[None])
    assert float(cmp["max_abs_diff"][0]) < 1e-4                # second order in h (1.9e-5 for this mesh)
    # a further block of steps does not move it
    w_more = model.theta_run(w, 1.0, 200)
    assert np.max(np.abs(w_more - w)) < 1e-11
    # headless snapshots
    path = tmp_path / "snaps.npy"
    _, ck = model.theta_run(w0, 1.0, 40, checkpoint_every=10, save=str(path))
    assert np.array_equal(np.load(path), ck) and ck.shape == (4, J)


@pytest.mark.parametrize("variant,J", [("fused", 200), ("fused", 1024), ("thomas", 200)])
def test_variable_time_step_matches_per_step_oracle(variant, J):
    """A time step that changes during the run (per problem, too): re-factored where it changes --
    inside the kernel for the on-chip variant."""
    P, n = 2, 18
    f = W.graded_faces(J)
    a, d = np.array([1.0, 0.6]), np.array([0.02, 0.05])
    cells = 0.5 * (f[1:] + f[:-1])
    w0 = np.stack([np.exp(-((cells - 0.4) / 0.1) ** 2)] * P)
    dt = np.empty((n, P))
    dt[:6] = [2e-3, 1e-3]; dt[6:7] = [5e-3, 1e-3]; dt[7:] = [1e-3, 4e-3]      # runs of 6, 1 and 11 steps
    bc = dict(left_flux=0.2, right_flux=0.0)
    m = BatchedAdvectionDiffusionModel(f, a, d, 1.0, "exponential", P=P)      # the constructor's k is not used
    m.set_boundary_conditions(**bc)
    w, ck = m.run_variable_dt(w0, 0.5, dt, checkpoint_every=4, variant=variant)
    assert ck.shape == (4, P, J)
    for p in range(P):
        asm = O.assemble(f, a[p], d[p], "exponential", **bc)
        x, snaps = w0[p].copy(), []
        for s in range(n):
            x = O.theta_run_scipy(asm, dt[s, p], 0.5, x, 1, every=1)[0]
            snaps.append(x)
        snaps = np.array(snaps)
        assert step_error(ck[:, p].cpu().numpy(), snaps[3::4][:4], 1.0) < TOL, p
        assert step_error(w[p].cpu().numpy()[None], snaps[-1][None], 1.0) < TOL, p
    # the model's own time step and factors are untouched afterwards
    assert float(m.k[0]) == 1.0
    # a constant history is run()
    w1, _ = m.run_variable_dt(w0, 1.0, np.full(n, 1.0), variant=variant)
    w2, _ = m.run(w0, 1.0, n, variant=variant)
    assert np.array_equal(w1.cpu().numpy(), w2.cpu().numpy())
