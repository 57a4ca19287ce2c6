"""The operator tracking behind the script-level drop-in (fvm_b200/tracked.py), without a GPU: a
stand-in model supplies the diagonals; the tests check that the reference's driver expressions
[fvm_classes.py:373-377,410-411; fvm_examples.py:29-45,85] keep their tags, have the same
numbers as plain scipy arithmetic, are recognised by the spsolve wrapper, and that anything
else falls through to scipy."""
import numpy as np
from scipy import sparse
from scipy.sparse import linalg as scipy_linalg

from fvm_b200 import tracked as T


class FakeModel(object):
    def __init__(self, J=12, alpha_ends=(0.0, 0.0), beta_ends=(0.0, 0.0)):
        rng = np.random.default_rng(J)
        self.mesh = type("M", (), {"J": J})()
        self.lo, self.up = rng.uniform(0.5, 1.5, J), rng.uniform(0.5, 1.5, J)
        self.lo[0] = self.up[-1] = 0.0
        self.di = -(self.lo + self.up) - 0.1
        self.al = np.ones(J); self.al[0], self.al[-1] = alpha_ends
        self.beta = np.zeros(J); self.beta[0], self.beta[-1] = beta_ends
        self.calls = []

    def _alpha_diagonal(self): return self.al
    def _m_diagonals(self): return self.lo, self.di, self.up
    def beta_vector(self): return self.beta.copy()
    def alpha_matrix(self): return T.TrackedDia.build(self, "alpha", 0.0, 1.0)
    def coefficient_matrix(self): return T.TrackedDia.build(self, "M")

    def _gpu_theta_step(self, w, theta, dt):
        self.calls.append((theta, dt))
        J = self.mesh.J                                          # the same step, on the CPU
        M = sparse.diags([self.lo[1:], self.di, self.up[:-1]], [-1, 0, 1])
        al, I = sparse.diags(self.al, 0), sparse.identity(J)
        return scipy_linalg.spsolve((I - dt * theta * al * M).tocsc(), (I + dt * (1 - theta) * al * M) * w + self.beta)


def plain(m):
    return sparse.diags([m.lo[1:], m.di, m.up[:-1]], [-1, 0, 1]), sparse.diags(m.al, 0)


def test_reference_loop_is_recognised_and_matches_scipy():
    for theta, ends, beta in ((0.5, (1.0, 1.0), (0.3, -0.2)), (1.0, (0.0, 0.0), (0.0, 0.0)), (0.3, (1.0, 1.0), (0.0, 0.0))):
        m = FakeModel(alpha_ends=ends, beta_ends=beta)
        k = 0.01
        M, alpha, beta_v, I = m.coefficient_matrix(), m.alpha_matrix(), m.beta_vector(), sparse.identity(m.mesh.J)
        A = I - k * theta * alpha * M                                     # fvm_classes.py:376
        Mp, ap = plain(m)
        assert sparse.issparse(A) and isinstance(A, sparse.dia_matrix)
        assert np.array_equal(A.toarray(), (I - k * theta * ap * Mp).toarray())
        w = np.linspace(0.0, 1.0, m.mesh.J) ** 2
        w_ref = w.copy()
        for _ in range(3):
            d = (I + k * (1 - theta) * alpha * M) * w + beta_v                # :410
            w = T.linalg.spsolve(A, d)                                        # :411
            w_ref = scipy_linalg.spsolve((I - k * theta * ap * Mp).tocsc(), (I + k * (1 - theta) * ap * Mp) * w_ref + beta_v)
        assert len(m.calls) == 3
        th, dt = m.calls[0]
        assert abs(th - theta) < 1e-15 and abs(dt - k) < 1e-17
        assert np.allclose(w, w_ref, rtol=0, atol=1e-13)


def test_old_api_loop_with_tocsc_is_recognised():
    m = FakeModel(beta_ends=(0.1, 0.0), alpha_ends=(1.0, 1.0))
    k, theta = 0.02, 0.5
    A1 = T.TrackedDia.build(m, "lin", 1.0, -(k * theta))
    M1 = T.TrackedDia.build(m, "lin", 1.0, k * (1 - theta))
    b1 = m.beta_vector()
    w1 = np.ones(m.mesh.J)
    w1 = T.linalg.spsolve(A1.tocsc(), M1 * w1 + b1)                          # fvm_examples.py:85
    assert m.calls == [(0.5, k)]
    assert np.all(np.isfinite(w1))


def test_everything_else_falls_through_to_scipy():
    m = FakeModel(alpha_ends=(1.0, 1.0))
    M, alpha, I = m.coefficient_matrix(), m.alpha_matrix(), sparse.identity(m.mesh.J)
    Mp, ap = plain(m)
    A = I - 0.01 * alpha * M
    w = np.arange(m.mesh.J, dtype=float)
    # a right-hand side that is not B*w + beta of the same model
    x = T.linalg.spsolve(A, w)
    assert not m.calls and np.allclose(x, scipy_linalg.spsolve((I - 0.01 * ap * Mp).tocsc(), w))
    # a beta that is not the model's
    rhs = (I + 0.01 * alpha * M) * w + np.ones(m.mesh.J)
    x = T.linalg.spsolve(A, rhs)
    assert not m.calls and np.allclose(x, scipy_linalg.spsolve((I - 0.01 * ap * Mp).tocsc(), (I + 0.01 * ap * Mp) * w + 1.0))
    # untracked algebra keeps scipy's numbers
    assert np.array_equal((M * M).toarray(), (Mp * Mp).toarray())
    assert np.array_equal((M + M).toarray(), (Mp + Mp).toarray())
    assert np.array_equal((3.0 * M).toarray(), (3.0 * Mp).toarray())
    assert np.array_equal((M * w), Mp * w)
    assert np.array_equal(np.asarray((I + 0.5 * alpha * M) * w + 1.0), (I + 0.5 * ap * Mp) * w + 1.0)
    assert np.array_equal(1.0 + np.asarray((I + 0.5 * alpha * M) * w), 1.0 + (I + 0.5 * ap * Mp) * w)
    # the other names of scipy.sparse.linalg are still there
    assert T.linalg.splu is scipy_linalg.splu
