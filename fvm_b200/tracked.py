"""The script-level drop-in: make the reference's driver idiom

    A = I - k*theta*alpha*M                                   [fvm_classes.py:376]
    d = (I + k*(1-theta)*alpha*M)*w + beta ;  w = linalg.spsolve(A, d)        [:410-411]
    w1 = linalg.spsolve(A1.tocsc(), M1 * w1 + b1)             [fvm_examples.py:85]

advance on the GPU without the script changing a line.

How: ``alpha_matrix()``, ``coefficient_matrix()`` (and the old API's ``A_matrix()`` /
``M_matrix()``) return real scipy ``dia_matrix`` objects -- every scipy function accepts them --
of a subclass that also remembers WHAT they are: ``c_I * I + c_M * (alpha M)`` of a given model.
The subclass overrides the arithmetic operators so that the expressions above keep that
knowledge (anything else falls through to scipy and yields a plain matrix), a matrix-vector
product with such an operator is a lazy right-hand side that remembers ``(B, w, beta)``, and
``linalg.spsolve`` (the wrapper exported by this package) recognises the triple: it reads
theta and the time step off the two operators and runs ONE theta step of the model's CUDA
path on ``w`` (state cached on the device between calls).  Whatever it does not recognise is
solved by scipy as before.
"""
from __future__ import annotations

import numpy as np
from scipy import sparse
from scipy.sparse import dia_matrix
from scipy.sparse import linalg as _scipy_linalg

# kinds of tracked matrices
_ALPHA, _M, _LIN = "alpha", "M", "lin"      # c * alpha ; M ; c_I * I + c_M * (alpha M)


class TrackedDia(dia_matrix):
    """A scipy ``dia_matrix`` that knows which operator of which model it is."""

    _fvm = None         # (kind, model, c_I, c_M)

    @classmethod
    def build(cls, model, kind, c_I=0.0, c_M=1.0):
        """Numeric value from the model's GPU-assembled diagonals; tag attached."""
        J = model.mesh.J
        if kind == _ALPHA:
            mat = sparse.diags(c_M * model._alpha_diagonal(), 0)
        else:
            lo, di, up = model._m_diagonals()
            if kind == _LIN:
                al = model._alpha_diagonal()
                lo, di, up = c_M * (al * lo), c_I + c_M * (al * di), c_M * (al * up)
            mat = sparse.diags([lo[1:], di, up[:-1]], [-1, 0, 1], shape=(J, J))
        mat = dia_matrix(mat)
        out = cls((mat.data, mat.offsets), shape=mat.shape)
        out._fvm = (kind, model, float(c_I), float(c_M))
        return out

    # ---- conversions keep the tag (fvm_examples.py:85 calls .tocsc() inside the loop)
    def _tag(self, other):
        try:
            other._fvm = self._fvm
        except AttributeError:
            pass
        return other

    def tocsc(self, copy=False):
        return self._tag(dia_matrix.tocsc(self, copy=copy))

    def tocsr(self, copy=False):
        return self._tag(dia_matrix.tocsr(self, copy=copy))

    def copy(self):
        return self._tag(dia_matrix.copy(self))

    # ---- arithmetic that stays inside the family c_I I + c_M alpha M
    def _plain(self):
        return dia_matrix((self.data, self.offsets), shape=self.shape)

    def __mul__(self, other):
        f = self._fvm
        if f is not None:
            kind, model, cI, cM = f
            if np.isscalar(other):
                return self.__rmul__(other)
            g = getattr(other, "_fvm", None)
            if g is not None and g[1] is model and kind == _ALPHA and g[0] == _M:
                return TrackedDia.build(model, _LIN, 0.0, cM)              # (c alpha) * M
            if kind == _LIN and isinstance(other, np.ndarray) and other.shape == (self.shape[0],):
                return LazyRhs(self, other)                                 # B * w
        return self._plain() * other

    def __rmul__(self, other):
        f = self._fvm
        if f is not None and np.isscalar(other):
            kind, model, cI, cM = f
            if kind == _ALPHA:
                return TrackedDia.build(model, _ALPHA, 0.0, other * cM)
            if kind == _LIN:
                return TrackedDia.build(model, _LIN, other * cI, other * cM)
        return other * self._plain()

    def _identity_like(self, other):
        """True for an untracked J x J identity (what ``sparse.identity(J)`` returns)."""
        if getattr(other, "_fvm", None) is not None or not sparse.issparse(other) or other.shape != self.shape:
            return False
        o = other.todia() if other.format != "dia" else other
        return o.offsets.tolist() == [0] and bool(np.all(o.data[0] == 1.0))

    def _combine(self, other, sign_self, sign_other):
        """sign_self * self + sign_other * other when both are in the family (other may be I)."""
        f = self._fvm
        if f is None or f[0] != _LIN:
            return None
        _, model, cI, cM = f
        g = getattr(other, "_fvm", None)
        if g is not None and g[0] == _LIN and g[1] is model:
            return TrackedDia.build(model, _LIN, sign_self * cI + sign_other * g[2], sign_self * cM + sign_other * g[3])
        if self._identity_like(other):
            return TrackedDia.build(model, _LIN, sign_self * cI + sign_other, sign_self * cM)
        return None

    def __add__(self, other):
        r = self._combine(other, 1.0, 1.0)
        return r if r is not None else self._plain() + other

    def __radd__(self, other):
        r = self._combine(other, 1.0, 1.0)
        return r if r is not None else other + self._plain()

    def __sub__(self, other):
        r = self._combine(other, 1.0, -1.0)
        return r if r is not None else self._plain() - other

    def __rsub__(self, other):
        r = self._combine(other, -1.0, 1.0)
        return r if r is not None else other - self._plain()

    def __neg__(self):
        f = self._fvm
        if f is not None and f[0] == _LIN:
            return TrackedDia.build(f[1], _LIN, -f[2], -f[3])
        return -self._plain()


class LazyRhs(object):
    """``B * w (+ beta)`` that has not been evaluated: the spsolve wrapper turns the pair
    (A, this) into a theta step on ``w``.  Behaves as the (J,) array it stands for wherever
    numpy asks for one."""

    __array_ufunc__ = None          # ndarray + LazyRhs defers to __radd__ instead of converting

    def __init__(self, op, w, extra=None):
        self.op, self.w, self.extra = op, w, extra

    @property
    def shape(self):
        return self.w.shape

    def __len__(self):
        return len(self.w)

    def _value(self):
        v = self.op._plain() * self.w
        return v if self.extra is None else v + self.extra

    def __array__(self, dtype=None, copy=None):
        v = self._value()
        return v if dtype is None else v.astype(dtype)

    def __getitem__(self, idx):
        return self._value()[idx]

    def __add__(self, other):
        if isinstance(other, np.ndarray) and other.shape == self.w.shape:
            return LazyRhs(self.op, self.w, other if self.extra is None else self.extra + other)
        return self._value() + other

    __radd__ = __add__

    def __sub__(self, other):
        return self._value() - other

    def __rsub__(self, other):
        return other - self._value()

    def __mul__(self, other):
        return self._value() * other

    __rmul__ = __mul__


def _recognise(A, rhs):
    """(model, theta, dt, w) when (A, rhs) is the triple of one theta step of a model, else None."""
    fa = getattr(A, "_fvm", None)
    if fa is None or not isinstance(rhs, LazyRhs):
        return None
    fb = rhs.op._fvm
    kind_a, model, aI, aM = fa
    if kind_a != _LIN or fb is None or fb[0] != _LIN or fb[1] is not model:
        return None
    bI, bM = fb[2], fb[3]
    # A = I - (k theta) alpha M ,  B = I + (k (1-theta)) alpha M
    if aI != 1.0 or bI != 1.0 or not (aM < 0.0) or bM < 0.0:
        return None
    beta = model.beta_vector()
    extra = rhs.extra if rhs.extra is not None else np.zeros_like(beta)
    if extra.shape != beta.shape or not np.array_equal(extra, beta):
        return None
    dt = bM - aM
    theta = -aM / dt
    w = np.asarray(rhs.w, dtype=float)
    if w.shape != (model.mesh.J,):
        return None
    return model, theta, dt, w


def spsolve(A, b, *args, **kwargs):
    """``scipy.sparse.linalg.spsolve`` with a fast path: when ``A`` and ``b`` are the operators and
    the right-hand side of ONE theta step of a model of this package [fvm_classes.py:376-377,
    410-411; fvm_examples.py:85], the step runs on the GPU (the model's fused / streaming kernels)
    and the new state comes back as an ndarray.  Everything else goes to scipy unchanged."""
    hit = None if (args or kwargs) else _recognise(A, b)
    if hit is None:
        if isinstance(b, LazyRhs):
            b = np.asarray(b)
        if isinstance(A, TrackedDia):
            A = A._plain()
        return _scipy_linalg.spsolve(A, b, *args, **kwargs)
    model, theta, dt, w = hit
    return model._gpu_theta_step(w, theta, dt)


class _LinalgProxy(object):
    """``scipy.sparse.linalg`` with :func:`spsolve` replaced (scripts say ``linalg.spsolve``)."""

    spsolve = staticmethod(spsolve)

    def __getattr__(self, name):
        return getattr(_scipy_linalg, name)

    def __dir__(self):
        return dir(_scipy_linalg)


linalg = _LinalgProxy()
