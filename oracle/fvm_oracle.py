"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's theta-step hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` /
``--impl reference`` legs of ``bench.py`` may import this module.  The product
path (``fvm_b200/``) never does: it runs sm_100a CUDA or fails loudly.

What is restated (all citations are into ``/root/reference/fvm_classes.py``):

* ``Mesh``                       :32-41   -> :func:`mesh_arrays`
* Peclet / CFL                   :135-139 -> :func:`peclet`, :func:`cfl`
* kappa and modified diffusivity :116-131 -> :func:`kappa_reference`, :func:`kappa_fitted`
* ``CellVariable.m/p``           :81-87   -> :func:`face_interp`
* interior rows                  :149-154 -> :func:`assemble`
* Robin rows / beta entries      :156-189 -> :func:`assemble`
* Dirichlet rows / beta entries  :191-231 -> :func:`assemble`
* ``alpha_matrix``               :233-242 -> :func:`alpha_ends`
* ``beta_vector``                :244-263 -> :func:`assemble`
* ``coefficient_matrix``         :265-310 -> :func:`assemble`
* theta operators                :373-377 -> :func:`theta_operators`
* the time loop                  :408-411 -> :func:`theta_run_scipy` (literal scipy loop)

Third-party arithmetic on the path: ``scipy.sparse.linalg.spsolve`` (SuperLU
``gssv``, COLAMD, partial pivoting).  The reference pins no scipy version
(it has no requirements file); this container has scipy 1.18.1 / numpy 2.3.5.

PINNING: the reference ships no tests, golden vectors or fixtures (SURVEY.md
section 4).  The restatement is therefore pinned against OUTPUTS OF THE REFERENCE
ITSELF, produced in the build container by ``tests/golden/make_golden.py``
through ``oracle/ref_loader.py`` and committed under ``tests/golden/``; the test
``tests/test_oracle_golden.py`` checks assembly bit-for-bit and the time loop
to round-off against those files, and (container only) against the live reference.

The module also holds two helpers that are NOT in the reference and say so:
:func:`kappa_fitted` (the cancellation-safe Bernoulli form the CUDA path uses, so
that the GPU can be checked where the reference's formula returns NaN, Pe >= 709.8)
and :func:`theta_run_thomas` (a no-pivot Thomas loop in numpy ``longdouble``, the
arbiter when SuperLU and the GPU disagree by O(eps * cond)) with its C twin
:func:`theta_run_arbiter` (``oracle/thomas_arbiter.c``: __float128, for 10^4-step runs).
"""
from __future__ import annotations

import numpy as np
from scipy import sparse
from scipy.sparse import linalg as splinalg

SCHEMES = ("central", "upwind", "exponential")


# --------------------------------------------------------------------------- geometry
def mesh_arrays(faces):
    """faces (J+1,) -> cells (J,), cell_widths (J,)   [fvm_classes.py:38-41]."""
    f = np.array(faces, dtype=float)
    if len(f) != len(set(f.tolist())):                       # :36-37
        raise ValueError("The faces array contains duplicated positions.")
    cells = 0.5 * (f[0:-1] + f[1:])
    widths = f[1:] - f[0:-1]
    return f, cells, widths


def _cellvar(x, J):
    """CellVariable broadcast of a scalar [fvm_classes.py:66-70]."""
    x = np.asarray(x, dtype=float)
    if x.ndim == 0:
        return x * np.ones(J)
    if x.shape != (J,):
        raise ValueError("cell variable must be scalar or of length J")
    return x.copy()


def peclet(a, d, h):
    """mu_i = a_i h_i / d_i  [fvm_classes.py:135-136]."""
    return a * h / d


def cfl(a, k, h):
    """a_i k / h_i  [fvm_classes.py:138-139]."""
    return a * k / h


# --------------------------------------------------------------------------- kappa
def kappa_reference(mu, a, scheme):
    """kappa exactly as the reference computes it [fvm_classes.py:116-126].
    NaN for mu >= ~709.8 in the exponential branch (exp overflows, inf/inf)."""
    if scheme == "exponential":
        with np.errstate(all="ignore"):
            kappa = (np.exp(mu) + 1) / (np.exp(mu) - 1) - 2 / mu      # :117
        kappa[np.where(mu == 0.0)] = 0                                # :118
        kappa[np.where(np.isposinf(mu))] = 1                          # :119
        kappa[np.where(np.isneginf(mu))] = -1                         # :120
        return kappa
    if scheme == "upwind":
        return (np.where(a < 0, -1, 0) + np.where(a > 0, 1, 0)).astype(float)  # :122-124
    if scheme == "central":
        return np.zeros(len(mu))                                      # :126
    raise AttributeError("unknown discretisation %r (the reference prints and then fails "
                         "on self.kappa, fvm_classes.py:128,131)" % (scheme,))


def fitted_D(a, d, h):
    """NOT IN THE REFERENCE.  Modified diffusivity of the exponential scheme in the
    cancellation-safe form  D = d * (B(mu) + mu/2),  B(mu) = mu / expm1(mu),  B(0) = 1,
    which equals d + a h kappa / 2 with kappa = coth(mu/2) - 2/mu for every finite mu
    and never overflows.  Evaluated in longdouble and rounded, as an accuracy reference
    for the CUDA path's fp64 evaluation."""
    a = np.asarray(a, dtype=np.longdouble)
    d = np.asarray(d, dtype=np.longdouble)
    h = np.asarray(h, dtype=np.longdouble)
    with np.errstate(all="ignore"):
        mu = a * h / d
        B = np.where(mu == 0, np.longdouble(1), mu / np.expm1(mu))
        D = d * (B + mu / 2)
        D = np.where(np.isinf(mu), np.abs(a) * h / 2, D)
    return np.asarray(D, dtype=float)


def kappa_fitted(mu):
    """NOT IN THE REFERENCE.  kappa = coth(mu/2) - 2/mu without overflow or
    cancellation (longdouble; series below |mu| = 0.5)."""
    mu = np.asarray(mu, dtype=np.longdouble)
    x = mu / 2
    x2 = x * x
    ser = x * (1 / np.longdouble(3) + x2 * (-1 / np.longdouble(45) + x2 * (2 / np.longdouble(945)
          + x2 * (-1 / np.longdouble(4725) + x2 * (2 / np.longdouble(93555)
          + x2 * (-1382 / np.longdouble(638512875) + x2 * (4 / np.longdouble(18243225)
          + x2 * (-3617 / np.longdouble(325641566250)))))))))
    with np.errstate(all="ignore"):
        big = 1 / np.tanh(x) - 1 / x
    out = np.where(np.abs(mu) < 0.5, ser, big)
    out = np.where(np.isposinf(mu), 1, np.where(np.isneginf(mu), -1, out))
    return np.asarray(out, dtype=float)


# --------------------------------------------------------------------------- face interpolation
def face_interp(v, h, dx):
    """Value of cell variable v at interior face f (between cells f and f+1), f = 0..J-2.
    ``CellVariable.p(f)`` and ``CellVariable.m(f+1)`` are the same expression
    [fvm_classes.py:83,87]:  h[f+1]/(2 dx_f) * v[f] + h[f]/(2 dx_f) * v[f+1]."""
    return h[1:] / (2 * dx) * v[:-1] + h[:-1] / (2 * dx) * v[1:]


# --------------------------------------------------------------------------- boundary flags
def resolve_bc(left_flux=None, right_flux=None, left_value=None, right_value=None):
    """Which row type each end gets.  A value beats a flux when both are given
    (later ``if`` wins, fvm_classes.py:277-287).  Returns ('dirichlet'|'robin') x 2;
    raises UnboundLocalError like the reference (:304) when an end has neither."""
    if left_value is not None:
        left = "dirichlet"
    elif left_flux is not None:
        left = "robin"
    else:
        raise UnboundLocalError("left boundary condition not set (fvm_classes.py:260,304)")
    if right_value is not None:
        right = "dirichlet"
    elif right_flux is not None:
        right = "robin"
    else:
        raise UnboundLocalError("right boundary condition not set (fvm_classes.py:260,304)")
    return left, right


def alpha_ends(left_flux=None, right_flux=None, left_value=None, right_value=None, strict_bc=False):
    """End entries of the alpha mask.  The reference keys BOTH ends off ``left_flux``
    [fvm_classes.py:237-238]; ``strict_bc`` (not in the reference) keys each end off
    its own row type."""
    if strict_bc:
        left, right = resolve_bc(left_flux, right_flux, left_value, right_value)
        return (0.0 if left == "dirichlet" else 1.0), (0.0 if right == "dirichlet" else 1.0)
    a1 = 0.0 if left_flux is None else 1.0
    return a1, a1


# --------------------------------------------------------------------------- assembly
def assemble(faces, a, d, scheme="central", left_flux=None, right_flux=None, left_value=None,
             right_value=None, reaction=None, strict_bc=False, exponential="reference"):
    """M = coefficient_matrix() as three diagonals (M[i,i-1]=lo[i], M[i,i]=di[i],
    M[i,i+1]=up[i]; lo[0]=up[J-1]=0), the alpha diagonal, beta, kappa and the modified
    diffusivity D.  Operation order follows the reference so that the result is
    bit-identical to it [fvm_classes.py:116-131,149-177,191-310].

    ``exponential``: 'reference' = literal kappa formula (:117-120);
    'fitted' = :func:`fitted_D` (not in the reference, what the CUDA path evaluates).
    ``reaction`` (not in the reference): -r_i added to di on non-Dirichlet rows.
    """
    f, c, h = mesh_arrays(faces)
    J = len(c)
    if J < 3:
        raise IndexError("the reference needs J >= 3 (fvm_classes.py:292)")
    a = _cellvar(a, J)
    d = _cellvar(d, J)
    with np.errstate(all="ignore"):
        mu = peclet(a, d, h)
    if scheme == "exponential" and exponential == "fitted":
        kappa = kappa_fitted(mu)
        D = fitted_D(a, d, h)
    else:
        kappa = kappa_reference(mu, a, scheme)
        D = d + 0.5 * a * h * kappa                                   # :131
    left, right = resolve_bc(left_flux, right_flux, left_value, right_value)

    dx = c[1:] - c[:-1]                       # hp(f) = hm(f+1) = cells[f+1]-cells[f]  (:51,57)
    af = face_interp(a, h, dx)
    Df = face_interp(D, h, dx)

    lo = np.zeros(J)
    di = np.zeros(J)
    up = np.zeros(J)
    i = np.arange(1, J - 1)
    # interior rows (:151-153); left face of cell i is face i-1, right face is face i
    lo[i] = 1. / h[i] * (af[i - 1] * h[i] / (2 * dx[i - 1]) + Df[i - 1] / dx[i - 1])
    di[i] = 1. / h[i] * (af[i - 1] * h[i - 1] / (2 * dx[i - 1]) - af[i] * h[i + 1] / (2 * dx[i])
                         - Df[i - 1] / dx[i - 1] - Df[i] / dx[i])
    up[i] = 1. / h[i] * (-af[i] * h[i] / (2 * dx[i]) + Df[i] / dx[i])

    beta = np.zeros(J)
    if left == "robin":                                               # :158-159,182
        di[0] = 1. / h[0] * (-af[0] * h[1] / (2 * dx[0]) - Df[0] / dx[0])
        up[0] = 1. / h[0] * (-af[0] * h[0] / (2 * dx[0]) + Df[0] / dx[0])
        beta[0] = left_flux / h[0]
    else:                                                             # :200-201,223
        di[0], up[0] = 0.0, 1.0
    if right == "robin":                                              # :169-170,188
        lo[J - 1] = 1. / h[J - 1] * (af[J - 2] * h[J - 1] / (2 * dx[J - 2]) + Df[J - 2] / dx[J - 2])
        di[J - 1] = 1. / h[J - 1] * (af[J - 2] * h[J - 2] / (2 * dx[J - 2]) - Df[J - 2] / dx[J - 2])
        beta[J - 1] = -right_flux / h[J - 1]
    else:                                                             # :214-215,230
        lo[J - 1], di[J - 1] = 0.0, 1.0

    if reaction is not None:
        r = _cellvar(reaction, J)
        mask = np.ones(J)
        if left == "dirichlet":
            mask[0] = 0
        if right == "dirichlet":
            mask[J - 1] = 0
        di = di - mask * r

    a1, aJ = alpha_ends(left_flux, right_flux, left_value, right_value, strict_bc)
    alpha = np.ones(J)
    alpha[0] = a1
    alpha[-1] = aJ
    return dict(lo=lo, di=di, up=up, alpha=alpha, beta=beta, kappa=kappa, D=D,
                cells=c, h=h, faces=f, left=left, right=right)


def coefficient_matrix(asm):
    """The J x J DIA matrix the reference returns [fvm_classes.py:303,310]."""
    J = len(asm["di"])
    return sparse.dia_matrix(sparse.diags([asm["lo"][1:], asm["di"], asm["up"][:-1]], [-1, 0, 1],
                                          shape=(J, J)))


# --------------------------------------------------------------------------- theta method
def theta_operators(asm, k, theta):
    """Diagonals of  A = I - k*theta*alpha*M  and  B = I + k*(1-theta)*alpha*M
    [fvm_classes.py:376-377], each entry rounded as scipy rounds it:
    ((k*theta)*alpha_i)*M_ij, then 1 -/+ that on the diagonal."""
    al = asm["alpha"]
    sA = (k * theta) * al
    sB = (k * (1 - theta)) * al
    A = dict(lo=-(sA * asm["lo"]), di=1.0 - sA * asm["di"], up=-(sA * asm["up"]))
    B = dict(lo=sB * asm["lo"], di=1.0 + sB * asm["di"], up=sB * asm["up"])
    return A, B


def theta_run_scipy(asm, k, theta, w_init, nsteps, every=1):
    """The literal loop of fvm_classes.py:373-377,408-411 on the restated matrices:
    the RHS operator is rebuilt and SuperLU refactorises on every step, exactly as the
    reference does.  This is the timed CPU baseline (kind='port')."""
    import warnings
    J = len(asm["di"])
    M = coefficient_matrix(asm)
    alpha = sparse.diags(asm["alpha"], 0)
    beta = asm["beta"]
    I = sparse.identity(J)
    A = I - k * theta * alpha * M
    w = np.array(w_init, dtype=float)
    out = []
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")          # SparseEfficiencyWarning, as in the reference
        for i in range(nsteps):
            rhs = (I + k * (1 - theta) * alpha * M) * w + beta
            w = splinalg.spsolve(A, rhs)
            if (i + 1) % every == 0:
                out.append(w.copy())
    return np.array(out)


def theta_run_thomas(A, B, beta, w_init, nsteps, every=1, dtype=np.longdouble):
    """NOT IN THE REFERENCE.  Same recurrence with a no-pivot Thomas solve, vectorised
    over a leading batch axis, in ``dtype`` (longdouble = arbiter; float64 = a fast
    oracle where SuperLU is impractical).  A, B: dicts of (…, J) diagonals."""
    lo = np.asarray(A["lo"], dtype=dtype); di = np.asarray(A["di"], dtype=dtype)
    up = np.asarray(A["up"], dtype=dtype)
    blo = np.asarray(B["lo"], dtype=dtype); bdi = np.asarray(B["di"], dtype=dtype)
    bup = np.asarray(B["up"], dtype=dtype)
    beta = np.asarray(beta, dtype=dtype)
    w = np.array(w_init, dtype=dtype)
    J = w.shape[-1]
    # factor once
    inv = np.empty_like(di); m = np.zeros_like(di)
    inv[..., 0] = 1 / di[..., 0]
    for j in range(1, J):
        m[..., j] = lo[..., j] * inv[..., j - 1]
        inv[..., j] = 1 / (di[..., j] - m[..., j] * up[..., j - 1])
    out = []
    for n in range(nsteps):
        rhs = bdi * w + beta
        rhs[..., 1:] += blo[..., 1:] * w[..., :-1]
        rhs[..., :-1] += bup[..., :-1] * w[..., 1:]
        for j in range(1, J):
            rhs[..., j] -= m[..., j] * rhs[..., j - 1]
        w = np.empty_like(rhs)
        w[..., J - 1] = rhs[..., J - 1] * inv[..., J - 1]
        for j in range(J - 2, -1, -1):
            w[..., j] = (rhs[..., j] - up[..., j] * w[..., j + 1]) * inv[..., j]
        if (n + 1) % every == 0:
            out.append(np.asarray(w, dtype=float))
    return np.array(out)


def theta_run_arbiter(A, B, beta, w_init, nsteps, snaps=None, precision="quad"):
    """NOT IN THE REFERENCE.  The recurrence of :func:`theta_run_thomas` in C (oracle/thomas_arbiter.c,
    built by oracle/Makefile), carried in __float128 ('quad') or long double: the exact solution of
    the reference's fp64 operators, fast enough for 10^4-step runs.  A, B: dicts of (J,) diagonals;
    returns the states after the steps listed in ``snaps`` (default: the last one), shape [n, J]."""
    import ctypes
    import os
    so = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_build", "libthomas_arbiter.so")
    if not os.path.exists(so):
        import subprocess
        subprocess.run(["make", "-C", os.path.dirname(so)[:-len("/_build")], "-s"], check=True)
    lib = ctypes.CDLL(so)
    fn = lib.arbiter_run_quad if precision == "quad" else lib.arbiter_run_longdouble
    dp = ctypes.POINTER(ctypes.c_double)
    ip = ctypes.POINTER(ctypes.c_int64)
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_int64, ctypes.c_int64] + [dp] * 8 + [ctypes.c_int64, ip, dp]
    arrs = [np.ascontiguousarray(x, dtype=np.float64) for x in
            (A["lo"], A["di"], A["up"], B["lo"], B["di"], B["up"], beta, w_init)]
    J = arrs[1].shape[0]
    snaps = np.ascontiguousarray([nsteps] if snaps is None else snaps, dtype=np.int64)
    out = np.zeros((len(snaps), J))
    st = fn(J, int(nsteps), *[a.ctypes.data_as(dp) for a in arrs], len(snaps),
            snaps.ctypes.data_as(ip), out.ctypes.data_as(dp))
    if st != 0:
        raise MemoryError("thomas_arbiter: allocation failed")
    return out


# --------------------------------------------------------------------------- synthetic inputs
def graded_faces(J, ratio=100.0):
    """Smoothly graded mesh of SURVEY.md 8(d): h_i proportional to r^i with total ratio
    ``ratio``; faces = cumsum normalised to [0, 1]."""
    r = ratio ** (-1.0 / (J - 1))
    h = r ** np.arange(J)
    f = np.concatenate([[0.0], np.cumsum(h)])
    return f / f[-1]


def geo_series(n, r):
    """The geometric mesh generator of the example driver [fvm_examples.py:108-122]."""
    total = 1.0
    series = []
    for i in range(n):
        if i:
            total = total - total * r
        series.append(total)
    series = np.array(series)
    norm = series / (np.max(series) - np.min(series))
    series = norm - np.min(norm)
    return np.abs(series - 1)


def example_faces(kind):
    """The three meshes of fvm_examples.py:105-107,124."""
    if kind == "uniform":
        return np.linspace(0, 1, 50)
    if kind == "random":
        rs = np.random.RandomState(1)            # np.random.seed(seed=1) at import (:7), first draw
        return np.concatenate((np.array([0]), np.sort(rs.uniform(0, 1, 48)), np.array([1])))
    if kind == "geometric":
        return geo_series(50, 0.15)
    raise ValueError(kind)


def demo_faces():
    """Mesh of the shipped demo [fvm_classes.py:346] (J = 149)."""
    return np.concatenate([np.linspace(0, 0.99, 50), np.logspace(np.log10(0.991), np.log10(1.0), 100)])
