"""TEST INFRASTRUCTURE ONLY -- never imported by the product path (fvm_b200/).

Loads the UNMODIFIED reference library from ``/root/reference/fvm_classes.py`` in
memory so that golden vectors can be generated and the numpy restatement in
``oracle/fvm_oracle.py`` can be pinned against the reference's own code.

The reference is Python 2 (print statements, ``collections.Iterable``), so it
cannot be imported as-is under Python 3.12.  The loader takes lines 1-311 only
(the library; the ``__main__`` movie driver at fvm_classes.py:313-422 needs
matplotlib+ffmpeg) and applies the mechanical fixes SURVEY.md section 8(c) lists:

* the ``print`` statements at fvm_classes.py:22,25,128,132,133,308 get parentheses;
* ``collections.Iterable`` (fvm_classes.py:20) becomes ``collections.abc.Iterable``.

Nothing is written to disk: no reference text enters this repository.
``/root/reference`` exists only in the build container, not on the GPU box, so
only ``tests/golden/make_golden.py`` and the container-only pinning tests use
this module (they skip when the reference is absent).
"""
from __future__ import annotations

import contextlib
import io
import os
import re
import types
import warnings

REFERENCE_PATH = os.environ.get("FVM_REFERENCE_PATH", "/root/reference/fvm_classes.py")
_LIBRARY_LINES = 311  # fvm_classes.py:1-311 is the library; 313-422 is the demo driver

_PRINT_STMT = re.compile(r"^(\s*)print (.*)$")


def reference_available() -> bool:
    return os.path.isfile(REFERENCE_PATH)


def _py3_source() -> str:
    with open(REFERENCE_PATH, "r") as fh:
        lines = fh.read().split("\n")[:_LIBRARY_LINES]
    out = []
    for line in lines:
        m = _PRINT_STMT.match(line)
        if m:
            line = "%sprint(%s)" % (m.group(1), m.group(2))
        line = line.replace("collections.Iterable", "collections.abc.Iterable")
        out.append(line)
    # line 1 is the __future__ import and must stay first; slot the abc import after it
    out.insert(1, "import collections.abc")
    return "\n".join(out) + "\n"


_cached = None


def load_reference() -> types.ModuleType:
    """Return the reference library as a module object (cached)."""
    global _cached
    if _cached is not None:
        return _cached
    if not reference_available():
        raise FileNotFoundError(
            "%s not found: the reference only exists in the build container" % REFERENCE_PATH)
    import numpy as np
    state = np.random.get_state()  # the reference reseeds numpy at import (fvm_classes.py:7)
    mod = types.ModuleType("fvm_classes_reference")
    mod.__file__ = REFERENCE_PATH
    code = compile(_py3_source(), REFERENCE_PATH, "exec")
    with contextlib.redirect_stdout(io.StringIO()):
        exec(code, mod.__dict__)
    np.random.set_state(state)
    _cached = mod
    return mod


@contextlib.contextmanager
def quiet():
    """Silence the reference's prints (kappa array, BC entries) and its warnings."""
    with contextlib.redirect_stdout(io.StringIO()), warnings.catch_warnings():
        warnings.simplefilter("ignore")
        yield


def reference_model(faces, a, d, k, discretisation, bc):
    """Construct the reference model and set BCs; ``bc`` is a dict of
    set_boundary_conditions keywords (fvm_classes.py:141-147)."""
    ref = load_reference()
    with quiet():
        model = ref.AdvectionDiffusionModel(faces, a, d, k, discretisation=discretisation)
        model.set_boundary_conditions(**bc)
    return model


def reference_assembly(faces, a, d, k, discretisation, bc):
    """M (as lo/di/up with M[i,i-1]=lo[i], M[i,i+1]=up[i]), alpha diagonal, beta,
    kappa and the modified diffusivity, straight from the reference classes."""
    import numpy as np
    model = reference_model(faces, a, d, k, discretisation, bc)
    with quiet():
        M = model.coefficient_matrix().toarray()
        alpha = model.alpha_matrix().diagonal()
        beta = model.beta_vector()
    J = model.mesh.J
    lo = np.zeros(J); up = np.zeros(J)
    lo[1:] = np.diag(M, -1)
    up[:-1] = np.diag(M, 1)
    return dict(lo=lo, di=np.diag(M).copy(), up=up, alpha=np.asarray(alpha, dtype=float),
                beta=np.asarray(beta, dtype=float), kappa=np.asarray(model.kappa, dtype=float),
                D=np.asarray(model.d, dtype=float))


def reference_theta_run(faces, a, d, k, discretisation, bc, theta, w_init, nsteps, every=1):
    """The literal time loop of fvm_classes.py:373-377,408-411 on the reference's
    own matrices.  Returns the states after steps every, 2*every, ... (array [n, J])."""
    import numpy as np
    ref = load_reference()
    model = reference_model(faces, a, d, k, discretisation, bc)
    with quiet():
        M = model.coefficient_matrix()
        alpha = model.alpha_matrix()
        beta = model.beta_vector()
        I = ref.sparse.identity(model.mesh.J)
        A = I - k * theta * alpha * M
        w = np.array(w_init, dtype=float)
        out = []
        for i in range(nsteps):
            rhs = (I + k * (1 - theta) * alpha * M) * w + beta
            w = ref.linalg.spsolve(A, rhs)
            if (i + 1) % every == 0:
                out.append(w.copy())
    return np.array(out)
