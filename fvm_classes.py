"""Drop-in module for scripts written against the reference: ``from fvm_classes import *``
(fvm_examples.py:1) gives ``Mesh``, ``CellVariable``, ``AdvectionDiffusionModel`` (and the older
``Model`` API fvm_examples.py:29-45 uses), ``np``, ``sparse``, ``linalg``, ``dia_matrix`` and the
helper shapes -- the names the reference module exports -- backed by the CUDA path of fvm_b200.
``linalg`` is ``scipy.sparse.linalg`` with an ``spsolve`` that sends the reference's time-loop
idiom  ``w = linalg.spsolve(A, (I + k*(1-theta)*alpha*M)*w + beta)``  to the GPU (fvm_b200/tracked.py).
Python 3; the reference itself is Python 2.
"""
import os as _os
import sys as _sys

_sys.path.insert(0, _os.path.dirname(_os.path.abspath(__file__)))

from fvm_b200.classes import (AdvectionDiffusionModel, CellVariable, H, Mesh, Model, TH,  # noqa: E402,F401
                              check_index_within_bounds, dia_matrix, gaussian, linalg, np, sparse)

np.random.seed(seed=1)      # the reference seeds numpy when it is imported (fvm_classes.py:7); its random example mesh depends on it

__all__ = ["AdvectionDiffusionModel", "CellVariable", "H", "Mesh", "Model", "TH", "check_index_within_bounds",
           "dia_matrix", "gaussian", "linalg", "np", "sparse"]
