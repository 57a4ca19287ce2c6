"""fvm_b200 -- B200-native implementation of the FVM theta-step hot path.

``from fvm_b200 import *`` gives the names a script gets from the reference's
``from fvm_classes import *`` (fvm_examples.py:1), plus the batched GPU model.
"""
from .classes import (AdvectionDiffusionModel, CellVariable, H, Mesh, Model, TH,  # noqa: F401
                      check_index_within_bounds, dia_matrix, gaussian, linalg, np, sparse)
from .batched import BatchedAdvectionDiffusionModel  # noqa: F401
from .partitioned import PartitionedAdvectionDiffusionModel, plan_partition  # noqa: F401

__all__ = ["AdvectionDiffusionModel", "BatchedAdvectionDiffusionModel", "PartitionedAdvectionDiffusionModel", "CellVariable", "H", "Mesh",
           "Model", "TH", "check_index_within_bounds", "dia_matrix", "gaussian", "linalg", "np", "sparse"]
