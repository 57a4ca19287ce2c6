"""The reference's class surface -- ``Mesh``, ``CellVariable``, ``AdvectionDiffusionModel``
(fvm_classes.py:29-310) -- with the arithmetic of the model in CUDA.

Geometry and the ndarray-like cell variable stay host objects (they are how user scripts
hand data in); everything the reference computes in ``AdvectionDiffusionModel`` -- kappa,
the modified diffusivity, the matrix rows, beta -- comes from ``fvm_assemble`` on the GPU,
and the objects handed back (``dia_matrix``, ``sparse.diags``, ndarray) are the types the
reference returns, so  ``I - k*theta*alpha*M``  and ``spsolve`` in existing scripts keep
working.  ``theta_run`` is the accelerated replacement of the script-level time loop.
"""
from __future__ import annotations

import collections.abc
import warnings

import numpy as np
from scipy import sparse
from scipy.sparse import dia_matrix  # noqa: F401  (re-exported: scripts use `import *`)

from . import _lib
from .batched import BatchedAdvectionDiffusionModel
from .tracked import TrackedDia, linalg  # noqa: F401  (linalg: scipy.sparse.linalg with the spsolve fast path)

# helper shapes of the reference module (fvm_classes.py:10-12), kept for `import *` users
gaussian = lambda z, height, position, hwhm: height * np.exp(-np.log(2) * ((z - position) / hwhm) ** 2)  # noqa: E731
H = lambda z: 0.5 * (1 - np.sign(z))  # noqa: E731
TH = lambda x, sigma, mu: np.where(x > (mu - sigma), 1, 0) * np.where(x < (mu + sigma), 1, 0)  # noqa: E731


def check_index_within_bounds(i, min_i, max_i):
    """True when the index (number or array) lies in [min_i, max_i] (fvm_classes.py:14-26)."""
    i_arr = np.asarray(i)
    ok = (i_arr >= min_i) & (i_arr <= max_i)
    if np.all(ok):
        return True
    if isinstance(i, collections.abc.Iterable):
        print("Index is out of bounds.\ni=%s" % i_arr[np.logical_not(ok)])
    else:
        print("Index is out of bounds.\ni=%s" % i)
    return False


class Mesh(object):
    """1-D cell-centred mesh defined by its faces (fvm_classes.py:29-57)."""

    def __init__(self, faces):
        faces = np.array(faces, dtype=float)
        if faces.ndim != 1:
            raise ValueError("faces must be one-dimensional")
        if np.unique(faces).size != faces.size:                      # :36-37
            raise ValueError("The faces array contains duplicated positions. No cell can have "
                             "zero volume so please update with unique face positions.")
        self.faces = faces
        self.cells = 0.5 * (faces[0:-1] + faces[1:])
        self.J = len(self.cells)
        self.cell_widths = faces[1:] - faces[0:-1]

    def h(self, i):
        """Width of cell i."""
        return self.cell_widths[i]

    def hm(self, i):
        """Distance between centroids in the backwards direction (valid for 1..J-1)."""
        if not check_index_within_bounds(i, 1, self.J - 1):
            raise ValueError("hm index runs out of bounds")
        return self.cells[i] - self.cells[i - 1]

    def hp(self, i):
        """Distance between centroids in the forward direction (valid for 0..J-2)."""
        if not check_index_within_bounds(i, 0, self.J - 2):
            raise ValueError("hp index runs out of bounds")
        return self.cells[i + 1] - self.cells[i]


class CellVariable(np.ndarray):
    """A variable defined at cell centres, with linear interpolation to the faces
    (fvm_classes.py:59-87).  A scalar is broadcast over the mesh."""

    def __new__(cls, input_array, mesh=None):
        if np.ndim(input_array) == 0:
            input_array = input_array * np.ones(len(mesh.cells))
        obj = np.asarray(input_array).view(cls)
        obj.mesh = mesh
        return obj

    def __array_finalize__(self, obj):
        if obj is None:
            return
        self.mesh = getattr(obj, "mesh", None)

    def m(self, i):
        """Value at the left face of cell i."""
        mesh = self.mesh
        return mesh.h(i) / (2 * mesh.hm(i)) * self[i - 1] + mesh.h(i - 1) / (2 * mesh.hm(i)) * self[i]

    def p(self, i):
        """Value at the right face of cell i."""
        mesh = self.mesh
        return mesh.h(i + 1) / (2 * mesh.hp(i)) * self[i] + mesh.h(i) / (2 * mesh.hp(i)) * self[i + 1]


class AdvectionDiffusionModel(object):
    """Drop-in for the reference's model class (fvm_classes.py:89-310).

    ``k`` is the time step (fvm_classes.py:92,351).  Extra keywords, absent from the
    reference: ``reaction`` (decay rate on the diagonal of M), ``strict_bc`` (alpha mask per
    end instead of the reference's both-ends-follow-left_flux), ``device``."""

    def __init__(self, faces, a, d, k, discretisation="central", reaction=None, strict_bc=False,
                 device=None):
        self.mesh = Mesh(faces)
        self.a = CellVariable(a, mesh=self.mesh)
        self.d = CellVariable(d, mesh=self.mesh)
        self.k = k
        self.discretisation = discretisation

        mu = self.peclet_number()                                     # warnings of :103-114
        mu_max = np.max(np.abs(mu))
        if 1.5 <= mu_max < 2.0:
            warnings.warn("\n\nThe Peclet number is %g, this is getting close to the limit of mod 2." % mu_max)
        elif mu_max > 2:
            warnings.warn("\n\nThe Peclet number (%g) has exceeded the maximum value of mod 2 for the "
                          "central discretisation scheme." % (np.max(mu),))
        cfl = self.CFL_condition()
        cfl_max = np.max(np.abs(cfl))
        if 0.5 < cfl_max < 1.0:
            warnings.warn("\n\nThe CFL condition value is %g, it is getting close to the upper limit."
                          % (np.max(cfl),))
        elif cfl_max > 1:
            warnings.warn("\n\nThe CFL condition value is %g, and has gone above the upper limit."
                          % (np.max(cfl),))

        self._gpu = BatchedAdvectionDiffusionModel(
            self.mesh.faces, np.asarray(self.a), np.asarray(self.d), float(k), discretisation, P=1,
            reaction=None if reaction is None else np.asarray(CellVariable(reaction, mesh=self.mesh)),
            strict_bc=strict_bc, device=device)
        # kappa and the modified diffusivity (:116-131) do not depend on the boundary rows;
        # assemble once with placeholder flags to get them from the GPU.
        self._gpu.set_boundary_conditions(left_value=0.0, right_value=0.0)
        asm = self._gpu.assemble()
        self.kappa = asm["kappa"][0].cpu().numpy()
        self.d = CellVariable(asm["D"][0].cpu().numpy(), mesh=self.mesh)
        del self._gpu.left_flux, self._gpu.right_flux, self._gpu.left_value, self._gpu.right_value
        self._gpu._asm = None
        self._gpu._pr = None

    def peclet_number(self):
        return self.a * self.mesh.cell_widths / self.d

    def CFL_condition(self):
        return self.a * self.k / self.mesh.cell_widths

    def set_boundary_conditions(self, left_flux=None, right_flux=None, left_value=None, right_value=None):
        """Make sure this function is used sensibly otherwise the matrix will be ill posed."""
        self.left_flux = left_flux
        self.right_flux = right_flux
        self.left_value = left_value
        self.right_value = right_value
        self._gpu.set_boundary_conditions(left_flux, right_flux, left_value, right_value)
        self._m_host = None
        self._dev_state = None

    def _assembled(self):
        if not hasattr(self, "left_flux"):
            raise AttributeError("'AdvectionDiffusionModel' object has no attribute 'left_flux' "
                                 "(call set_boundary_conditions first)")
        return self._gpu.assemble()

    def _alpha_diagonal(self):
        asm = self._assembled()
        diagonals = np.ones(self.mesh.J)
        diagonals[0] = asm["alpha_left"]
        diagonals[-1] = asm["alpha_right"]
        return diagonals

    def _m_diagonals(self):
        """lo, di, up of M as host arrays (read back once per set_boundary_conditions)."""
        if getattr(self, "_m_host", None) is None:
            asm = self._assembled()
            self._m_host = tuple(asm[n][0].cpu().numpy() for n in ("lo", "di", "up"))
        return self._m_host

    def alpha_matrix(self):
        """Mask that freezes Dirichlet rows; identity for a fully Robin system (:233-242).  A real
        scipy ``dia_matrix`` (of a subclass that remembers which model it belongs to, tracked.py)."""
        return TrackedDia.build(self, "alpha", 0.0, 1.0)

    def beta_vector(self):
        """The Robin boundary-condition vector (:244-263)."""
        asm = self._assembled()
        b = np.zeros(self.mesh.J)
        b[0] = float(asm["beta_left"][0])
        b[-1] = float(asm["beta_right"][0])
        return b

    def coefficient_matrix(self):
        """The J x J tridiagonal matrix M as a ``dia_matrix`` (:265-310), tracked like alpha."""
        return TrackedDia.build(self, "M")

    def _gpu_theta_step(self, w, theta, dt):
        """ONE theta step on the GPU for the spsolve fast path (tracked.py).  The state stays on the
        device between calls: when ``w`` is the array the previous call returned (and nobody
        changed it) nothing is uploaded."""
        torch = _lib.require_cuda()
        g = self._gpu
        if float(g.k[0]) != float(dt):                 # the script's own k in I - k*theta*alpha*M wins
            g.k = g.k.new_full((1,), float(dt))
            g._factors = {}
        st = getattr(self, "_dev_state", None)
        if st is None or not np.array_equal(st[1], w):
            dev = torch.as_tensor(w[None, :], device=g.device)
        else:
            dev = st[0]
        new, _ = g.run(dev, theta, 1)
        out = new[0].cpu().numpy()
        self._dev_state = (new, out.copy())
        return out

    # -------------------------------------------------------------- accelerated time loop
    def theta_run(self, w_init, theta, nsteps, checkpoint_every=0, variant="auto", save=None):
        """``nsteps`` steps of  A w_new = B w + beta  (fvm_classes.py:373-377,408-411) on the
        GPU.  Returns the final state (J,), or (final, checkpoints (n, J)) when
        ``checkpoint_every`` > 0.  ``save``: path of a .npy file that receives the checkpoints
        (the final state when there are none) -- the headless stand-in for the demo's movie
        writer, which grabs a frame per step (fvm_classes.py:406-422)."""
        if not hasattr(self, "left_flux"):
            raise AttributeError("call set_boundary_conditions first")
        w, ck = self._gpu.run(np.asarray(w_init, dtype=float)[None, :], theta, nsteps,
                              checkpoint_every, variant)
        w = w[0].cpu().numpy()
        # checkpoint_every > nsteps: no checkpoint was due, an empty (0, J) array comes back
        snaps = None if not checkpoint_every else (ck[:, 0].cpu().numpy() if ck is not None else np.zeros((0, self.mesh.J)))
        if save is not None:
            np.save(save, snaps if snaps is not None else w[None, :])
        if checkpoint_every:
            return w, snaps
        return w

    def theta_run_until_steady(self, w_init, theta, tol=1e-9, check_every=100, max_steps=1000000):
        """Step until max|w_n - w_{n-check_every}| <= tol max|w| ; returns (w, steps taken)."""
        if not hasattr(self, "left_flux"):
            raise AttributeError("call set_boundary_conditions first")
        w, steps, _ = self._gpu.run_until_steady(np.asarray(w_init, dtype=float)[None, :], theta, tol,
                                                 check_every, max_steps)
        return w[0].cpu().numpy(), steps

    def theta_run_time_varying(self, w_init, theta, left_flux_t=None, right_flux_t=None, checkpoint_every=0):
        """Theta steps with per-step boundary fluxes (not in the reference; see
        BatchedAdvectionDiffusionModel.run_time_varying)."""
        w, ck = self._gpu.run_time_varying(np.asarray(w_init, dtype=float)[None, :], theta, left_flux_t,
                                           right_flux_t, checkpoint_every)
        w = w[0].cpu().numpy()
        if not checkpoint_every:
            return w
        return w, (ck[:, 0].cpu().numpy() if ck is not None else np.zeros((0, self.mesh.J)))


class Model(AdvectionDiffusionModel):
    """The older API that fvm_examples.py:29-45 is written against:
    ``A_matrix() = I - k theta alpha M``, ``M_matrix() = I + k (1-theta) alpha M``,
    ``b_vector() = beta``.  theta is not recoverable from the reference snapshot;
    Crank-Nicolson by default."""

    theta = 0.5

    def A_matrix(self):
        return TrackedDia.build(self, "lin", 1.0, -(self.k * self.theta))

    def M_matrix(self):
        return TrackedDia.build(self, "lin", 1.0, self.k * (1 - self.theta))

    def b_vector(self):
        return self.beta_vector()
