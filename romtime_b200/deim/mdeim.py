"""Matrix DEIM on B200 -- drop-in for the reference's deim/mdeim.py.

A matrix operator is interpolated through the vector of its non-zeros; ``rows`` / ``cols`` hold the
(row-sorted) sparsity pattern, ``dofs`` the selected (row, col) entries.  ``project_basis`` computes
``vec(V^T A_q V)`` for every collateral mode on the device (rt_project_mat_f64) instead of the
reference's Python loop of sparse-dense products (deim/mdeim.py:153-192)."""
from copy import deepcopy

import numpy as np

from ..conventions import EmpiricalInterpolation
from ..utils import bilinear_to_csr, eliminate_zeros, get_nonzero_entries, vector_to_csr
from .deim import DiscreteEmpiricalInterpolation


def _sorted_topology(Ah):
    """(rows, cols) of the non-zeros of ``Ah`` sorted by row with a STABLE sort, as the reference does
    with ``sorted(zip(rows, cols), key=row)`` (deim/mdeim.py:141-147)."""
    rows, cols, _ = get_nonzero_entries(Ah)
    order = np.argsort(rows, kind="stable")
    return [int(r) for r in rows[order]], [int(c) for c in cols[order]]


class MatrixDiscreteEmpiricalInterpolation(DiscreteEmpiricalInterpolation):

    TYPE = EmpiricalInterpolation.MDEIM

    def __init__(self, assemble, name=None, grid=None, tree_walk_params=None):
        super().__init__(assemble=assemble, name=name, grid=grid, tree_walk_params=tree_walk_params)
        self.rows = None  # matrix topology (COO pattern of the vectorised non-zeros)
        self.cols = None

    def copy(self):
        new = super().copy()
        for attr in ("rows", "cols"):
            value = getattr(self, attr)
            if value is not None:
                setattr(new, attr, deepcopy(value))
        return new

    def setup(self, rnd):
        """Seed + matrix topology from one sampled parameter at t = 1 (reference deim/mdeim.py:79-100)."""
        super().setup(rnd=rnd)
        mu = list(self.build_sampling_space(num=1))[0]
        self.rows, self.cols = self.get_matrix_topology(mu=mu, t=1.0)

    def get_entry(self, idx):
        """(row, col) of entry ``idx`` of the vectorised matrix (reference deim/mdeim.py:102-117)."""
        return self.rows[idx], self.cols[idx]

    def store_dofs(self, dofs):
        self.dofs = [self.get_entry(int(dof)) for dof in dofs]

    def get_matrix_topology(self, mu, t):
        return _sorted_topology(eliminate_zeros(self._assemble_matrix(mu, t)))

    def _assemble_matrix(self, mu, t):
        return bilinear_to_csr(matrix=self.assemble(mu=mu, t=t))

    def assemble_snapshot(self, mu, t):
        """Non-zeros of the assembled operator, CSR order (reference deim/mdeim.py:194-212)."""
        return eliminate_zeros(self._assemble_matrix(mu, t)).data

    def project_basis(self, V):
        """``basis_rom[(i*N_V + j), q] = (V^T A_q V)[i, j]`` with ``A_q = csr(basis_fom[:, q]; rows, cols)``
        (reference deim/mdeim.py:153-192; row-major flattening, :183).  Sets ``N_V``."""
        from .. import kernels

        self.N_V = V.shape[1]
        out = kernels.project_mat(self.basis_fom, self.rows, self.cols, V)
        self.basis_rom = np.asfortranarray(out.cpu().numpy())

    def interpolate(self, mu, t, which=None):
        """Reduced operator as an (N_V, N_V) array for ``which="rom"``, CSR matrix in FOM space otherwise
        (reference deim/mdeim.py:230-261)."""
        approximation = self._interpolate(mu, t, which=which)
        if which == self.ROM:
            return approximation.reshape((self.N_V, self.N_V))
        return vector_to_csr(entries=approximation, rows=self.rows, cols=self.cols)
