"""(M)DEIM hyper-reductors (reference deim/__init__.py)."""
from .deim import DiscreteEmpiricalInterpolation, SelectionMatrix
from .mdeim import MatrixDiscreteEmpiricalInterpolation
from .nonlinear import MatrixDiscreteEmpiricalInterpolationNonlinear

__all__ = [
    "DiscreteEmpiricalInterpolation",
    "MatrixDiscreteEmpiricalInterpolation",
    "MatrixDiscreteEmpiricalInterpolationNonlinear",
    "SelectionMatrix",
]
