"""tlfea -- host-side mirror of the reference's element / solver interface over libtlfea_b200.so."""
from ._lib import NewtonParams, TlfeaError, lib  # noqa: F401
from .elements import (  # noqa: F401
    TYPE_3243,
    TYPE_3443,
    TYPE_T10,
    ElementBase,
    GPU_ANCF3243_Data,
    GPU_ANCF3443_Data,
    GPU_FEAT10_Data,
    tet5pt_quadrature,
)
from .solver import SolverBase, SyncedNewtonParams, SyncedNewtonSolver  # noqa: F401
