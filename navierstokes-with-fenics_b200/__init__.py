"""nsb200 -- B200-native hot path (assembly, Dirichlet rows, Krylov solves, BDF history) behind
the solver classes of LKM-code-base/NavierStokes-with-Fenics.

The directory is named ``navierstokes-with-fenics_b200`` (not an importable identifier); the
repo-root module ``nsb200.py`` loads it under the package name ``nsb200``.  Host code is Python,
PyTorch only provides device buffers / streams, all per-step arithmetic is in libnsb200.so
(csrc/, C-ABI in include/nsb200.h).  No CPU fallback: importing is cheap, creating a solver
without the built library or without a GPU raises.
"""
from . import _lib                                                            # noqa: F401
from .grid import (SimplexMesh, HyperCubeBoundaryMarkers, hyper_cube, hyper_rectangle,   # noqa: F401
                   channel_with_cylinder, DFG_BOUNDARY_IDS)
from .spaces import TaylorHoodSpaces                                          # noqa: F401
from .timestepping import BDFTimeStepping                                     # noqa: F401
from .solver_base import (VelocityBCType, PressureBCType, TractionBCType, WeakFormConvectiveTerm,      # noqa: F401
                          WeakFormViscousTerm, Expression, NewtonNotConverged)


def __getattr__(name):
    # torch-dependent modules are imported lazily so that host-only users stay light
    if name in ("Device", "Pattern"):
        from . import device
        return getattr(device, name)
    if name == "IPCSSolver":
        from .ipcs import IPCSSolver
        return IPCSSolver
    if name == "ImplicitBDFSolver":
        from .bdf import ImplicitBDFSolver
        return ImplicitBDFSolver
    if name == "StationarySolver":
        from .stationary import StationarySolver
        return StationarySolver
    raise AttributeError(name)
