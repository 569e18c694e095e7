"""Host-side mirror of the reference solver interface (source/ns_solver_base.py): the BC-type
enums, the boundary-condition tuple format, equation coefficients, and the dolfin
NewtonSolver stopping rule -- everything below it runs in libnsb200.so on the GPU.
"""
import math
from enum import Enum, auto

import numpy as np

from .grid import SimplexMesh


class VelocityBCType(Enum):
    """source/ns_solver_base.py:16-23"""
    no_slip = auto()
    no_normal_flux = auto()
    no_tangential_flux = auto()
    constant = auto()
    constant_component = auto()
    function = auto()
    function_component = auto()


class PressureBCType(Enum):
    """source/ns_solver_base.py:26-29"""
    constant = auto()
    function = auto()
    mean_value = auto()


class WeakFormConvectiveTerm(Enum):
    """source/ns_solver_base.py:39-48 (only the standard form is on the GPU path so far)"""
    standard_form = auto()
    rotational_form = auto()
    divergence_form = auto()
    skew_symmetric_form = auto()


class Expression:
    """Stand-in for a ``dolfin.Expression`` boundary/initial value: ``fn(x[n,d], t)`` evaluated
    at the dof points; ``time`` is set to the next time level before every step
    (source/ns_solver_base.py:1033-1104, quirk Q3)."""

    def __init__(self, fn, time=0.0, degree=2):
        self.fn, self.time, self.degree = fn, float(time), degree

    def __call__(self, x):
        return np.asarray(self.fn(x, self.time), dtype=np.float64)


class NewtonNotConverged(RuntimeError):
    pass


def newton_converged(r, r0, atol, rtol):
    """dolfin NewtonSolver, convergence_criterion="residual" (SURVEY.md App. A.5)."""
    return (r0 > 0.0 and r / r0 < rtol) or r < atol


class SolverBase:
    """Common part of the three solver mirrors: BC tuples -> (marker, comp, value) tables,
    coefficient dictionary (source/ns_solver_base.py:722-827, :828-857)."""

    def __init__(self, mesh, boundary_markers, form_convective_term="standard"):
        assert isinstance(mesh, SimplexMesh)
        if isinstance(form_convective_term, str):
            form_convective_term = WeakFormConvectiveTerm[form_convective_term.lower() + "_form"] \
                if not form_convective_term.endswith("_form") else WeakFormConvectiveTerm[form_convective_term]
        if form_convective_term is not WeakFormConvectiveTerm.standard_form:
            raise NotImplementedError("only the standard convective form is implemented on the GPU path")
        self._mesh = mesh
        self._boundary_markers = boundary_markers
        self._space_dim = mesh.dim
        self._velocity_bcs, self._pressure_bcs = [], []
        self._equation_coefficients = dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0,
                                           body_force_term=None)

    # ---- boundary conditions ------------------------------------------------------------------
    def set_boundary_conditions(self, bcs, internal_constraints=None):
        """Same tuple format as source/ns_solver_base.py:722-827:
        (type, boundary_id, value) or (type, boundary_id, component, value)."""
        assert isinstance(bcs, (list, tuple))
        assert internal_constraints is None, "internal constraints are not on the GPU path"
        ids = self._mesh.boundary_ids()
        vel, prs = [], []
        d = self._space_dim
        for bc in bcs:
            assert isinstance(bc, tuple) and len(bc) in (3, 4)
            kind, bid = bc[0], bc[1]
            if isinstance(kind, VelocityBCType):
                assert bid in ids, "unknown boundary id %r" % (bid,)
                if kind is VelocityBCType.no_slip:
                    vel.append((bid, None, (0.0,) * d))
                elif kind is VelocityBCType.constant:
                    assert len(bc[2]) == d
                    vel.append((bid, None, tuple(float(v) for v in bc[2])))
                elif kind is VelocityBCType.function:
                    assert callable(bc[2])
                    vel.append((bid, None, bc[2]))
                elif kind is VelocityBCType.constant_component:
                    assert len(bc) == 4 and 0 <= bc[2] < d
                    vel.append((bid, int(bc[2]), float(bc[3])))
                elif kind is VelocityBCType.function_component:
                    assert len(bc) == 4 and 0 <= bc[2] < d and callable(bc[3])
                    vel.append((bid, int(bc[2]), bc[3]))
                elif kind in (VelocityBCType.no_normal_flux, VelocityBCType.no_tangential_flux):
                    n = self._planar_normal_axis(bid)
                    comps = [n] if kind is VelocityBCType.no_normal_flux else [c for c in range(d) if c != n]
                    vel.extend((bid, c, 0.0) for c in comps)
            elif isinstance(kind, PressureBCType):
                if kind is PressureBCType.mean_value:
                    assert bid is None and isinstance(bc[2], float)
                    self._mean_pressure_value = bc[2]
                else:
                    assert bid in ids, "unknown boundary id %r" % (bid,)
                    prs.append((bid, "p", bc[2] if callable(bc[2]) else float(bc[2])))
            else:
                raise NotImplementedError("traction boundary conditions are not on the GPU path")
        self._velocity_bcs, self._pressure_bcs = vel, prs

    def _planar_normal_axis(self, bid):
        """``boundary_normal`` (source/auxiliary_methods.py:8-53) for axis-aligned planes."""
        m = self._mesh
        X = m.coords[m.bfacets[m.bfacet_marker == bid]].reshape(-1, m.dim)
        spread = X.max(axis=0) - X.min(axis=0)
        axis = int(np.argmin(spread))
        assert spread[axis] < 5.0e-15, "boundary is not an axis-aligned plane"
        return axis

    def _set_time(self, t):
        """source/ns_solver_base.py:1033-1104: every Expression gets time = next_time."""
        for _, _, v in self._velocity_bcs + self._pressure_bcs:
            if isinstance(v, Expression):
                v.time = t

    # ---- coefficients ---------------------------------------------------------------------------
    def set_equation_coefficients(self, coefficients):
        """source/ns_solver_base.py:828-857; values from EquationCoefficientHandler
        (source/auxiliary_classes.py:251-306): convective 1, pressure 1, viscous 1/Re."""
        for key, value in coefficients.items():
            assert key in ("convective_term", "coriolis_term", "euler_term", "pressure_term",
                           "viscous_term", "body_force_term")
            if key in ("coriolis_term", "euler_term"):
                assert value is None, "rotating-frame terms are not on the GPU path"
                continue
            if value is not None:
                assert isinstance(value, float) and math.isfinite(value) and value > 0.0
            self._equation_coefficients[key] = value

    def set_dimensionless_numbers(self, Re=1.0):
        self.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
