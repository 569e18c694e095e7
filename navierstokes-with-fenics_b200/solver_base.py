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


class TractionBCType(Enum):
    """source/ns_solver_base.py:32-37"""
    constant = auto()
    constant_component = auto()
    function = auto()
    function_component = auto()
    free = auto()


class WeakFormConvectiveTerm(Enum):
    """source/ns_solver_base.py:39-48; the enum order is the ``conv`` code of nsb_set_forms"""
    standard_form = auto()
    rotational_form = auto()
    divergence_form = auto()
    skew_symmetric_form = auto()


class WeakFormViscousTerm(Enum):
    """source/ns_solver_base.py:51-56"""
    reduced_form = auto()
    traction_form = auto()


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

    def __init__(self, mesh, boundary_markers, form_convective_term="standard", form_viscous_term="standard"):
        """source/ns_solver_base.py:78-115: ``form_convective_term`` in {"standard", "rotational", "divergence",
        "skew_symmetric"}, ``form_viscous_term`` in {"standard", "reduced", "traction"}"""
        assert isinstance(mesh, SimplexMesh)
        if isinstance(form_convective_term, str):
            form_convective_term = WeakFormConvectiveTerm[form_convective_term.lower() + "_form"] \
                if not form_convective_term.endswith("_form") else WeakFormConvectiveTerm[form_convective_term]
        if isinstance(form_viscous_term, str):
            form_viscous_term = WeakFormViscousTerm.traction_form if form_viscous_term.lower().startswith("traction") \
                else WeakFormViscousTerm.reduced_form
        self._form_convective_term, self._form_viscous_term = form_convective_term, form_viscous_term
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
        vel, prs, trc = [], [], []
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
            elif isinstance(kind, TractionBCType):
                assert bid in ids, "unknown boundary id %r" % (bid,)
                if kind is TractionBCType.free:
                    continue
                if kind in (TractionBCType.constant, TractionBCType.function):
                    assert len(bc) == 3
                    trc.append((bid, None, bc[2] if callable(bc[2]) else tuple(float(v) for v in bc[2])))
                else:
                    assert len(bc) == 4 and 0 <= bc[2] < d
                    trc.append((bid, int(bc[2]), bc[3] if callable(bc[3]) else float(bc[3])))
            else:
                raise TypeError("unknown boundary condition type %r" % (kind,))
        self._velocity_bcs, self._pressure_bcs = vel, prs
        if trc:
            # source/ns_solver_base.py:823-825 (quirk Q12): any traction BC switches the viscous term to its traction form
            self._traction_bcs = trc
            self._form_viscous_term = WeakFormViscousTerm.traction_form
            self._forms_dirty = True

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
        for _, _, v in self._velocity_bcs + self._pressure_bcs + getattr(self, "_traction_bcs", []):
            if isinstance(v, Expression):
                v.time = t
        if isinstance(getattr(self, "_body_force", None), Expression):
            self._body_force.time = t
        if callable(getattr(self, "_angular_velocity", None)):
            self._omega_now = self._angular_velocity(t)
            self._forms_dirty = True

    # ---- body force, Coriolis term, weak-form switches ----------------------------------------------
    def set_body_force(self, body_force):
        """source/ns_solver_base.py:711-720: an Expression (vector valued) or a tuple of floats"""
        assert isinstance(body_force, Expression) or len(body_force) == self._space_dim
        self._body_force = body_force
        self._load_vector = None

    def set_angular_velocity(self, omega):
        """rotating frame (source/ns_solver_base.py:173-191): ``omega`` = Omega_z (2D) or a 3-vector, or a callable
        of time returning one; the Coriolis term 2 c_Omega Omega x u needs ``coriolis_term`` among the coefficients"""
        self._angular_velocity = omega
        self._omega_now = omega(0.0) if callable(omega) else omega
        self._forms_dirty = True

    def _apply_forms(self, dev):
        """hand the weak-form switches to the device (nsb_set_forms) when they changed"""
        if not getattr(self, "_forms_dirty", True):
            return
        cor = None
        c_cor = self._equation_coefficients.get("coriolis_term")
        if getattr(self, "_omega_now", None) is not None and c_cor is not None:
            om = np.atleast_1d(np.asarray(self._omega_now, dtype=np.float64))
            om3 = np.zeros(3)
            if len(om) == 1:
                om3[2] = om[0]
            else:
                om3[:] = om
            cor = 2.0 * c_cor * om3
        dev.set_forms(self._form_convective_term.value - 1, self._form_viscous_term.value - 1, cor)
        self._forms_dirty = False

    def _external_load(self, dev):
        """V-numbered load vector  -c_b int f.w dx + int_Gamma t.w ds  (source/ns_solver_base.py:121-171), or None.
        The body force is integrated on the device (cell-wise P_k interpolant, csrc/project.cu); boundary tractions
        live on O(boundary) facets and are integrated on the host like every other boundary datum."""
        f, trc = getattr(self, "_body_force", None), getattr(self, "_traction_bcs", None)
        if f is None and not trc:
            return None
        timedep = isinstance(f, Expression) or any(isinstance(v, Expression) for _, _, v in (trc or []))
        if getattr(self, "_load_vector", None) is not None and not timedep:
            return self._load_vector
        sp, d = dev.spaces, dev.dim
        load = dev.zeros(sp.ndofs_V)
        if f is not None:
            c_b = self._equation_coefficients["body_force_term"]
            assert c_b is not None, "set_body_force needs the body_force_term coefficient"
            fn = f if isinstance(f, Expression) else (lambda x: np.broadcast_to(np.asarray(f, dtype=np.float64), (len(x), d)))
            rhs = dev.cell_rhs(fn, "V", f.degree if isinstance(f, Expression) else 1)
            dev.axpy(-c_b, rhs, load[:rhs.numel()])
        if trc:
            from .tractions import traction_load
            load += dev.upload(traction_load(sp, trc))
        self._load_vector = load
        return load

    # ---- coefficients ---------------------------------------------------------------------------
    def set_equation_coefficients(self, coefficients):
        """source/ns_solver_base.py:828-857; values from EquationCoefficientHandler
        (source/auxiliary_classes.py:251-306): convective 1, pressure 1, viscous 1/Re."""
        for key, value in coefficients.items():
            assert key in ("convective_term", "coriolis_term", "euler_term", "pressure_term",
                           "viscous_term", "body_force_term")
            if key == "euler_term":
                assert value is None, "the Euler acceleration (time-dependent Omega) is not on the GPU path"
                continue
            if key == "coriolis_term":
                self._forms_dirty = True
            if value is not None:
                assert isinstance(value, float) and math.isfinite(value) and value > 0.0
            self._equation_coefficients[key] = value

    def set_dimensionless_numbers(self, Re=1.0):
        self.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
