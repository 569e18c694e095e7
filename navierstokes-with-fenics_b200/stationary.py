"""StationarySolver: mirror of ``StationarySolverBase`` / ``StationaryNavierStokesSolver``
(source/ns_solver_base.py:860-988, source/ns_solver_base.py:894-949 for the forms): hybrid
Picard -> Newton iteration on the mixed space with dolfin ``SystemAssembler`` semantics
(symmetric local Dirichlet elimination with values x - g, source/auxiliary_classes.py:120-164).
Assembly of F, J_picard and J_newton with the elimination fused in runs in ``k_mixed_rows``;
every update is the K16 direct solve (library leg, timed apart).
"""
import math

from . import _lib as L
from .monolithic import MixedSolverBase


class StationarySolver(MixedSolverBase):
    def __init__(self, mesh, boundary_markers, form_convective_term="standard", tol=1e-10, maxiter=50,
                 tol_picard=1e-2, maxiter_picard=10, device_index=0, reorder=1, linear_solver="auto", linear_rtol=1e-13,
                 form_viscous_term="standard", device=None):
        super().__init__(mesh, boundary_markers, form_convective_term, device_index, reorder, linear_solver, linear_rtol,
                         form_viscous_term=form_viscous_term, device=device)
        assert isinstance(maxiter, int) and maxiter > 0 and isinstance(tol, float) and tol > 0.0
        assert isinstance(maxiter_picard, int) and maxiter_picard > 0 and isinstance(tol_picard, float)
        self._tol, self._maxiter = tol, maxiter
        self._tol_picard, self._maxiter_picard = tol_picard, maxiter_picard
        self._solution = self.dev.zeros(self._nW)
        # dolfin's NewtonSolver starts with error_on_nonconvergence = True; the reference switches it
        # off before the Newton stage and never back on (source/ns_solver_base.py:979)
        self._error_on_nonconvergence = True

    def _system(self, picard, matrix):
        """SystemAssembler(J, F, bcs).assemble(...): b (and A) with the symmetric elimination;
        returns ||b||"""
        self._assemble_system(picard, matrix)
        return self.dev.norm(self._b)

    def _assemble_system(self, picard, matrix):
        flags = L.VEC | L.SYMBC | (L.PICARD if picard else 0) | (L.MAT if matrix else 0)
        self.dev.assemble_mixed(flags, self._solution, b=self._b, A=self._A if matrix else None)
        if self._loadW is not None:
            # the external load enters b_e linearly and the local elimination overwrites only Dirichlet rows:
            # adding the global load on all other rows equals assembling it cell by cell
            self.dev.axpy(1.0, self._loadW, self._b)

    def solve(self):
        """source/ns_solver_base.py:951-988"""
        if not self._bc_ready:
            co = self._equation_coefficients
            self.dev.set_coefficients(1.0, (1.0, -1.0, 0.0), co["convective_term"], co["pressure_term"],
                                      co["viscous_term"])
            self._setup_boundary_conditions()
        self._prepare_forms_and_load()
        if self._loadW is not None:
            self._loadW = self._loadW.clone()
            self._loadW[self._bc_dofs.long()] = 0.0          # setup-time masking of the Dirichlet rows
        residual = self._system(True, False)
        if residual < self._tol_picard:          # :966-970: lower the Picard tolerance below the start residual
            order = math.floor(math.log10(residual))
            self._tol_picard = (residual / 10.0 ** order - 1.0) * 10.0 ** order
        it_p = self._newton(lambda: self._system(True, False), lambda: self._assemble_system(True, True), self._solution,
                            atol=self._tol_picard, rtol=1e-9, maxit=self._maxiter_picard,
                            error_on_nonconvergence=self._error_on_nonconvergence)
        self._error_on_nonconvergence = False
        it_n = self._newton(lambda: self._system(False, False), lambda: self._assemble_system(False, True), self._solution,
                            atol=self._tol, rtol=1e-9, maxit=self._maxiter, error_on_nonconvergence=False)
        self.newton_iterations = [it_p, it_n]
        residual = self._system(False, False)
        assert residual <= self._tol, "Newton iteration did not converge."
        return residual

    @property
    def solution(self):
        return self._solution

    @property
    def velocity(self):
        return self.split(self._solution)[0]

    @property
    def pressure(self):
        return self.split(self._solution)[1]
