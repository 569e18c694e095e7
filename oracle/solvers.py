"""Oracle solvers: dolfin NewtonSolver semantics + sparse direct LU, and the step
drivers of the reference solver classes.

* ``newton``            dolfin ``NewtonSolver``
* ``IPCSOracle``        ``IPCSSolver``  source/ns_ipcs_solver.py:35-43,106-247
* ``BDFOracle``         ``ImplicitBDFSolver``  source/ns_bdf_solver.py:36-126
* ``StationaryOracle``  ``StationarySolverBase``  source/ns_solver_base.py:894-988
                        (+ ``CustomNonlinearProblem`` source/auxiliary_classes.py:120-164)
* ``project_V/Q``       ``dolfin.project``  source/ns_solver_base.py:1151,1168

LU = ``scipy.sparse.linalg.splu`` (SuperLU) standing in for PETSc/UMFPACK, freshly
factorised on every solve like the reference.

Singular all-Neumann pressure problems (cavity): the reference leaves
them to LU pivoting, which fixes the pressure only up to an arbitrary constant.  The
oracle (and the CUDA path) pick the representative whose dof vector has zero arithmetic
mean; parity of p is "up to a constant" by construction.

Test infrastructure only (see ``oracle/__init__.py``).
"""
import math
import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from . import bc as obc
from .dofmap import Spaces
from .forms import Assembler


def lu_solve(A, b):
    return spla.splu(A.tocsc()).solve(b)


def solve_singular_mean_zero(A, b):
    """Solve A x = b for a symmetric A whose null space is the constants; return the
    solution with zero arithmetic mean (bordered system)."""
    n = A.shape[0]
    one = sp.csr_matrix(np.ones((n, 1)))
    Ab = sp.bmat([[A, one], [one.T, None]], format="csc")
    x = spla.splu(Ab).solve(np.concatenate([b, [0.0]]))
    return x[:n]


class NewtonNotConverged(RuntimeError):
    pass


def newton(F, J, x, atol=1e-10, rtol=1e-9, maxit=50, error_on_nonconvergence=True, solve=lu_solve):
    """dolfin NewtonSolver, convergence_criterion="residual", relaxation 1.
    F(x)->b, J(x)->A.  Mutates x.  Returns (iterations, residual history)."""
    b = F(x)
    r0 = r = float(np.linalg.norm(b))
    hist = [r]
    it = 0
    converged = (r / r0 < rtol if r0 > 0 else False) or r < atol
    while not converged and it < maxit:
        A = J(x)
        dx = solve(A, b)
        x -= dx
        b = F(x)
        r = float(np.linalg.norm(b))
        hist.append(r)
        it += 1
        converged = (r / r0 < rtol) or (r < atol)
    if not converged and error_on_nonconvergence:
        raise NewtonNotConverged("Newton solver did not converge (%d its, r=%g)" % (it, r))
    return it, hist


def project_V(asm, f):
    """L2 projection of f(x)->[n,d] onto V after nodal P2 interpolation (App. A.7 for degree-2 data)."""
    s = asm.sp
    g = np.asarray(f(s.node_coords()), dtype=np.float64).reshape(-1)
    M = asm.velocity_mass_matrix()
    return lu_solve(M, M @ g)


def project_Q(asm, f):
    s = asm.sp
    g = np.asarray(f(s.mesh.coords), dtype=np.float64).reshape(-1)
    M = asm.pressure_mass_matrix()
    return lu_solve(M, M @ g)


def _initial_fields(asm, velocity, pressure, degree):
    from . import projection
    d = asm.sp.dim
    vel = velocity if callable(velocity) else (lambda x: np.broadcast_to(np.asarray(velocity, float), (len(x), d)))
    u0 = project_V(asm, vel) if degree is None else projection.project(asm, vel, "V", degree)
    p0 = None
    if pressure is not None:
        prs = pressure if callable(pressure) else (lambda x: np.full(len(x), float(pressure)))
        p0 = project_Q(asm, prs) if degree is None else projection.project(asm, prs, "Q", degree)
    return u0, p0


class _Base:
    def __init__(self, mesh, Re=None, coefficients=None, assembler="numpy", threads=None):
        self.mesh = mesh
        self.sp = Spaces(mesh)
        if assembler == "c":      # plain-C restatement (oracle/c/nsoracle.c), same interface
            from .native import CAssembler
            self.asm = CAssembler(self.sp, threads=threads)
        else:
            self.asm = Assembler(self.sp)
        # EquationCoefficientHandler (source/auxiliary_classes.py:251-306): conv=1, p=1, visc=1/Re
        co = dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re if Re else 1.0)
        if coefficients:
            co.update(coefficients)
        self.c_conv, self.c_p, self.c_visc = co["convective_term"], co["pressure_term"], co["viscous_term"]
        self.newton_iterations = []


class IPCSOracle(_Base):
    """State: velocities[0..2] (V), intermediate_velocity (V), pressure, old_pressure (Q)."""

    def __init__(self, mesh, Re, velocity_bcs, pressure_bcs=(), tol=1e-10, max_iter=50, time_fn=None,
                 assembler="numpy", threads=None):
        super().__init__(mesh, Re, assembler=assembler, threads=threads)
        s = self.sp
        self.velocities = [np.zeros(s.ndofs_V) for _ in range(3)]
        self.intermediate_velocity = np.zeros(s.ndofs_V)
        self.pressure = np.zeros(s.ndofs_Q)
        self.old_pressure = np.zeros(s.ndofs_Q)
        self.velocity_bcs = list(velocity_bcs)        # [(marker, comp, value(x,t) or const)]
        self.pressure_bcs = list(pressure_bcs)        # [(marker, value)]
        self.tol, self.max_iter = tol, max_iter
        self.alpha, self.k = (1.0, -1.0, 0.0), None
        self.linear_solver = "lu"          # "lu" (reference-faithful) | "krylov" / "krylov-c" (algorithm-matched)
        self.krylov_rtol = 1e-12
        self.krylov_iterations = []
        self.load = None                   # V-numbered external load  -c_b int f.w + int_Gamma t.w ds, or callable(t)

    def _bcs_at(self, t):
        def bind(v):
            return (lambda x, v=v: v(x, t)) if callable(v) else v
        vb = obc.dirichlet_values(self.sp, [("V", m, c, bind(v)) for m, c, v in self.velocity_bcs])
        pb = obc.dirichlet_values(self.sp, [("Q", m, "p", bind(v)) for m, v in self.pressure_bcs])
        return vb, pb

    def set_initial_conditions(self, velocity, pressure=None, degree=None):
        """``degree``: the degree of the dolfin Expression (App. A.7: interpolated into P_degree per cell, then
        L2-projected); None = nodal data of the space's own degree"""
        u0, p0 = _initial_fields(self.asm, velocity, pressure, degree)
        self.velocities[0][:] = u0
        self.velocities[1][:] = u0
        if p0 is not None:
            self.pressure[:] = p0
            self.old_pressure[:] = p0

    def solution_pressure(self, mean_value=None):
        """pressure part of ``solution`` incl. the mean-value shift of source/ns_solver_base.py:1190-1203"""
        from . import projection
        return self.pressure if mean_value is None else projection.mean_pressure_shift(self.asm, self.pressure, mean_value)

    def solve(self, alpha, k, t_next):
        """One ``_solve_time_step`` (source/ns_ipcs_solver.py:198-208)."""
        asm, a0 = self.asm, alpha[0]
        (vd, vv), (pd, pv) = self._bcs_at(t_next)
        un, unm1, pn = self.velocities[1], self.velocities[2], self.old_pressure

        # (1) tentative velocity: Newton with NonlinearVariationalSolver BC semantics
        load = self.load(t_next) if callable(self.load) else self.load

        def F(x):
            b = asm.ipcs_momentum_residual(x, un, unm1, pn, alpha, k, self.c_conv, self.c_p, self.c_visc)
            if load is not None:
                b = b + load
            return obc.apply_rhs_newton(b, x, vd, vv)

        def J(x):
            return obc.apply_ident(asm.ipcs_momentum_jacobian(x, a0, k, self.c_conv, self.c_visc), vd)

        krylov = self.linear_solver in ("krylov", "krylov-c")
        kit = dict(momentum=[], pressure=0, correction=0)
        if krylov:
            from . import krylov as _okr

            class okr:                      # "krylov": numpy statement; "krylov-c": the same in C/OpenMP
                gmres = staticmethod(_okr.gmres_c if self.linear_solver == "krylov-c" else _okr.gmres)
                pcg = staticmethod(_okr.pcg_c if self.linear_solver == "krylov-c" else _okr.pcg)

            def lin(A, b):
                x, n_it = okr.gmres(A, b, rtol=self.krylov_rtol)
                kit["momentum"].append(n_it)
                return x
        else:
            lin = lu_solve
        its, hist = newton(F, J, self.intermediate_velocity, atol=self.tol, rtol=10 * self.tol,
                           maxit=self.max_iter, solve=lin)
        self.newton_iterations.append(its)
        self.newton_history = hist
        # (2) pressure Poisson
        A = asm.poisson_matrix()
        b = asm.poisson_rhs(pn, self.intermediate_velocity, a0, k)
        if krylov:
            if len(pd):
                x0 = self.pressure.copy()
                x0[pd] = pv
                self.pressure[:], kit["pressure"] = okr.pcg(obc.apply_ident(A, pd), obc.apply_rhs_linear(b, pd, pv),
                                                            x0=x0, rtol=self.krylov_rtol)
            else:
                self.pressure[:], kit["pressure"] = okr.pcg(A, b, x0=self.pressure, rtol=self.krylov_rtol, nullspace=True)
        elif len(pd):
            self.pressure[:] = lu_solve(obc.apply_ident(A, pd), obc.apply_rhs_linear(b, pd, pv))
        else:
            self.pressure[:] = solve_singular_mean_zero(A, b - b.mean())
        # (3) velocity correction
        M = asm.velocity_mass_matrix()
        b = asm.correction_rhs(self.intermediate_velocity, self.pressure, self.old_pressure, a0, k)
        if krylov:
            x0 = self.intermediate_velocity.copy()
            x0[vd] = vv
            self.velocities[0][:], kit["correction"] = okr.pcg(obc.apply_ident(M, vd),
                                                               obc.apply_rhs_linear(b, vd, vv), x0=x0, rtol=self.krylov_rtol)
        else:
            self.velocities[0][:] = lu_solve(obc.apply_ident(M, vd), obc.apply_rhs_linear(b, vd, vv))
        self.krylov_iterations.append(kit)

    def advance_time(self):
        """``_advance_solution`` source/ns_ipcs_solver.py:35-43."""
        self.velocities[2][:] = self.velocities[1]
        self.velocities[1][:] = self.velocities[0]
        self.old_pressure[:] = self.pressure


class BDFOracle(_Base):
    """Monolithic BDF-2 Newton on W; solutions[0..2]."""

    def __init__(self, mesh, Re, velocity_bcs, pressure_bcs=(), tol=1e-10, max_iter=50, assembler="numpy",
                 threads=None):
        super().__init__(mesh, Re, assembler=assembler, threads=threads)
        self.solutions = [np.zeros(self.sp.ndofs_W) for _ in range(3)]
        self.velocity_bcs, self.pressure_bcs = list(velocity_bcs), list(pressure_bcs)
        self.tol, self.max_iter = tol, max_iter
        self.load = None                   # V-numbered external load, or callable(t)

    def _bcs_at(self, t):
        def bind(v):
            return (lambda x, v=v: v(x, t)) if callable(v) else v
        lst = [("W", m, c, bind(v)) for m, c, v in self.velocity_bcs]
        lst += [("W", m, "p", bind(v)) for m, v in self.pressure_bcs]
        return obc.dirichlet_values(self.sp, lst)

    def set_initial_conditions(self, velocity, pressure=None, degree=None):
        s = self.sp
        u0, p0 = _initial_fields(self.asm, velocity, pressure, degree)
        for w in self.solutions[:2]:
            w[s.V_to_W] = u0
            if p0 is not None:
                w[s.Q_to_W] = p0

    def solve(self, alpha, k, t_next, pin_pressure=False):
        asm = self.asm
        dofs, vals = self._bcs_at(t_next)
        w1, w2 = self.solutions[1], self.solutions[2]

        load = self.load(t_next) if callable(self.load) else self.load

        def F(x):
            b = asm.mixed_residual(x, self.c_conv, self.c_p, self.c_visc, bdf=(alpha, k, w1, w2))
            if load is not None:
                b[self.sp.V_to_W] += load
            return obc.apply_rhs_newton(b, x, dofs, vals)

        def J(x):
            A = asm.mixed_jacobian(x, self.c_conv, self.c_p, self.c_visc, mass_coef=alpha[0] / k)
            return obc.apply_ident(A, dofs)

        its, hist = newton(F, J, self.solutions[0], atol=self.tol, rtol=10 * self.tol, maxit=self.max_iter)
        self.newton_iterations.append(its)
        self.newton_history = hist

    def advance_time(self):
        """``_advance_solution`` source/ns_solver_base.py:1012-1016."""
        self.solutions[2][:] = self.solutions[1]
        self.solutions[1][:] = self.solutions[0]


class StationaryOracle(_Base):
    """Hybrid Picard -> Newton with SystemAssembler (symmetric local) BC elimination."""

    def __init__(self, mesh, Re, velocity_bcs, pressure_bcs=(), tol=1e-10, maxiter=50,
                 tol_picard=1e-2, maxiter_picard=10):
        super().__init__(mesh, Re)
        self.solution = np.zeros(self.sp.ndofs_W)
        lst = [("W", m, c, v) for m, c, v in velocity_bcs] + [("W", m, "p", v) for m, v in pressure_bcs]
        self.bc_dofs, self.bc_vals = obc.dirichlet_values(self.sp, lst)
        self.tol, self.maxiter = tol, maxiter
        self.tol_picard, self.maxiter_picard = tol_picard, maxiter_picard

    def system(self, x, newton_lin):
        """SystemAssembler(J, F, bcs): returns (A, b) with symmetric local elimination,
        BC values x - g (App. A.5)."""
        return symmetric_local_elimination(self.asm, x, self.bc_dofs, self.bc_vals,
                                           lambda: self._element_tensors(x, newton_lin))

    def _element_tensors(self, x, newton_lin):
        """Generator of (start, Ae, be) chunks for the SystemAssembler restatement."""
        asm, s = self.asm, self.sp
        d, nl = s.dim, s.nloc2
        nv = d * nl
        for st, en in asm._chunks():
            wdet, G2, G1 = asm._geo(st, en)
            U = asm._cell_vel_W(x, st, en)
            P = asm._cell_p_W(x, st, en)
            A = asm._momentum_jacobian(wdet, G2, U, self.c_conv, self.c_visc, 0.0, newton_lin)
            A = asm._to_V_local(A)
            A = np.swapaxes(A, 2, 3).reshape(A.shape[0], nv, nv)
            B = asm._to_V_local(asm._div_block(wdet, G2, self.c_p))
            Ae = np.zeros((A.shape[0], nv + d + 1, nv + d + 1))
            Ae[:, :nv, :nv] = A
            Ae[:, :nv, nv:] = B
            Ae[:, nv:, :nv] = np.swapaxes(B, 1, 2)
            F = asm._momentum_residual(wdet, G2, G1, U, None, P, self.c_conv, self.c_p, self.c_visc, 0.0)
            divu = np.einsum("cqIb,cIb->cq", G2, U)
            Fp = -self.c_p * np.einsum("cq,cq,qM->cM", wdet, divu, asm.phi1)
            yield st, Ae, np.concatenate([asm._to_V_local(F), Fp], axis=1)

    def solve(self):
        """``StationarySolverBase.solve`` source/ns_solver_base.py:951-988."""
        x = self.solution
        r = float(np.linalg.norm(self.system(x, False)[1]))
        tol_picard = self.tol_picard
        if r < tol_picard:
            order = math.floor(math.log10(r))
            tol_picard = (r / 10.0 ** order - 1.0) * 10.0 ** order
        it_p, _ = newton(lambda y: self.system(y, False)[1], lambda y: self.system(y, False)[0], x,
                         atol=tol_picard, rtol=1e-9, maxit=self.maxiter_picard,
                         error_on_nonconvergence=False)
        it_n, hist = newton(lambda y: self.system(y, True)[1], lambda y: self.system(y, True)[0], x,
                            atol=self.tol, rtol=1e-9, maxit=self.maxiter, error_on_nonconvergence=False)
        self.newton_iterations = [it_p, it_n]
        r = float(np.linalg.norm(self.system(x, True)[1]))
        assert r <= self.tol, "Newton iteration did not converge."
        return r


def symmetric_local_elimination(asm, x, bc_dofs, bc_g, element_tensors):
    """dolfin ``SystemAssembler`` restated (App. A.5): per cell, for each local BC column j
    with value v_j = x_j - g_j:  b_e -= A_e[:,j] v_j; zero row j and column j;
    A_e[j,j] = 1; b_e[j] = v_j.  Then add into the global tensors."""
    s = asm.sp
    n = s.ndofs_W
    is_bc = np.zeros(n, dtype=bool)
    is_bc[bc_dofs] = True
    v_glob = np.zeros(n)
    v_glob[bc_dofs] = x[bc_dofs] - bc_g
    rowptr, col, pos = asm.pattern("WW")
    vals = np.zeros(len(col))
    b = np.zeros(n)
    for st, Ae, be in element_tensors():
        dofs = s.cell_dofs_W[st:st + Ae.shape[0]]
        m = is_bc[dofs]                                    # [c, nloc]
        v = np.where(m, v_glob[dofs], 0.0)
        be = be - np.einsum("cij,cj->ci", Ae, v)
        Ae = Ae * (~m)[:, :, None] * (~m)[:, None, :]
        idx = np.arange(Ae.shape[1])
        Ae[:, idx, idx] = np.where(m, 1.0, Ae[:, idx, idx])
        be = np.where(m, v, be)
        vals += np.bincount(pos[st:st + Ae.shape[0]].ravel(), weights=Ae.ravel(), minlength=len(col))
        b += np.bincount(dofs.ravel(), weights=be.ravel(), minlength=n)
    return sp.csr_matrix((vals, col, rowptr), shape=(n, n)), b
