"""Binding recognised variational problems to the device code of the nsb200 mirrors (``ipcs.IPCSSolver``,
``bdf.ImplicitBDFSolver``, ``stationary.StationarySolver``).

The reference's ``IPCSSolver`` creates three dolfin solver objects and calls ``.solve()`` on each of them once per time
step (source/ns_ipcs_solver.py:198-208).  Here the three problems of one mesh share ONE engine -- an instance of the
mirror class working on the very tensors of the compat ``Function`` objects -- and each ``.solve()`` runs the matching
sub-step (``_solve_diffusion_step`` / ``_solve_projection_step`` / ``_solve_correction_step``).  The coefficients are
read from the recorded ``Constant`` objects at every solve (``Constant.assign`` is the reference's update mechanism,
source/ns_ipcs_solver.py:210-227), the Dirichlet data from the ``DirichletBC`` objects (their Expressions carry the
time the reference's ``_set_time`` gave them).
"""
import numpy as np

from . import recognise as R


class _Levels:
    """what the mirror's constructor asks of a time-stepping object; the compat layer feeds the coefficients itself"""

    def __init__(self, n):
        self._n = n

    def n_levels(self):
        return self._n

    def coefficients(self, derivative=1):
        return (1.0, -1.0, 0.0)

    def get_next_step_size(self):
        return 1.0

    next_time = 0.0


def _engine(mesh, n_levels=2):
    """the mirror IPCSSolver of this mesh on the mesh's Device (created once)"""
    if mesh._engine is None:
        from ..ipcs import IPCSSolver
        eng = IPCSSolver(mesh._m, mesh._m.bfacet_marker, "standard", _Levels(n_levels), device=mesh._dev())
        eng._compat = dict(diffusion=None, projection=None, correction=None, bcs_version=None)
        mesh._engine = eng
    return mesh._engine


def _bc_tuples(bcs):
    return [bc.as_tuple() for bc in bcs]


def _sync_bcs(eng, kind, bcs):
    """hand the DirichletBC lists to the engine (velocity list from the diffusion / correction problems, pressure list
    from the projection problem); re-run the engine's table setup when the lists change"""
    tuples = _bc_tuples(bcs)
    attr = "_velocity_bcs" if kind == "velocity" else "_pressure_bcs"
    if getattr(eng, "_compat_" + attr, None) is not bcs:
        setattr(eng, attr, tuples)
        setattr(eng, "_compat_" + attr, bcs)
        eng._bc_ready = False


def _apply_coefficients(eng, force=False):
    """k, alpha_i, c_conv, c_p, c_visc from the Constants recorded in the tentative-velocity residual"""
    d = eng._compat["diffusion"]
    if d is None:
        raise RuntimeError("the IPCS sub-problems must be set up before they are solved")
    k_alpha = {}
    for key, monos in d["acc"].items():
        k_alpha[key] = sum(m.value() for m in monos)          # alpha_i / k
    # the step size is the Constant that divides the acceleration term (source/ns_ipcs_solver.py:33)
    inv = [c for c, pw in d["acc"]["u"][0].consts if pw == -1]
    if len(inv) != 1:
        raise R.Unrecognised("the acceleration term must be divided by exactly one Constant (the step size)")
    k = float(inv[0])
    a0_over_k = k_alpha["u"]
    c = eng._compat["correction"]
    if c is not None:
        k_over_a0 = max(abs(m.value()) for _, m in c["grads"])
        if abs(k_over_a0 * a0_over_k - 1.0) > 1e-12:
            raise R.Unrecognised("the correction step's k / alpha_0 does not match the momentum equation's alpha_0 / k")
    p = eng._compat["projection"]
    if p is not None and abs(abs(p["c_div"].value()) - a0_over_k) > 1e-12 * a0_over_k:
        raise R.Unrecognised("the projection step's alpha_0 / k does not match the momentum equation's")
    hist = sorted(k for k in k_alpha if k != "u")
    alpha = [a0_over_k * k] + [k_alpha[h] * k for h in hist]
    while len(alpha) < 3:
        alpha.append(0.0)
    eng._alpha, eng._next_step_size = tuple(alpha), k
    new = (k,) + eng._alpha + (d["c_conv"].value(), -d["c_p"].value(), d["c_visc"].value())
    if force or eng.dev.coef is None or tuple(eng.dev.coef[:7]) != new:
        eng.dev.set_coefficients(k, eng._alpha, new[4], new[5], new[6])


def _adopt(eng):
    """make the engine work on the tensors of the compat Functions (no copies, no second set of fields)"""
    d, p, c = eng._compat["diffusion"], eng._compat["projection"], eng._compat["correction"]
    if d is not None:
        eng._intermediate_velocity = d["ustar"]._tensor
        eng._old_pressure = d["p_old"]._tensor
        hist = [d["history"][k] for k in sorted(d["history"])]
        for i, f in enumerate(hist):
            eng._velocities[i + 1] = f._tensor
    if p is not None:
        eng._pressure = p["p_new"]._tensor
    if c is not None:
        eng._velocities[0] = c["u_new"]._tensor


def _prepare(eng, first_of_step=False):
    """what ``InstationarySolverBase.solve`` does before ``_solve_time_step`` in the mirror; the tentative-velocity
    solve opens a time step (source/ns_ipcs_solver.py:202), so the coefficients are handed over there unconditionally
    (the reference re-assigns its Constants every step, quirk Q2) and only on change by the two linear solves"""
    if not eng._bc_ready:
        eng._setup_boundary_conditions()
    eng._update_bc_values()
    eng._apply_forms(eng.dev)
    _apply_coefficients(eng, force=first_of_step)
    if not (eng._have_constant and eng._cache):
        eng._assemble_constant()
    eng._load = None


def bind_nonlinear(problem):
    u = problem.u
    if u.function_space().kind == "W":
        return bind_mixed_bdf(problem)
    if u.function_space().kind != "V":
        raise R.Unrecognised("nonlinear problems live on the velocity space (IPCS) or on the mixed space (BDF)")
    rec = R.recognise_ipcs_diffusion(problem.F, u, problem.J)
    eng = _engine(u.function_space().mesh(), max(len(rec["history"]), 1))
    eng._compat["diffusion"] = rec
    _sync_bcs(eng, "velocity", problem.bcs)
    _adopt(eng)

    def run(newton_parameters):
        eng._tol = float(newton_parameters["absolute_tolerance"])
        eng._maxiter = int(newton_parameters["maximum_iterations"])
        if abs(float(newton_parameters["relative_tolerance"]) - 10.0 * eng._tol) > 1e-15 * eng._tol:
            raise R.Unrecognised("the Newton relative tolerance must be 10 x the absolute one (source/ns_ipcs_solver.py:146)")
        _adopt(eng)
        _prepare(eng, first_of_step=True)
        eng._solve_diffusion_step()         # raises RuntimeError on non-convergence, like dolfin
        return eng.newton_iterations[-1], True
    return run


def bind_linear(problem):
    u = problem.u
    mesh = u.function_space().mesh()
    kind = u.function_space().kind
    if kind == "Q":
        rec = R.recognise_ipcs_projection(problem.a, problem.L, u)
        eng = _engine(mesh)
        eng._compat["projection"] = rec
        _sync_bcs(eng, "pressure", problem.bcs)
        step = "_solve_projection_step"
    elif kind == "V":
        rec = R.recognise_ipcs_correction(problem.a, problem.L, u)
        eng = _engine(mesh)
        eng._compat["correction"] = rec
        _sync_bcs(eng, "velocity", problem.bcs)
        step = "_solve_correction_step"
    else:
        raise R.Unrecognised("linear problems on the mixed space are outside the compat vocabulary")
    _adopt(eng)

    def run():
        _adopt(eng)
        _prepare(eng)
        getattr(eng, step)()
    return run


# =====================================================================================================
# monolithic problems on the mixed space (source/ns_bdf_solver.py:88-106, source/ns_solver_base.py:938-983)
# =====================================================================================================
def _split_bcs(bcs):
    tuples = _bc_tuples(bcs)
    return [t for t in tuples if t[1] != "p"], [t for t in tuples if t[1] == "p"]


def _mixed_engine(mesh, kind, n_levels=2):
    """the mirror ImplicitBDFSolver / StationarySolver of this mesh on the mesh's Device (created once)"""
    if mesh._engine is None:
        if kind == "bdf":
            from ..bdf import ImplicitBDFSolver
            eng = ImplicitBDFSolver(mesh._m, mesh._m.bfacet_marker, "standard", _Levels(n_levels), device=mesh._dev())
        else:
            from ..stationary import StationarySolver
            eng = StationarySolver(mesh._m, mesh._m.bfacet_marker, "standard", device=mesh._dev())
        eng._compat = dict(kind=kind, residual=None, coefficients=None)
        mesh._engine = eng
    if mesh._engine._compat.get("kind") != kind:
        raise R.Unrecognised("one mesh carries one solver family on the compat layer (found %s, asked for %s)"
                             % (mesh._engine._compat.get("kind"), kind))
    return mesh._engine


def _bcs_signature(bcs):
    """what a DirichletBC list says, by content: the reference rebuilds its problem (and the list) on every step
    (``_required_objects = ("_solver")`` is a string) -- an equal list must not cost a new table"""
    from .dolfin import Constant
    sig = []
    for bc in bcs:
        v = bc.value
        sig.append((bc.space.key(), bc.marker_id,
                    ("const", tuple(np.atleast_1d(np.asarray(v._v, dtype=np.float64)))) if isinstance(v, Constant) else ("obj", id(v))))
    return tuple(sig)


def _sync_mixed_bcs(eng, bcs):
    sig = _bcs_signature(bcs)
    if getattr(eng, "_compat_bcs_sig", None) != sig:
        eng._velocity_bcs, eng._pressure_bcs = _split_bcs(bcs)
        eng._compat_bcs_sig = sig
        eng._compat_bcs = bcs             # keeps the value objects alive: their ids are part of the signature
        eng._bc_ready = False


def _apply_mixed_coefficients(eng):
    """k, alpha_i (BDF) and c_conv, c_p, c_visc from the Constants recorded in the mixed residual, at every solve"""
    rec = eng._compat["residual"]
    c_conv, c_p, c_visc = rec["c_conv"].value(), -rec["c_p"].value(), rec["c_visc"].value()
    if abs(rec["c_div"].value() + c_p) > 1e-14 * abs(c_p):
        raise R.Unrecognised("the mass balance and the pressure term must carry the same coefficient c_p")
    if rec["acc"]:
        inv = [c for c, pw in rec["acc"]["s"][0].consts if pw == -1]
        if len(inv) != 1:
            raise R.Unrecognised("the acceleration term must be divided by exactly one Constant (the step size)")
        k = float(inv[0])
        alpha = [sum(m.value() for m in rec["acc"]["s"]) * k]
        alpha += [sum(m.value() for m in rec["acc"][h]) * k for h in sorted(rec["history"])]
        while len(alpha) < 3:
            alpha.append(0.0)
    else:
        k, alpha = 1.0, [1.0, -1.0, 0.0]
    new = (k,) + tuple(alpha) + (c_conv, c_p, c_visc)
    eng._equation_coefficients = dict(convective_term=c_conv, pressure_term=c_p, viscous_term=c_visc)
    if eng._compat["coefficients"] != new:
        eng._compat["coefficients"] = new
        eng._alpha, eng._next_step_size = tuple(alpha), k
        if rec["acc"]:
            eng._mass_coefficient = alpha[0] / k
            eng._vel_lmax = None
        eng.dev.set_coefficients(k, tuple(alpha), c_conv, c_p, c_visc)


def bind_mixed_bdf(problem):
    """NonlinearVariationalProblem(F, w, bcs, derivative(F, w)) of source/ns_bdf_solver.py:88-94"""
    w = problem.u
    rec = R.recognise_mixed_residual(problem.F, w)
    if R.recognise_mixed_jacobian(problem.J, problem.F, w) != "newton" or not rec["acc"]:
        raise R.Unrecognised("instationary mixed problem: expected the BDF acceleration term and J = derivative(F, w)")
    mesh = w.function_space().mesh()
    eng = _mixed_engine(mesh, "bdf", max(len(rec["history"]), 1))
    eng._compat["residual"] = rec
    _sync_mixed_bcs(eng, problem.bcs)

    def adopt():
        eng._solutions[0] = w._tensor
        for i, h in enumerate(sorted(rec["history"])):
            eng._solutions[i + 1] = rec["history"][h]._tensor
    adopt()

    def run(newton_parameters):
        eng._tol = float(newton_parameters["absolute_tolerance"])
        eng._maxiter = int(newton_parameters["maximum_iterations"])
        if abs(float(newton_parameters["relative_tolerance"]) - 10.0 * eng._tol) > 1e-15 * eng._tol:
            raise R.Unrecognised("the Newton relative tolerance must be 10 x the absolute one (source/ns_bdf_solver.py:99)")
        adopt()
        if not eng._bc_ready:
            eng._setup_boundary_conditions()
        eng._prepare_forms_and_load()
        eng._update_bc_values()
        _apply_mixed_coefficients(eng)
        eng._solve_time_step()                # raises RuntimeError on non-convergence, like dolfin
        return eng.newton_iterations[-1], True
    return run


class StationaryBinding:
    """what ``SystemAssembler(J, F, bcs)`` stands for on the device: the assembly calls of one linearisation of the
    stationary problem (source/auxiliary_classes.py:141-164), on the mirror StationarySolver of the mesh"""

    def __init__(self, J, F, bcs):
        from .expr import Derivative
        w = J.u if isinstance(J, Derivative) else None
        parents = R._split_parents(R._monomials(F))
        if w is None:
            if len(parents) != 1:
                raise R.Unrecognised("stationary residual: expected exactly one mixed function")
            w = parents[0]
        self.w, self.F_form = w, F
        self.rec = R.recognise_mixed_residual(F, w)
        if self.rec["acc"]:
            raise R.Unrecognised("SystemAssembler is bound to the stationary problem (no acceleration term)")
        self.picard = R.recognise_mixed_jacobian(J, F, w) == "picard"
        self.eng = _mixed_engine(w.function_space().mesh(), "stationary")
        self.eng._compat["residual"] = self.rec
        _sync_mixed_bcs(self.eng, bcs)

    def _prepare(self):
        eng = self.eng
        eng._solution = self.w._tensor
        _apply_mixed_coefficients(eng)
        if not eng._bc_ready:
            eng._setup_boundary_conditions()
            eng._prepare_forms_and_load()

    def assemble_vector(self, b_tensor, x_tensor):
        """b = F(x) with the symmetric local elimination (values x - g on the Dirichlet rows)"""
        eng = self.eng
        if x_tensor.data_ptr() != self.w._tensor.data_ptr():
            raise R.Unrecognised("SystemAssembler.assemble(b, x): x must be the vector of the Function the forms hold")
        self._prepare()
        keep, eng._b = eng._b, b_tensor
        try:
            eng._assemble_system(self.picard, False)
        finally:
            eng._b = keep

    def assemble_matrix(self):
        """A = J(x) with the symmetric elimination, into the engine's matrix storage (the solver owns it); the fused
        kernel fills the engine's b with F(x) again, the same values assemble_vector gave for this x"""
        self._prepare()
        self.eng._assemble_system(self.picard, True)

    def newton(self, problem, x_vec, parameters):
        """dolfin's NewtonSolver loop, driven through the problem's F / J callbacks (source/ns_solver_base.py:974-983)"""
        from .dolfin import _Vector, _RawVectorOwner
        eng = self.eng
        if x_vec._f._tensor.data_ptr() != self.w._tensor.data_ptr():
            raise R.Unrecognised("NewtonSolver.solve(problem, x): x must be the vector of the Function the forms hold")
        self._prepare()
        b_vec = _Vector(_RawVectorOwner(eng._b, x_vec._f))
        handle = _MatrixHandle(self)

        def residual():
            problem.F(b_vec, x_vec)
            return eng.dev.norm(eng._b)

        def jacobian():
            problem.J(handle, x_vec)

        atol = float(parameters.get("absolute_tolerance", 1e-10))
        rtol = float(parameters.get("relative_tolerance", 1e-9))
        maxit = int(parameters.get("maximum_iterations", 50))
        err = bool(parameters.get("error_on_nonconvergence", True))
        it = eng._newton(residual, jacobian, self.w._tensor, atol=atol, rtol=rtol, maxit=maxit,
                         error_on_nonconvergence=err)
        eng.newton_iterations.append(it)
        r, r0 = eng.newton_history[-1], eng.newton_history[0]
        return it, bool(r < atol or (r0 > 0 and r / r0 < rtol))


class _MatrixHandle:
    """the ``A`` a NewtonSolver hands to ``NonlinearProblem.J(A, x)``: names the engine's matrix storage"""

    def __init__(self, binding):
        self.binding = binding
