"""Binding recognised variational problems to the device code of the nsb200 mirror (``ipcs.IPCSSolver``).

The reference's ``IPCSSolver`` creates three dolfin solver objects and calls ``.solve()`` on each of them once per time
step (source/ns_ipcs_solver.py:198-208).  Here the three problems of one mesh share ONE engine -- an instance of the
mirror class working on the very tensors of the compat ``Function`` objects -- and each ``.solve()`` runs the matching
sub-step (``_solve_diffusion_step`` / ``_solve_projection_step`` / ``_solve_correction_step``).  The coefficients are
read from the recorded ``Constant`` objects at every solve (``Constant.assign`` is the reference's update mechanism,
source/ns_ipcs_solver.py:210-227), the Dirichlet data from the ``DirichletBC`` objects (their Expressions carry the
time the reference's ``_set_time`` gave them).
"""
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
    if u.function_space().kind != "V":
        raise R.Unrecognised("nonlinear problems on the mixed space (monolithic BDF / stationary solvers) run through the "
                             "nsb200 mirror classes ImplicitBDFSolver / StationarySolver; the compat layer covers IPCS")
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
