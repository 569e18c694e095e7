"""The dolfin-named compat layer (navierstokes-with-fenics_b200/compat) on the real device: an IPCS time loop written
against the dolfin API -- function spaces, Functions, Constants, recorded forms, DirichletBC objects, the three
variational solvers, Function.assign history shifts -- gives the same fields as the nsb200 mirror class (same kernels:
bit-identical) and matches the LU oracle to 1e-8.  tests/test_compat_reference_classes.py runs the reference's own
IPCSSolver source on the same layer where /root/reference exists."""
import numpy as np
import pytest

from helpers import rel_max_err

pytestmark = pytest.mark.gpu


def test_ipcs_through_the_dolfin_api(built_library):
    import nsb200
    import oracle
    from nsb200 import compat
    D = compat.install()
    try:
        nx, ny, Re, dt, steps = 12, 4, 10.0, 0.01, 3
        mesh = D.RectangleMesh(D.Point(0.0, 0.0), D.Point(3.0, 1.0), nx, ny)
        markers = D.MeshFunction("size_t", mesh, 1)
        markers.set_all(0)
        for code, marker, val in (("near(x[0], val) && on_boundary", 1, 0.0), ("near(x[0], val) && on_boundary", 2, 3.0),
                                  ("near(x[1], val) && on_boundary", 3, 0.0), ("near(x[1], val) && on_boundary", 4, 1.0)):
            D.CompiledSubDomain(code, val=val).mark(markers, marker)
        cell = mesh.ufl_cell()
        W = D.FunctionSpace(mesh, D.MixedElement([D.VectorElement("CG", cell, 2), D.FiniteElement("CG", cell, 1)]))
        Vh, Qh = W.sub(0).collapse(), W.sub(1).collapse()
        u0, u1, u2, us = (D.Function(Vh) for _ in range(4))
        p, p_old = D.Function(Qh), D.Function(Qh)
        k, alpha = D.Constant(dt), [D.Constant(1.0), D.Constant(-1.0), D.Constant(0.0)]
        c_conv, c_p, c_visc = D.Constant(1.0), D.Constant(1.0), D.Constant(1.0 / Re)
        inlet = D.Expression(("6.0*x[1]*(1.0-x[1])*(1.0+0.5*sin(pi*time))", "0.0"), time=0.0, degree=2)
        zero = D.Constant((0.0, 0.0))
        vbcs = [D.DirichletBC(Vh, inlet, markers, 1), D.DirichletBC(Vh, zero, markers, 3), D.DirichletBC(Vh, zero, markers, 4)]
        pbcs = [D.DirichletBC(Qh, D.Constant(0.0), markers, 2)]
        dV = D.Measure("dx", domain=mesh)
        w, v = D.TestFunction(Vh), D.TrialFunction(Vh)
        q, r = D.TestFunction(Qh), D.TrialFunction(Qh)
        acc = (alpha[0] * D.dot(us, w) + alpha[1] * D.dot(u1, w) + alpha[2] * D.dot(u2, w)) / k
        F = (acc + c_conv * D.dot(D.dot(D.grad(us), us), w) - c_p * D.div(w) * p_old + c_visc * D.inner(D.grad(us), D.grad(w))) * dV
        diffusion = D.NonlinearVariationalSolver(D.NonlinearVariationalProblem(F, us, vbcs, D.derivative(F, us)))
        diffusion.parameters["newton_solver"]["relative_tolerance"] = 1e-9
        projection = D.LinearVariationalSolver(D.LinearVariationalProblem(
            D.dot(D.grad(r), D.grad(q)) * dV,
            (D.dot(D.grad(p_old), D.grad(q)) - alpha[0] / k * D.dot(D.div(us), q)) * dV, p, pbcs))
        correction = D.LinearVariationalSolver(D.LinearVariationalProblem(
            D.dot(v, w) * dV, (D.dot(us, w) - (k / alpha[0]) * D.dot(D.grad(p - p_old), w)) * dV, u0, vbcs))
        # the mirror and the oracle on the same scenario
        m2, mk2 = nsb200.hyper_rectangle((0.0, 0.0), (3.0, 1.0), (nx, ny))
        V, P = nsb200.VelocityBCType, nsb200.PressureBCType

        def inflow(x, t):
            return np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + 0.5 * np.sin(np.pi * t)), 0.0 * x[:, 0]], axis=1)
        ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
        mirror = nsb200.IPCSSolver(m2, mk2, "standard", ts)
        mirror.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
        mirror.set_boundary_conditions([(V.function, 1, nsb200.Expression(inflow)), (V.no_slip, 3, None), (V.no_slip, 4, None),
                                        (P.constant, 2, 0.0)])
        om = oracle.mesh.hyper_rectangle((0.0, 0.0), (3.0, 1.0), (nx, ny))
        orc = oracle.solvers.IPCSOracle(om, Re, [(1, None, inflow), (3, None, (0.0, 0.0)), (4, None, (0.0, 0.0))], [(2, 0.0)])
        for _ in range(steps):
            ts.update_coefficients()
            a = list(ts.coefficients(1)) + [0.0]
            for c, val in zip(alpha, a[:3]):
                c.assign(val)
            k.assign(ts.get_next_step_size())
            inlet.time = ts.next_time
            diffusion.solve()
            projection.solve()
            correction.solve()
            mirror.solve()
            orc.solve(tuple(a[:3]), ts.get_next_step_size(), ts.next_time)
            u2.assign(u1)
            u1.assign(u0)
            p_old.assign(p)
            ts.advance_time()
            mirror.advance_time()
            orc.advance_time()
        got_u, got_p = u1.vector().get_local(), p_old.vector().get_local()
        assert np.array_equal(got_u, mirror._velocities[1].cpu().numpy())          # same kernels, same order: same bits
        assert np.array_equal(got_p, mirror._old_pressure.cpu().numpy())
        assert rel_max_err(got_u, orc.velocities[1]) < 1e-8 and rel_max_err(got_p, orc.old_pressure) < 1e-8
        # FunctionAssigner + project + assemble (mean-pressure functionals of source/ns_solver_base.py:1190-1203)
        wfun = D.Function(W)
        D.FunctionAssigner(W, [Vh, Qh]).assign(wfun, [u1, p_old])
        _, pw = wfun.split()
        mean = D.assemble(pw * dV) / D.assemble(D.Constant(1.0) * dV)
        assert abs(mean - orc.asm.integral_Q(orc.old_pressure) / orc.asm.volume()) < 1e-8 * max(abs(mean), 1.0)
        proj = D.project(D.Expression(("x[0]*x[0]", "x[1]"), degree=2), Vh)
        xs = mesh.spaces().node_coordinates()
        assert rel_max_err(proj.vector().get_local(), np.stack([xs[:, 0] ** 2, xs[:, 1]], axis=1).reshape(-1)) < 1e-11
    finally:
        compat.uninstall()


def _channel(D, nx, ny, length=3.0):
    mesh = D.RectangleMesh(D.Point(0.0, 0.0), D.Point(length, 1.0), nx, ny)
    markers = D.MeshFunction("size_t", mesh, 1)
    markers.set_all(0)
    for code, marker, val in (("near(x[0], val) && on_boundary", 1, 0.0), ("near(x[0], val) && on_boundary", 2, length),
                              ("near(x[1], val) && on_boundary", 3, 0.0), ("near(x[1], val) && on_boundary", 4, 1.0)):
        D.CompiledSubDomain(code, val=val).mark(markers, marker)
    cell = mesh.ufl_cell()
    W = D.FunctionSpace(mesh, D.MixedElement([D.VectorElement("CG", cell, 2), D.FiniteElement("CG", cell, 1)]))
    return mesh, markers, W


def test_monolithic_bdf_through_the_dolfin_api(built_library):
    """the monolithic BDF-2 problem of source/ns_bdf_solver.py:36-106 written against the dolfin API (mixed space,
    split, TestFunctions, derivative, NonlinearVariationalSolver, Constant.assign, Function.assign): same bits as the
    mirror class, 1e-8 from the LU oracle, same Newton counts"""
    import nsb200
    import oracle
    from nsb200 import compat
    D = compat.install()
    try:
        nx, ny, Re, dt, steps = 12, 4, 10.0, 0.01, 3
        mesh, markers, W = _channel(D, nx, ny)
        sols = [D.Function(W) for _ in range(3)]
        k, alpha = D.Constant(dt), [D.Constant(1.0), D.Constant(-1.0), D.Constant(0.0)]
        c_conv, c_p, c_visc = D.Constant(1.0), D.Constant(1.0), D.Constant(1.0 / Re)
        inlet = D.Expression(("6.0*x[1]*(1.0-x[1])*(1.0+0.5*sin(pi*time))", "0.0"), time=0.0, degree=2)
        zero = D.Constant((0.0, 0.0))
        bcs = [D.DirichletBC(W.sub(0), inlet, markers, 1), D.DirichletBC(W.sub(0), zero, markers, 3),
               D.DirichletBC(W.sub(0), zero, markers, 4), D.DirichletBC(W.sub(1), D.Constant(0.0), markers, 2)]
        (w, q) = D.TestFunctions(W)
        u, p = D.split(sols[0])
        hist = [D.split(s)[0] for s in sols[1:]]
        dV = D.Measure("dx", domain=mesh)
        acc = D.dot(sum([alpha[0] * u, alpha[1] * hist[0], alpha[2] * hist[1]]), w) / k
        F = -(c_p * D.div(u) * q) * dV + (acc + c_conv * D.dot(D.dot(D.grad(u), u), w) - c_p * D.div(w) * p
                                           + c_visc * D.inner(D.grad(u), D.grad(w))) * dV
        solver = D.NonlinearVariationalSolver(D.NonlinearVariationalProblem(F, sols[0], bcs, D.derivative(F, sols[0])))
        solver.parameters["newton_solver"]["relative_tolerance"] = 1e-9

        m2, mk2 = nsb200.hyper_rectangle((0.0, 0.0), (3.0, 1.0), (nx, ny))
        V, P = nsb200.VelocityBCType, nsb200.PressureBCType

        def inflow(x, t):
            return np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + 0.5 * np.sin(np.pi * t)), 0.0 * x[:, 0]], axis=1)
        ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
        mirror = nsb200.ImplicitBDFSolver(m2, mk2, "standard", ts)
        mirror.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
        mirror.set_boundary_conditions([(V.function, 1, nsb200.Expression(inflow)), (V.no_slip, 3, None), (V.no_slip, 4, None),
                                        (P.constant, 2, 0.0)])
        om = oracle.mesh.hyper_rectangle((0.0, 0.0), (3.0, 1.0), (nx, ny))
        orc = oracle.solvers.BDFOracle(om, Re, [(1, None, inflow), (3, None, (0.0, 0.0)), (4, None, (0.0, 0.0))], [(2, 0.0)])
        counts = []
        for _ in range(steps):
            ts.update_coefficients()
            a = list(ts.coefficients(1)) + [0.0]
            for c, val in zip(alpha, a[:3]):
                c.assign(val)
            k.assign(ts.get_next_step_size())
            inlet.time = ts.next_time
            counts.append(solver.solve()[0])
            mirror.solve()
            orc.solve(tuple(a[:3]), ts.get_next_step_size(), ts.next_time)
            sols[2].assign(sols[1])
            sols[1].assign(sols[0])
            ts.advance_time()
            mirror.advance_time()
            orc.advance_time()
        got = sols[1].vector().get_local()
        assert np.array_equal(got, mirror._solutions[1].cpu().numpy())              # same kernels, same order: same bits
        osp = orc.sp
        assert rel_max_err(got[osp.V_to_W], orc.solutions[1][osp.V_to_W]) < 1e-8
        assert rel_max_err(got[osp.Q_to_W], orc.solutions[1][osp.Q_to_W]) < 1e-8
        assert counts == orc.newton_iterations == mirror.newton_iterations
    finally:
        compat.uninstall()


def test_stationary_newton_solver_through_the_dolfin_api(built_library):
    """the stationary problem of source/ns_solver_base.py:894-988 written against the dolfin API: NewtonSolver driving a
    NonlinearProblem subclass whose F / J call a SystemAssembler (Picard, then Newton linearisation)"""
    import nsb200
    import oracle
    from nsb200 import compat
    D = compat.install()
    try:
        n, Re = 8, 10.0
        mesh = D.RectangleMesh(D.Point(0.0, 0.0), D.Point(1.0, 1.0), n, n)
        markers = D.MeshFunction("size_t", mesh, 1)
        markers.set_all(0)
        for code, marker, val in (("near(x[0], val) && on_boundary", 1, 0.0), ("near(x[0], val) && on_boundary", 2, 1.0),
                                  ("near(x[1], val) && on_boundary", 3, 0.0), ("near(x[1], val) && on_boundary", 4, 1.0)):
            D.CompiledSubDomain(code, val=val).mark(markers, marker)
        cell = mesh.ufl_cell()
        W = D.FunctionSpace(mesh, D.MixedElement([D.VectorElement("CG", cell, 2), D.FiniteElement("CG", cell, 1)]))
        sol = D.Function(W)
        c_conv, c_p, c_visc = D.Constant(1.0), D.Constant(1.0), D.Constant(1.0 / Re)
        zero, lid = D.Constant((0.0, 0.0)), D.Constant((1.0, 0.0))
        bcs = [D.DirichletBC(W.sub(0), zero, markers, m) for m in (1, 2, 3)] + [D.DirichletBC(W.sub(0), lid, markers, 4)]
        (v, r), (w, q) = D.TrialFunctions(W), D.TestFunctions(W)
        u, p = D.split(sol)
        dV = D.Measure("dx", domain=mesh)
        F = -(c_p * D.div(u) * q) * dV + (c_conv * D.dot(D.dot(D.grad(u), u), w) - c_p * D.div(w) * p
                                           + c_visc * D.inner(D.grad(u), D.grad(w))) * dV
        J_picard = -(c_p * D.div(v) * q) * dV + (c_conv * D.dot(D.dot(D.grad(v), u), w) - c_p * D.div(w) * r
                                                  + c_visc * D.inner(D.grad(v), D.grad(w))) * dV
        J_newton = D.derivative(F, sol)

        class Problem(D.NonlinearProblem):
            calls = 0

            def __init__(self, J):
                super().__init__()
                self.assembler = D.SystemAssembler(J, F, bcs)

            def F(self, b, x):
                Problem.calls += 1
                self.assembler.assemble(b, x)

            def J(self, A, x):
                self.assembler.assemble(A)

        newton = D.NewtonSolver(D.MPI.comm_world, D.PETScLUSolver(), D.PETScFactory.instance())
        picard_problem, newton_problem = Problem(J_picard), Problem(J_newton)
        newton.parameters["maximum_iterations"] = 10
        newton.parameters["absolute_tolerance"] = 1e-2
        it_p, _ = newton.solve(picard_problem, sol.vector())
        newton.parameters["absolute_tolerance"] = 1e-10
        newton.parameters["maximum_iterations"] = 50
        newton.parameters["error_on_nonconvergence"] = False
        it_n, ok = newton.solve(newton_problem, sol.vector())
        res = D.Vector(sol.vector())
        newton_problem.F(res, sol.vector())
        assert ok and res.norm("l2") <= 1e-10 and Problem.calls >= it_p + it_n + 3     # the callbacks were honoured

        m2, mk2 = nsb200.hyper_cube(2, n)
        mirror = nsb200.StationarySolver(m2, mk2, "standard")
        mirror.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
        V = nsb200.VelocityBCType
        mirror.set_boundary_conditions([(V.no_slip, m, None) for m in (1, 2, 3)] + [(V.constant, 4, (1.0, 0.0))])
        mirror.solve()
        got = sol.vector().get_local()
        assert [it_p, it_n] == mirror.newton_iterations
        assert np.array_equal(got, mirror.solution.cpu().numpy())
        orc = oracle.solvers.StationaryOracle(oracle.mesh.hyper_cube(2, n), Re,
                                              [(m, None, (0.0, 0.0)) for m in (1, 2, 3)] + [(4, None, (1.0, 0.0))])
        orc.bc_dofs = np.append(orc.bc_dofs, orc.sp.Q_to_W[0])          # the pressure pin of the enclosed flow
        orc.bc_vals = np.append(orc.bc_vals, 0.0)
        order = np.argsort(orc.bc_dofs)
        orc.bc_dofs, orc.bc_vals = orc.bc_dofs[order], orc.bc_vals[order]
        orc.solve()
        osp = orc.sp
        assert rel_max_err(got[osp.V_to_W], orc.solution[osp.V_to_W]) < 1e-8
        assert rel_max_err(got[osp.Q_to_W], orc.solution[osp.Q_to_W]) < 1e-8
    finally:
        compat.uninstall()
