"""SURVEY.md 8f rows 1-3 on the device against their oracle restatements (oracle/projection.py): dolfin.project
of degree-k Expressions (initial conditions, source/ns_solver_base.py:1123-1171), the mean-pressure shift
(:1190-1203), dolfin.errornorm (convergence_test/taylor_green_vortex.py:101-119), the CFL diagnostic
(source/ns_problem.py:554-587) and the DG post-processing fields (vorticity, pressure gradient, :55-101)."""
import numpy as np
import pytest

from helpers import meshes, random_fields, rel_max_err

pytestmark = pytest.mark.gpu


def _fields(d):
    g = 2.0 * np.pi

    def vel(x):
        cols = [np.cos(g * x[:, c]) * np.sin(g * x[:, (c + 1) % d]) + 0.2 * x[:, c] ** 3 for c in range(d)]
        return np.stack(cols, axis=1)

    def prs(x):
        return -0.25 * (np.cos(2.0 * g * x[:, 0]) + np.cos(2.0 * g * x[:, 1])) + x[:, 0] ** 3
    return vel, prs


@pytest.fixture(scope="module", params=[0, 1], ids=["2d", "3d"])
def case(request, built_library):
    import oracle
    from nsb200.device import Device
    name, m, om = meshes()[request.param]
    dev = Device(m)
    asm = oracle.forms.Assembler(oracle.dofmap.Spaces(om))
    return dict(dev=dev, asm=asm, d=m.dim)


@pytest.mark.parametrize("degree", [1, 2, 3])
def test_project_matches_oracle(case, degree):
    from oracle import projection as opj
    dev, asm, d = case["dev"], case["asm"], case["d"]
    vel, prs = _fields(d)
    u = dev.project(vel, "V", degree).cpu().numpy()
    p = dev.project(prs, "Q", degree).cpu().numpy()
    assert rel_max_err(u, opj.project(asm, vel, "V", degree)) < 1e-12
    assert rel_max_err(p, opj.project(asm, prs, "Q", degree)) < 1e-12
    if degree == 2:      # data inside the space: the projection is the nodal interpolant (what the old code assumed)
        assert rel_max_err(u, vel(asm.sp.node_coords()).reshape(-1)) < 1e-12


def test_errornorm_and_cfl_match_oracle(case):
    from oracle import projection as opj
    dev, asm, d = case["dev"], case["asm"], case["d"]
    vel, prs = _fields(d)
    us, _, _, p = random_fields(dev.spaces)
    for space, exact, uh in (("V", vel, us), ("Q", prs, p)):
        got = dev.errornorm(exact, dev.upload(uh), space)
        ref = opj.errornorm(asm, exact, uh, space)
        assert abs(got - ref) < 1e-11 * ref
    for dt in (1e-2, 0.3):
        got = dev.cfl(dev.upload(us), dt)
        ref = opj.cfl_number(asm, us, dt)
        assert abs(got - ref) < 1e-12 * ref


def test_dg_postprocessing_fields_match_oracle(case):
    """vorticity and pressure gradient of source/ns_problem.py:55-101 (dolfin.project onto DG_1 / [DG_0]^d): the device
    evaluates them in closed form per cell, the oracle by a genuine cell-wise L2 projection with quadrature"""
    from oracle import projection as opj
    dev, asm, d = case["dev"], case["asm"], case["d"]
    us, _, _, p = random_fields(dev.spaces)
    w = dev.vorticity(dev.upload(us)).cpu().numpy()
    assert w.shape == ((asm.sp.mesh.num_cells, d + 1) if d == 2 else (asm.sp.mesh.num_cells, d + 1, 3))
    assert rel_max_err(w, opj.vorticity_dg(asm, us)) < 1e-12
    g = dev.pressure_gradient(dev.upload(p)).cpu().numpy()
    assert rel_max_err(g, opj.pressure_gradient_dg(asm, p)) < 1e-12


def test_initial_conditions_are_projected(built_library):
    """Taylor-Green initial data as the demo declares it (Expression, degree 3; demo/taylor_green_vortex.py:61-69):
    IPCS and BDF mirrors hold the L2 projection of the P3 interpolant, not the nodal values"""
    import nsb200
    import oracle
    from oracle import projection as opj
    g = 2.0 * np.pi
    vel = lambda x, t=0.0: np.stack([np.cos(g * x[:, 0]) * np.sin(g * x[:, 1]),                 # noqa: E731
                                     -np.sin(g * x[:, 0]) * np.cos(g * x[:, 1])], axis=1)
    prs = lambda x, t=0.0: -0.25 * (np.cos(2.0 * g * x[:, 0]) + np.cos(2.0 * g * x[:, 1]))      # noqa: E731
    n = 8
    mesh, markers = nsb200.hyper_cube(2, n)
    asm = oracle.forms.Assembler(oracle.dofmap.Spaces(oracle.mesh.hyper_cube(2, n)))
    ou, op = opj.project(asm, vel, "V", 3), opj.project(asm, prs, "Q", 3)
    assert rel_max_err(ou, vel(asm.sp.node_coords()).reshape(-1)) > 1e-4      # the projection is NOT the interpolant
    ic = dict(velocity=nsb200.Expression(vel, degree=3), pressure=nsb200.Expression(prs, degree=3))
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=1e-3)
    ipcs = nsb200.IPCSSolver(mesh, markers, "standard", ts)
    ipcs.set_initial_conditions(ic)
    for f in (ipcs._velocities[0], ipcs._velocities[1]):
        assert rel_max_err(f.cpu().numpy(), ou) < 1e-12
    for f in (ipcs._pressure, ipcs._old_pressure):
        assert rel_max_err(f.cpu().numpy(), op) < 1e-12
    bdf = nsb200.ImplicitBDFSolver(mesh, markers, "standard", ts)
    bdf.set_initial_conditions(ic)
    osp = asm.sp
    for w in bdf._solutions[:2]:
        w = w.cpu().numpy()
        assert rel_max_err(w[osp.V_to_W], ou) < 1e-12 and rel_max_err(w[osp.Q_to_W], op) < 1e-12
    # constants are in the space
    ipcs.set_initial_conditions(dict(velocity=(0.5, -2.0), pressure=3.0))
    assert float((ipcs._velocities[0].view(-1, 2)[:, 1] + 2.0).abs().max()) == 0.0


def test_mean_pressure_shift_bdf_and_ipcs(built_library):
    """PressureBCType.mean_value (source/ns_solver_base.py:1190-1203): after solve() the mean of the returned pressure,
    measured with the ORACLE's mass matrix, equals the requested value -- for the monolithic BDF solver (ADVICE r1:
    it used to ignore the request) and for the IPCS solution copy"""
    import nsb200
    import oracle
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    n, target = 6, 0.75
    mesh, markers = nsb200.hyper_cube(2, n)
    asm = oracle.forms.Assembler(oracle.dofmap.Spaces(oracle.mesh.hyper_cube(2, n)))
    osp = asm.sp
    bcs = [(V.no_slip, w, None) for w in (1, 2, 3)] + [(V.constant, 4, (1.0, 0.0)), (P.mean_value, None, target)]
    for cls in (nsb200.ImplicitBDFSolver, nsb200.IPCSSolver):
        ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=1e-2)
        solver = cls(mesh, markers, "standard", ts)
        solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.1))
        solver.set_boundary_conditions(bcs)
        for _ in range(2):
            ts.update_coefficients()
            solver.solve()
            w = solver.solution.cpu().numpy()
            mean = asm.integral_Q(w[osp.Q_to_W]) / asm.volume()
            assert abs(mean - target) < 1e-12, (cls.__name__, mean)
            ts.advance_time()
            solver.advance_time()
