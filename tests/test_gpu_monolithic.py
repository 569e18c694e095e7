"""GPU parity of the monolithic path: mixed residual and
Jacobian entries (1e-12), SystemAssembler symmetric elimination, the direct solve, whole BDF-2 steps
and the stationary Picard -> Newton solve against the LU-based oracle (fields 1e-8), and the
drag / lift boundary functional."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from helpers import meshes, random_fields, rel_max_err

pytestmark = pytest.mark.gpu
TOL = 1e-12
COEF = dict(k=1e-2, alpha=(1.5, -2.0, 0.5), c_conv=1.0, c_p=1.0, c_visc=1.0 / 50.0)


@pytest.fixture(scope="module", params=[0, 1], ids=["2d", "3d"])
def case(request, built_library):
    import oracle
    from nsb200 import _lib as L
    from nsb200.device import Device
    name, m, om = meshes()[request.param]
    dev = Device(m, mixed=True)
    osp = oracle.dofmap.Spaces(om)
    asm = oracle.forms.Assembler(osp)
    dev.set_coefficients(COEF["k"], COEF["alpha"], COEF["c_conv"], COEF["c_p"], COEF["c_visc"])
    d, nn = m.dim, dev.spaces.n_nodes
    u0, u1, u2, p = random_fields(dev.spaces)

    def mixed(u, pp):
        w = np.zeros(dev.spaces.ndofs_W)
        w[osp.V_to_W] = u
        w[osp.Q_to_W] = pp
        return w
    ws = [mixed(u0, p), mixed(u1, 0.5 * p), mixed(u2, 0.25 * p)]
    return dict(dev=dev, asm=asm, osp=osp, ws=ws, d=d, L=L, om=om, mesh=m, fields=(u0, p))


def up(dev, a):
    return dev.upload(np.ascontiguousarray(a, dtype=np.float64))


def csr_of(dev, vals):
    pat = dev.patterns["WW"]
    return sp.csr_matrix((vals.cpu().numpy(), pat.col_host, pat.rowptr_host), shape=(pat.nrows, pat.nrows))


@pytest.mark.parametrize("bdf", [False, True], ids=["stationary", "bdf"])
def test_mixed_residual(case, bdf):
    dev, asm, L, ws = case["dev"], case["asm"], case["L"], case["ws"]
    b = dev.empty(dev.spaces.ndofs_W)
    dev.assemble_mixed(L.VEC | (L.BDF if bdf else 0), up(dev, ws[0]), up(dev, ws[1]), up(dev, ws[2]), b=b)
    ref = asm.mixed_residual(ws[0], COEF["c_conv"], COEF["c_p"], COEF["c_visc"],
                             bdf=(COEF["alpha"], COEF["k"], ws[1], ws[2]) if bdf else None)
    assert rel_max_err(b.cpu().numpy(), ref) < TOL


@pytest.mark.parametrize("newton", [True, False], ids=["newton", "picard"])
@pytest.mark.parametrize("bdf", [False, True], ids=["stationary", "bdf"])
def test_mixed_jacobian(case, bdf, newton):
    dev, asm, L, ws = case["dev"], case["asm"], case["L"], case["ws"]
    A = dev.empty(dev.patterns["WW"].nnz)
    flags = L.MAT | (L.BDF if bdf else 0) | (0 if newton else L.PICARD)
    dev.assemble_mixed(flags, up(dev, ws[0]), up(dev, ws[1]), up(dev, ws[2]), A=A)
    ref = asm.mixed_jacobian(ws[0], COEF["c_conv"], COEF["c_p"], COEF["c_visc"],
                             mass_coef=COEF["alpha"][0] / COEF["k"] if bdf else 0.0, newton=newton)
    ref = sp.csr_matrix(ref)
    diff = (csr_of(dev, A) - ref)
    assert abs(diff).max() < TOL * abs(ref).max()


def test_symmetric_elimination_matches_system_assembler(case):
    """flag NSB_SYMBC == dolfin SystemAssembler(J, F, bcs) (source/auxiliary_classes.py:141,160-164)"""
    import oracle
    dev, asm, L, ws, d = case["dev"], case["asm"], case["L"], case["ws"], case["d"]
    osp = case["osp"]
    lst = [("W", 1, None, lambda x: np.stack([1.0 + x[:, 1]] + [0.0 * x[:, 0]] * (d - 1), axis=1)),
           ("W", 3, None, (0.0,) * d), ("W", 2, "p", 0.5)]
    dofs, vals = oracle.bc.dirichlet_values(osp, lst)
    orc = oracle.solvers.StationaryOracle(case["om"], 50.0, [], [])
    orc.bc_dofs, orc.bc_vals = dofs, vals
    Aref, bref = orc.system(ws[0], True)
    mask = np.zeros(osp.ndofs_W, dtype=np.uint8)
    mask[dofs] = 1
    g = np.zeros(osp.ndofs_W)
    g[dofs] = vals
    dev.set_mixed_bc(dev.upload(mask), up(dev, g))
    A, b = dev.empty(dev.patterns["WW"].nnz), dev.empty(osp.ndofs_W)
    dev.assemble_mixed(L.VEC | L.MAT | L.SYMBC, up(dev, ws[0]), b=b, A=A)
    assert rel_max_err(b.cpu().numpy(), bref) < TOL
    assert abs(csr_of(dev, A) - Aref).max() < TOL * abs(Aref).max()


def test_direct_solve(case):
    """K16 library leg: x = A^-1 b on the mixed pattern against scipy's sparse LU"""
    dev, asm, L, ws = case["dev"], case["asm"], case["L"], case["ws"]
    pat = dev.patterns["WW"]
    rng = np.random.default_rng(7)
    A = sp.csr_matrix((rng.uniform(-1.0, 1.0, pat.nnz), pat.col_host, pat.rowptr_host), shape=(pat.nrows,) * 2)
    A = A + sp.diags(np.full(pat.nrows, 40.0))
    A.sort_indices()
    assert np.array_equal(A.indices, pat.col_host)
    b = rng.uniform(-1.0, 1.0, pat.nrows)
    x = dev.empty(pat.nrows)
    sec = dev.direct_solve("WW", up(dev, A.data), up(dev, b), x)
    ref = spla.spsolve(A.tocsc(), b)
    assert rel_max_err(x.cpu().numpy(), ref) < 1e-10
    total, n = dev.direct_seconds()
    assert sec > 0.0 and n >= 1 and total >= sec


def test_boundary_force_matches_oracle(case):
    import oracle
    from nsb200.monolithic import boundary_force
    dev, d = case["dev"], case["d"]
    u, p = case["fields"]
    for marker in (1, 4):
        F = boundary_force(dev, up(dev, u), up(dev, p), marker, 0.02)
        ref = oracle.functionals.boundary_force(case["osp"], u, p, marker, 0.02)
        assert rel_max_err(F, ref) < 1e-12


def _channel(d):
    import nsb200
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType

    def inflow(x, t):
        u = np.zeros((len(x), d))
        u[:, 0] = 6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + 0.5 * np.sin(np.pi * t))
        return u
    walls = [3, 4] + ([5, 6] if d == 3 else [])
    bcs = [(V.function, 1, nsb200.Expression(inflow))] + [(V.no_slip, w, None) for w in walls] + \
          [(P.constant, 2, 0.0)]
    o_vel = [(1, None, inflow)] + [(w, None, (0.0,) * d) for w in walls]
    return bcs, o_vel, [(2, 0.0)]


def test_bdf_steps_match_oracle(case):
    """3 monolithic BDF-2 steps (start-up BDF-1 included), source/ns_bdf_solver.py, against the LU oracle:
    mixed solution to 1e-8 relative, identical Newton iteration counts."""
    import nsb200
    import oracle
    d, m = case["d"], case["mesh"]
    # the unit-height channel of helpers.meshes() has y in [0, 1] (2D) / [0, 1] (3D): inflow profile fits
    Re, dt = 50.0, 0.01
    bcs, o_vel, o_prs = _channel(d)
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
    solver = nsb200.ImplicitBDFSolver(m, m.bfacet_marker, "standard", ts, tol=1e-10, max_iter=50)
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    solver.set_boundary_conditions(bcs)
    orc = oracle.solvers.BDFOracle(case["om"], Re, o_vel, o_prs, tol=1e-10, max_iter=50)
    ots = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
    for step in range(3):
        ts.update_coefficients()
        solver.solve()
        ots.update_coefficients()
        a = list(ots.coefficients(1)) + [0.0]
        orc.solve(tuple(a[:3]), ots.get_next_step_size(), ots.next_time)
        w = solver.solution.cpu().numpy()
        osp = orc.sp
        assert rel_max_err(w[osp.V_to_W], orc.solutions[0][osp.V_to_W]) < 1e-8, "velocity, step %d" % step
        assert rel_max_err(w[osp.Q_to_W], orc.solutions[0][osp.Q_to_W]) < 1e-8, "pressure, step %d" % step
        assert solver.newton_iterations[-1] == orc.newton_iterations[-1]
        ts.advance_time()
        solver.advance_time()
        ots.advance_time()
        orc.advance_time()
    sec, n = solver.direct_solve_seconds()
    assert n == sum(solver.newton_iterations) and sec > 0.0


def test_stationary_cavity_matches_oracle(built_library):
    """config C1 in small: stationary lid-driven cavity, hybrid Picard -> Newton with SystemAssembler
    elimination (source/ns_solver_base.py:894-988, demo/cavity_flow.py) against the oracle"""
    import nsb200
    import oracle
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    n, Re = 8, 10.0
    mesh, markers = nsb200.hyper_cube(2, n)
    om = oracle.mesh.hyper_rectangle((0.0, 0.0), (1.0, 1.0), (n, n))
    solver = nsb200.StationarySolver(mesh, markers, "standard")
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    bcs = [(V.no_slip, w, None) for w in (1, 2, 3)] + [(V.constant, 4, (1.0, 0.0))]
    solver.set_boundary_conditions(bcs)
    orc = oracle.solvers.StationaryOracle(om, Re, [(w, None, (0.0, 0.0)) for w in (1, 2, 3)] + [(4, None, (1.0, 0.0))])
    # singular in p (H4): both sides pin the first pressure dof
    orc.bc_dofs = np.append(orc.bc_dofs, orc.sp.Q_to_W[0])
    orc.bc_vals = np.append(orc.bc_vals, 0.0)
    order = np.argsort(orc.bc_dofs)
    orc.bc_dofs, orc.bc_vals = orc.bc_dofs[order], orc.bc_vals[order]
    orc.solve()
    solver.solve()
    w = solver.solution.cpu().numpy()
    osp = orc.sp
    assert rel_max_err(w[osp.V_to_W], orc.solution[osp.V_to_W]) < 1e-8
    pg, po = w[osp.Q_to_W], orc.solution[osp.Q_to_W]
    assert rel_max_err(pg - pg.mean(), po - po.mean()) < 1e-6
    assert solver.newton_iterations == orc.newton_iterations


def test_stationary_channel_matches_oracle(case):
    """stationary channel (tests/test_stationary_solvers.py:173-175 of the reference: inlet 6y(1-y)) with a
    pressure Dirichlet outlet: regular mixed system, Poiseuille steady state"""
    import nsb200
    import oracle
    d, m = case["d"], case["mesh"]
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    Re = 10.0

    def inflow(x):
        u = np.zeros((len(x), d))
        u[:, 0] = 6.0 * x[:, 1] * (1.0 - x[:, 1])
        return u
    walls = [3, 4]
    bcs = [(V.function, 1, inflow)] + [(V.no_slip, w, None) for w in walls] + [(P.constant, 2, 0.0)]
    o_vel = [(1, None, inflow)] + [(w, None, (0.0,) * d) for w in walls]
    if d == 3:      # free-slip side walls keep the flow planar
        bcs += [(V.constant_component, w, 2, 0.0) for w in (5, 6)]
        o_vel += [(w, 2, 0.0) for w in (5, 6)]
    solver = nsb200.StationarySolver(m, m.bfacet_marker, "standard")
    solver.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    solver.set_boundary_conditions(bcs)
    orc = oracle.solvers.StationaryOracle(case["om"], Re, o_vel, [(2, 0.0)])
    orc.solve()
    solver.solve()
    w = solver.solution.cpu().numpy()
    osp = orc.sp
    assert rel_max_err(w[osp.V_to_W], orc.solution[osp.V_to_W]) < 1e-8
    assert rel_max_err(w[osp.Q_to_W], orc.solution[osp.Q_to_W]) < 1e-8
    assert solver.newton_iterations == orc.newton_iterations
