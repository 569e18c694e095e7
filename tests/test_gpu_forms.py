"""every weak-form variant of the reference (convective forms source/ns_solver_base.py:370-390 /
:478-499, viscous forms :662-673, Coriolis term :173-191, body force :158-171, boundary tractions :121-156) through
the generic row kernels (csrc/assemble_forms.cuh) against the quadrature oracle, entry by entry to 1e-12; the generic
kernels against the tuned ones on the default form; one solver-level run with everything switched on."""
import numpy as np
import pytest
import scipy.sparse as sp

from helpers import meshes, random_fields, rel_max_err, bsr_to_csr

pytestmark = pytest.mark.gpu
TOL = 1e-12
COEF = dict(k=1e-2, alpha=(1.5, -2.0, 0.5), c_conv=0.8, c_p=1.1, c_visc=1.0 / 50.0)
CONV = ["standard", "rotational", "divergence", "skew_symmetric"]
VISC = ["reduced", "traction"]


def up(dev, a):
    return dev.upload(np.ascontiguousarray(a, dtype=np.float64))


@pytest.fixture(scope="module", params=[0, 1], ids=["2d", "3d"])
def case(request, built_library):
    import oracle
    from nsb200.device import Device
    name, m, om = meshes()[request.param]
    dev = Device(m)
    devW = Device(m, mixed=True)
    asm = oracle.forms.Assembler(oracle.dofmap.Spaces(om))
    for d_ in (dev, devW):
        d_.set_coefficients(COEF["k"], COEF["alpha"], COEF["c_conv"], COEF["c_p"], COEF["c_visc"])
    return dict(dev=dev, devW=devW, asm=asm, fields=random_fields(dev.spaces), d=m.dim)


def _set(case, conv, visc, coriolis):
    asm, d = case["asm"], case["d"]
    cor = None
    if coriolis:
        cor = (0.0, 0.0, 0.7) if d == 2 else (0.3, -0.2, 0.7)
    asm.conv_form, asm.visc_form, asm.coriolis = conv, visc, cor
    for dev in (case["dev"], case["devW"]):
        dev.set_forms(CONV.index(conv), VISC.index(visc), cor)


@pytest.mark.parametrize("coriolis", [False, True], ids=["", "coriolis"])
@pytest.mark.parametrize("visc", VISC)
@pytest.mark.parametrize("conv", CONV)
def test_ipcs_forms(case, conv, visc, coriolis):
    import nsb200._lib as L
    dev, asm, d = case["dev"], case["asm"], case["d"]
    us, un, um, p = case["fields"]
    _set(case, conv, visc, coriolis)
    try:
        b = dev.empty(len(us))
        dev.assemble_ipcs_momentum(L.VEC, up(dev, us), up(dev, un), up(dev, um), up(dev, p), b=b)
        ref = asm.ipcs_momentum_residual(us, un, um, p, COEF["alpha"], COEF["k"], COEF["c_conv"], COEF["c_p"], COEF["c_visc"])
        assert rel_max_err(b.cpu().numpy(), ref) < TOL
        pat = dev.patterns["P2P2"]
        for newton in (True, False):
            A = dev.empty(pat.nnz * d * d)
            dev.assemble_ipcs_momentum(L.MAT | (0 if newton else L.PICARD), up(dev, us), A=A)
            got = bsr_to_csr(pat.rowptr_host, pat.col_host, A.cpu().numpy(), d)
            refJ = asm.ipcs_momentum_jacobian(us, COEF["alpha"][0], COEF["k"], COEF["c_conv"], COEF["c_visc"], newton=newton)
            assert np.abs((got - refJ).tocoo().data).max() / np.abs(refJ.data).max() < TOL
    finally:
        _set(case, "standard", "reduced", False)


@pytest.mark.parametrize("conv,visc,coriolis", [("rotational", "reduced", False), ("divergence", "traction", True),
                                                ("skew_symmetric", "traction", False), ("standard", "traction", True)])
def test_mixed_forms(case, conv, visc, coriolis):
    import nsb200._lib as L
    dev, asm = case["devW"], case["asm"]
    us, un, um, p = case["fields"]
    osp = asm.sp
    w = [np.zeros(osp.ndofs_W) for _ in range(3)]
    for wi, ui in zip(w, (us, un, um)):
        wi[osp.V_to_W] = ui
        wi[osp.Q_to_W] = p
    _set(case, conv, visc, coriolis)
    try:
        b = dev.empty(osp.ndofs_W)
        dev.assemble_mixed(L.VEC | L.BDF, up(dev, w[0]), up(dev, w[1]), up(dev, w[2]), b=b)
        ref = asm.mixed_residual(w[0], COEF["c_conv"], COEF["c_p"], COEF["c_visc"], bdf=(COEF["alpha"], COEF["k"], w[1], w[2]))
        assert rel_max_err(b.cpu().numpy(), ref) < TOL
        pat = dev.patterns["WW"]
        for newton in (True, False):
            A = dev.empty(pat.nnz)
            dev.assemble_mixed(L.MAT | L.BDF | (0 if newton else L.PICARD), up(dev, w[0]), up(dev, w[1]), up(dev, w[2]), A=A)
            got = sp.csr_matrix((A.cpu().numpy(), pat.col_host, pat.rowptr_host), shape=(osp.ndofs_W,) * 2)
            refJ = asm.mixed_jacobian(w[0], COEF["c_conv"], COEF["c_p"], COEF["c_visc"],
                                      mass_coef=COEF["alpha"][0] / COEF["k"], newton=newton)
            assert np.abs((got - refJ).tocoo().data).max() / np.abs(refJ.data).max() < TOL
    finally:
        _set(case, "standard", "reduced", False)


def test_generic_kernels_equal_tuned_kernels(case):
    """the one-thread-per-row kernels (every weak form) against the tuned row-group / cell kernels on the default
    form: the fallback path of round 1 is now the reference implementation the tuned path is held to"""
    import nsb200._lib as L
    dev, d = case["dev"], case["d"]
    us, un, um, p = (up(dev, a) for a in case["fields"])
    pat = dev.patterns["P2P2"]
    out = {}
    for generic in (False, True):
        dev.set_forms(0, 0, None, generic=generic)
        b, A = dev.empty(us.numel()), dev.empty(pat.nnz * d * d)
        dev.assemble_ipcs_momentum(L.VEC, us, un, um, p, b=b)
        dev.assemble_ipcs_momentum(L.MAT, us, A=A)
        out[generic] = (b.cpu().numpy(), A.cpu().numpy())
    dev.set_forms(0, 0, None)
    assert rel_max_err(out[True][0], out[False][0]) < 1e-13
    assert rel_max_err(out[True][1], out[False][1]) < 1e-13


def test_body_force_load(case):
    from oracle import projection as opj
    dev, asm, d = case["dev"], case["asm"], case["d"]
    f = lambda x: np.stack([np.sin(x[:, 0]) * x[:, 1] + c for c in range(d)], axis=1)      # noqa: E731
    for degree in (1, 2, 3):
        got = dev.cell_rhs(f, "V", degree).cpu().numpy()
        assert rel_max_err(got, opj.rhs_vector(asm, f, "V", degree)) < TOL


@pytest.mark.parametrize("cls", ["ipcs", "bdf"])
def test_solvers_with_every_extra_term(built_library, cls):
    """channel flow with the skew-symmetric convective form, a body force, a traction BC at the outlet (which switches
    the viscous term to its traction form, quirk Q12) and a Coriolis term: 2 steps against the LU oracle"""
    import nsb200
    import oracle
    from oracle import projection as opj
    V, T = nsb200.VelocityBCType, nsb200.TractionBCType
    Re, dt = 20.0, 0.01
    mesh, markers = nsb200.hyper_rectangle((0.0, 0.0), (2.0, 1.0), (8, 4))
    om = oracle.mesh.hyper_rectangle((0.0, 0.0), (2.0, 1.0), (8, 4))

    def inflow(x, t):
        return np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + t), np.zeros(len(x))], axis=1)

    def force(x, t):
        return np.stack([0.3 * np.sin(np.pi * x[:, 1]) * (1.0 + t), -0.2 * x[:, 0]], axis=1)
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
    Solver = nsb200.IPCSSolver if cls == "ipcs" else nsb200.ImplicitBDFSolver
    solver = Solver(mesh, markers, "skew_symmetric", ts)
    co = dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re, body_force_term=0.5, coriolis_term=0.25)
    solver.set_equation_coefficients(co)
    solver.set_boundary_conditions([(V.function, 1, nsb200.Expression(inflow)), (V.no_slip, 3, None), (V.no_slip, 4, None),
                                    (T.constant, 2, (-0.1, 0.02))])
    solver.set_body_force(nsb200.Expression(force, degree=2))
    solver.set_angular_velocity(0.8)
    assert solver._form_viscous_term is nsb200.WeakFormViscousTerm.traction_form
    vel = [(1, None, inflow), (3, None, (0.0, 0.0)), (4, None, (0.0, 0.0))]
    orc = (oracle.solvers.IPCSOracle if cls == "ipcs" else oracle.solvers.BDFOracle)(om, Re, vel, [])
    orc.asm.conv_form, orc.asm.visc_form, orc.asm.coriolis = "skew_symmetric", "traction", (0.0, 0.0, 2.0 * 0.25 * 0.8)
    trac = opj.traction_vector(orc.asm, [(2, None, (-0.1, 0.02), 1)])
    orc.load = lambda t: -0.5 * opj.rhs_vector(orc.asm, lambda x: force(x, t), "V", 2) + trac
    for _ in range(2):
        ts.update_coefficients()
        solver.solve()
        a = list(ts.coefficients(1)) + [0.0]
        orc.solve(tuple(a[:3]), ts.get_next_step_size(), ts.next_time)
        ts.advance_time()
        solver.advance_time()
        orc.advance_time()
    assert solver.newton_iterations == orc.newton_iterations
    if cls == "ipcs":
        assert rel_max_err(solver._velocities[1].cpu().numpy(), orc.velocities[1]) < 1e-8
        p, po = solver._old_pressure.cpu().numpy(), orc.old_pressure
        assert rel_max_err(p - p.mean(), po - po.mean()) < 1e-8
    else:
        w, osp = solver._solutions[1].cpu().numpy(), orc.sp
        assert rel_max_err(w[osp.V_to_W], orc.solutions[1][osp.V_to_W]) < 1e-8
        assert rel_max_err(w[osp.Q_to_W], orc.solutions[1][osp.Q_to_W]) < 1e-8
