"""The drop-in boundary as the north star words it: the reference's OWN solver class (source/ns_ipcs_solver.py, imported
unchanged from /root/reference) runs on the dolfin-named compat layer, and the calls that reach the C-ABI are the ones
the nsb200 mirror class makes for the same scenario -- the channel flow of the reference's tests/test_ipcs_solver.py:14-54
(mesh 10 n x n on [0,10] x [0,1], Re = 10, inlet 6 y (1 - y), no-slip walls, p = 0 at the outlet, dt = 0.002).

Runs where /root/reference exists (the build container); the device is a recording fake (tests/recording_device.py), so
what is compared is the sequence of device calls, their scalar arguments and the identity pattern of their buffers.
The same compat objects on the real device are exercised by tests/test_gpu_compat.py."""
import os
import sys

import numpy as np
import pytest

REF = "/root/reference/source"
pytestmark = pytest.mark.skipif(not os.path.isdir(REF), reason="the reference source tree is only present in the build container")


@pytest.fixture()
def compat(built_library):
    from recording_device import RecordingDevice
    import nsb200
    from nsb200 import compat as C
    devices = []

    def factory(mesh):
        devices.append(RecordingDevice(mesh))
        return devices[-1]
    D = C.install(device_factory=factory)
    sys.path.insert(0, REF)
    try:
        yield D, devices
    finally:
        sys.path.remove(REF)
        C.uninstall()
        for name in ("ns_ipcs_solver", "ns_bdf_solver", "ns_solver_base", "auxiliary_classes", "auxiliary_methods", "grid_generator",
                     "bdf_time_stepping", "discrete_time", "grid_tools"):
            sys.modules.pop(name, None)
        del nsb200


def _normalise(log):
    """rename buffers by first appearance, so that two runs with different tensors compare equal"""
    names, out = {}, []
    for entry in log:
        row = []
        for a in entry:
            if isinstance(a, str) and a[:1] == "t" and a[1:].isdigit():
                row.append(names.setdefault(a, "b%d" % len(names)))
            else:
                row.append(a)
        out.append(tuple(row))
    return out


def test_reference_ipcs_solver_runs_unchanged_and_makes_the_mirrors_calls(compat):
    D, devices = compat
    import nsb200
    from recording_device import RecordingDevice
    n, Re, dt, steps = 3, 10.0, 0.002, 3
    # ---- the reference's classes, unchanged -----------------------------------------------------------------
    import grid_generator
    from auxiliary_classes import EquationCoefficientHandler
    from bdf_time_stepping import BDFTimeStepping
    from ns_ipcs_solver import IPCSSolver
    from ns_solver_base import VelocityBCType, PressureBCType
    assert IPCSSolver.__module__ == "ns_ipcs_solver" and "/root/reference" in sys.modules["ns_ipcs_solver"].__file__
    mesh, markers = grid_generator.hyper_rectangle((0.0, 0.0), (10.0, 1.0), (10 * n, n))
    M = grid_generator.HyperRectangleBoundaryMarkers
    ts = BDFTimeStepping(0.0, 1.0, desired_start_time_step=dt)
    solver = IPCSSolver(mesh, markers, "standard", ts)
    solver.set_equation_coefficients(EquationCoefficientHandler(Re=Re).equation_coefficients)
    inlet = D.Expression(("6.0*x[1]/h*(1.0-x[1]/h)", "0.0"), h=1.0, degree=2)
    solver.set_boundary_conditions([(PressureBCType.constant, M.right.value, 0.0),
                                    (VelocityBCType.function, M.left.value, inlet),
                                    (VelocityBCType.no_slip, M.bottom.value, None),
                                    (VelocityBCType.no_slip, M.top.value, None)])
    solver.set_initial_conditions({"velocity": (0.0, 0.0), "pressure": 0.0})
    ref_dev = devices[0]
    ref_dev.log.clear()
    for _ in range(steps):
        ts.update_coefficients()
        solver.solve()
        ts.advance_time()
        solver.advance_time()
    w = solver.solution                         # FunctionAssigner: sub -> mixed
    assert w.vector().size() == mesh._m.num_vertices + 2 * mesh.spaces().n_nodes
    ref_log = _normalise(ref_dev.log)
    # ---- the nsb200 mirror on the same scenario, same fake device type -------------------------------------------
    m2, mk2 = nsb200.hyper_rectangle((0.0, 0.0), (10.0, 1.0), (10 * n, n))
    assert np.array_equal(m2.cells, mesh._m.cells) and np.array_equal(m2.bfacet_marker, mesh._m.bfacet_marker)
    ts2 = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
    mir_dev = RecordingDevice(m2)
    mirror = nsb200.IPCSSolver(m2, mk2, "standard", ts2, device=mir_dev)
    mirror.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    mirror.set_boundary_conditions([(P.constant, 2, 0.0),
                                    (V.function, 1, lambda x: np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]), 0.0 * x[:, 0]], axis=1)),
                                    (V.no_slip, 3, None), (V.no_slip, 4, None)])
    mirror.set_initial_conditions(dict(velocity=(0.0, 0.0), pressure=0.0))
    mir_dev.log.clear()
    for _ in range(steps):
        ts2.update_coefficients()
        mirror.solve()
        ts2.advance_time()
        mirror.advance_time()
    mir_log = _normalise(mir_dev.log)
    # the compat run adds the sub->mixed copy of `solution`; everything else must be the same calls in the same order
    ref_core = [e for e in ref_log if e[0] != "scatter"]
    for i, (a, b) in enumerate(zip(ref_core, mir_log)):
        assert a == b, (i, a, b)
    assert len(ref_core) == len(mir_log), (len(ref_core), len(mir_log))
    kinds = {e[0] for e in mir_log}
    assert {"assemble_ipcs_momentum", "gmres", "cg", "assemble_poisson", "assemble_correction", "set_coefficients"} <= kinds
    # the coefficients that reached the device: k, alpha (BDF-1 start-up, then BDF-2), c_conv, c_p, c_visc
    coefs = [e[1:] for e in ref_log if e[0] == "set_coefficients"]
    assert coefs[0] == (dt, 1.0, -1.0, 0.0, 1.0, 1.0, 1.0 / Re) and coefs[1] == (dt, 1.5, -2.0, 0.5, 1.0, 1.0, 1.0 / Re)


def _compare_logs(ref_log, mir_log, ignore=("scatter", "gather"), buffers=True):
    """same device calls in the same order, with the same scalar arguments and (``buffers``) the same identity pattern
    of the buffers; the logs are raw (names by first appearance are given after the ignored calls are dropped)"""
    def strip(log):
        return [tuple(a for a in e if not (isinstance(a, str) and a[:1] == "b" and a[1:].isdigit())) for e in log]
    ref_core = _normalise([e for e in ref_log if e[0] not in ignore])
    mir_core = _normalise([e for e in mir_log if e[0] not in ignore])
    if not buffers:
        ref_core, mir_core = strip(ref_core), strip(mir_core)
    for i, (a, b) in enumerate(zip(ref_core, mir_core)):
        assert a == b, (i, a, b)
    assert len(ref_core) == len(mir_core), (len(ref_core), len(mir_core))
    return mir_core


def test_reference_bdf_solver_runs_unchanged_and_makes_the_mirrors_calls(compat):
    """source/ns_bdf_solver.py (monolithic BDF-2, Newton on the mixed space) imported unchanged: channel flow of the
    reference's tests/test_transient_solvers.py:52-78 at a small size"""
    D, devices = compat
    import nsb200
    from recording_device import RecordingDevice
    n, Re, dt, steps = 3, 10.0, 0.01, 3
    import grid_generator
    from auxiliary_classes import EquationCoefficientHandler
    from bdf_time_stepping import BDFTimeStepping
    from ns_bdf_solver import ImplicitBDFSolver
    from ns_solver_base import VelocityBCType, PressureBCType
    assert "/root/reference" in sys.modules["ns_bdf_solver"].__file__
    mesh, markers = grid_generator.hyper_rectangle((0.0, 0.0), (10.0, 1.0), (10 * n, n))
    M = grid_generator.HyperRectangleBoundaryMarkers
    ts = BDFTimeStepping(0.0, 1.0, desired_start_time_step=dt)
    solver = ImplicitBDFSolver(mesh, markers, "standard", ts)
    solver.set_equation_coefficients(EquationCoefficientHandler(Re=Re).equation_coefficients)
    inlet = D.Expression(("6.0*x[1]*(1.0-x[1])*(1.0+0.5*sin(M_PI*time))", "0.0"), time=0.0, degree=2)
    solver.set_boundary_conditions([(PressureBCType.constant, M.right.value, 0.0),
                                    (VelocityBCType.function, M.left.value, inlet),
                                    (VelocityBCType.no_slip, M.bottom.value, None),
                                    (VelocityBCType.no_slip, M.top.value, None)])
    solver.set_initial_conditions({"velocity": (0.0, 0.0), "pressure": 0.0})
    ref_dev = devices[0]
    ref_dev.log.clear()
    for _ in range(steps):
        ts.update_coefficients()
        solver.solve()
        ts.advance_time()
        solver.advance_time()
    assert abs(inlet.time - ts.current_time) < 1e-14          # the reference's _set_time reached the compat Expression
    ref_log = list(ref_dev.log)
    # ---- the nsb200 mirror, same scenario ---------------------------------------------------------------------
    m2, mk2 = nsb200.hyper_rectangle((0.0, 0.0), (10.0, 1.0), (10 * n, n))
    ts2 = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=dt)
    mir_dev = RecordingDevice(m2)
    mirror = nsb200.ImplicitBDFSolver(m2, mk2, "standard", ts2, device=mir_dev)
    mirror.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    inlet2 = nsb200.Expression(lambda x, time: np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]) * (1.0 + 0.5 * np.sin(np.pi * time)),
                                                         0.0 * x[:, 0]], axis=1), time=0.0)
    mirror.set_boundary_conditions([(P.constant, 2, 0.0), (V.function, 1, inlet2), (V.no_slip, 3, None), (V.no_slip, 4, None)])
    mirror.set_initial_conditions(dict(velocity=(0.0, 0.0), pressure=0.0))
    mir_dev.log.clear()
    for _ in range(steps):
        ts2.update_coefficients()
        mirror.solve()
        ts2.advance_time()
        mirror.advance_time()
    core = _compare_logs(ref_log, list(mir_dev.log))
    kinds = {e[0] for e in core}
    assert {"assemble_mixed", "direct_solve", "bc_apply_vector", "bc_apply_matrix", "set_coefficients"} <= kinds
    coefs = [e[1:] for e in ref_log if e[0] == "set_coefficients"]
    assert coefs[0] == (dt, 1.0, -1.0, 0.0, 1.0, 1.0, 1.0 / Re) and coefs[1] == (dt, 1.5, -2.0, 0.5, 1.0, 1.0, 1.0 / Re)


def test_reference_stationary_solver_runs_unchanged_and_makes_the_mirrors_calls(compat):
    """source/ns_solver_base.py StationarySolverBase (Picard -> Newton, NewtonSolver + CustomNonlinearProblem +
    SystemAssembler) imported unchanged: the lid-driven cavity of the reference's tests/test_stationary_solvers.py:35-58"""
    D, devices = compat
    import nsb200
    from recording_device import RecordingDevice
    import grid_generator
    from auxiliary_classes import EquationCoefficientHandler
    from ns_solver_base import StationarySolverBase, VelocityBCType
    n, Re = 4, 10.0
    # residual norms the fake device reports: start, Picard (2 iterations to 1e-3 < 1e-2), Newton (2 iterations), final check
    script = (1.0, 1.0, 1e-1, 1e-3, 1e-3, 1e-7, 1e-12, 1e-12)
    mesh, markers = grid_generator.hyper_cube(2, n)
    Mk = grid_generator.HyperCubeBoundaryMarkers
    solver = StationarySolverBase(mesh, markers, "standard")
    solver.set_equation_coefficients(EquationCoefficientHandler(Re=Re).equation_coefficients)
    solver.set_boundary_conditions([(VelocityBCType.no_slip, Mk.left.value, None),
                                    (VelocityBCType.no_slip, Mk.right.value, None),
                                    (VelocityBCType.no_slip, Mk.bottom.value, None),
                                    (VelocityBCType.constant, Mk.top.value, (1.0, 0.0))])
    ref_dev = mesh._dev()
    ref_dev._script, ref_dev._k = script, 0
    ref_dev.log.clear()
    solver.solve()
    ref_log = list(ref_dev.log)
    m2, mk2 = nsb200.hyper_cube(2, n)
    mir_dev = RecordingDevice(m2, norm_script=script)
    mirror = nsb200.StationarySolver(m2, mk2, "standard", device=mir_dev)
    mirror.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=1.0 / Re))
    V = nsb200.VelocityBCType
    mirror.set_boundary_conditions([(V.no_slip, 1, None), (V.no_slip, 2, None), (V.no_slip, 3, None),
                                    (V.constant, 4, (1.0, 0.0))])
    mir_dev.log.clear()
    mirror.solve()
    assert mirror.newton_iterations == [2, 2] and mesh._engine.newton_iterations[-2:] == [2, 2]
    # same device calls with the same flags and scalars; the reference keeps its own residual vector for the first and
    # the last F call (dlfn.Vector copy), so the buffer identities are not compared here
    core = _compare_logs(ref_log, list(mir_dev.log), ignore=("scatter", "gather", "copy"), buffers=False)
    kinds = {e[0] for e in core}
    assert {"assemble_mixed", "direct_solve", "set_mixed_bc", "set_coefficients"} <= kinds


def test_forms_outside_the_vocabulary_are_refused(compat):
    """an edited form must not silently run a different equation"""
    D, _ = compat
    import grid_generator
    mesh, markers = grid_generator.hyper_cube(2, 2)
    cell = mesh.ufl_cell()
    W = D.FunctionSpace(mesh, D.MixedElement([D.VectorElement("CG", cell, 2), D.FiniteElement("CG", cell, 1)]))
    Vh, Qh = W.sub(0).collapse(), W.sub(1).collapse()
    u, p, w = D.Function(Vh), D.Function(Qh), D.TestFunction(Vh)
    dV = D.Measure("dx", domain=mesh)
    k = D.Constant(0.1)
    F = (D.dot(u, w) / k + D.dot(D.dot(D.grad(u), u), w) - D.div(w) * p + D.inner(D.grad(u), D.grad(w))
         + D.dot(D.dot(D.grad(u), w), u)) * dV                                   # last term: not in the reference forms
    with pytest.raises(NotImplementedError):
        D.NonlinearVariationalSolver(D.NonlinearVariationalProblem(F, u, [], D.derivative(F, u)))
    with pytest.raises(NotImplementedError):
        D.FunctionSpace(mesh, D.FiniteElement("DG", cell, 1))


@pytest.mark.parametrize("dim,n", [(2, 4), (3, 3)])
def test_reference_grid_generator_marks_what_the_mirror_marks(compat, dim, n):
    """source/grid_generator.py hyper_cube / hyper_rectangle, imported unchanged and run on the compat layer (its
    CompiledSubDomain strings decide which exterior facet gets which id), against nsb200.grid: same vertices, same cells in
    the same order, same boundary id on every exterior facet -- the marker conventions of the mirror are the reference's own"""
    D, _ = compat
    import nsb200
    import grid_generator
    mesh, markers = grid_generator.hyper_cube(dim, n)
    m2, _ = nsb200.hyper_cube(dim, n)
    assert np.array_equal(mesh._m.coords, m2.coords) and np.array_equal(mesh._m.cells, m2.cells)
    assert np.array_equal(mesh._m.bfacets, m2.bfacets)
    ref_ids = np.array([markers[f] for f in D.facets(mesh) if f.exterior()])
    assert np.array_equal(ref_ids, m2.bfacet_marker)
    Mk = grid_generator.HyperCubeBoundaryMarkers
    assert set(int(i) for i in np.unique(ref_ids)) == {Mk.left.value, Mk.right.value, Mk.bottom.value, Mk.top.value} | \
        ({Mk.back.value, Mk.front.value} if dim == 3 else set())
    p1 = (2.0, 1.0) if dim == 2 else (2.0, 1.0, 0.5)
    nn = (2 * n, n) if dim == 2 else (2 * n, n, 2)
    mesh, markers = grid_generator.hyper_rectangle((0.0,) * dim, p1, nn)
    m2, _ = nsb200.hyper_rectangle((0.0,) * dim, p1, nn)
    assert np.array_equal(mesh._m.cells, m2.cells)
    assert np.array_equal(np.array([markers[f] for f in D.facets(mesh) if f.exterior()]), m2.bfacet_marker)
