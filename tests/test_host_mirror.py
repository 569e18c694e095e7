"""Host-side mirror of the reference interface (meshes, spaces, Dirichlet tables, BC tuple
format) against the independently written oracle equivalents.  Bit-exact (integer / index work)."""
import numpy as np
import pytest

import nsb200
import oracle

CASES = [((0.0, 0.0), (2.0, 1.0), (5, 3)), ((0.0, 0.0, 0.0), (1.0, 2.0, 0.5), (2, 3, 2))]


@pytest.mark.parametrize("box", CASES, ids=["2d", "3d"])
def test_mesh_spaces_and_bc_tables(box):
    m, markers = nsb200.hyper_rectangle(*box)
    om = oracle.mesh.hyper_rectangle(*box)
    assert np.array_equal(m.coords, om.coords) and np.array_equal(m.cells, om.cells)
    assert np.array_equal(m.edges, om.edges) and np.array_equal(m.cell_edges, om.cell_edges)
    assert markers is m.bfacet_marker and m.boundary_ids() == om.boundary_marker_ids()
    key = lambda f, mk: sorted(zip(map(tuple, f.tolist()), mk.tolist()))   # noqa: E731
    assert key(m.bfacets, m.bfacet_marker) == key(om.bfacets, om.bfacet_marker)
    sp, osp = nsb200.TaylorHoodSpaces(m), oracle.dofmap.Spaces(om)
    assert np.array_equal(sp.cell_nodes, osp.cell_nodes) and np.array_equal(sp.cell_dofs_W, osp.cell_dofs_W)
    assert np.array_equal(sp.cell_dofs_V, osp.cell_dofs_V)
    assert np.array_equal(sp.V_to_W(), osp.V_to_W) and np.array_equal(sp.Q_to_W(), osp.Q_to_W)
    assert np.array_equal(sp.node_coordinates(), osp.node_coords())
    d = m.dim

    def inflow(x):
        u = np.zeros((len(x), d))
        u[:, 0] = 6.0 * x[:, 1] * (1.0 - x[:, 1])
        return u
    bcs = [(1, None, inflow), (3, None, (0.0,) * d), (4, 0, 1.0), (4, 1, lambda x: x[:, 0] ** 2), (2, "p", 0.5)]
    for space in ("V", "Q", "W"):
        dofs, vals = sp.dirichlet_table(bcs, space)
        obcs = [(space, mk, c, v) for mk, c, v in bcs if (c == "p") == (space == "Q") or space == "W"]
        odofs, ovals = oracle.bc.dirichlet_values(osp, obcs)
        assert np.array_equal(dofs, odofs) and np.array_equal(vals, ovals), space


def test_bc_tuple_translation_matches_reference_format():
    """same tuples as the reference demos/tests (demo/cavity_flow.py:26-29,
    tests/test_ipcs_solver.py:27-37): validated against the markers, last wins."""
    V, P = nsb200.VelocityBCType, nsb200.PressureBCType
    m, markers = nsb200.hyper_cube(2, 4)
    base = nsb200.solver_base.SolverBase(m, markers, "standard")
    base.set_boundary_conditions([(V.no_slip, 1, None), (V.no_slip, 2, None), (V.no_slip, 3, None),
                                  (V.constant, 4, (1.0, 0.0)), (P.constant, 2, 0.0),
                                  (V.constant_component, 3, 1, 0.0), (V.no_normal_flux, 1, None)])
    assert base._velocity_bcs[3] == (4, None, (1.0, 0.0)) and base._pressure_bcs == [(2, "p", 0.0)]
    assert base._velocity_bcs[-1] == (1, 0, 0.0)          # planar normal of x = 0 is axis 0
    with pytest.raises(AssertionError):
        base.set_boundary_conditions([(V.no_slip, 9, None)])       # unknown boundary id
    # every weak-form variant of the reference constructor is accepted (source/ns_solver_base.py:78-115)
    W, Wv, T = nsb200.WeakFormConvectiveTerm, nsb200.WeakFormViscousTerm, nsb200.TractionBCType
    for name, form in (("standard", W.standard_form), ("rotational", W.rotational_form), ("divergence", W.divergence_form),
                       ("skew_symmetric", W.skew_symmetric_form)):
        assert nsb200.solver_base.SolverBase(m, markers, name)._form_convective_term is form
    other = nsb200.solver_base.SolverBase(m, markers, "standard", "traction")
    assert other._form_viscous_term is Wv.traction_form
    # a traction BC switches the viscous term to the traction form (quirk Q12, source/ns_solver_base.py:823-825)
    base2 = nsb200.solver_base.SolverBase(m, markers, "standard")
    base2.set_boundary_conditions([(V.no_slip, 3, None), (T.constant, 2, (1.0, 0.0)), (T.constant_component, 4, 1, 0.5)])
    assert base2._form_viscous_term is Wv.traction_form and base2._traction_bcs == [(2, None, (1.0, 0.0)), (4, 1, 0.5)]
    base.set_equation_coefficients(dict(convective_term=1.0, viscous_term=0.01))
    with pytest.raises(AssertionError):
        base.set_equation_coefficients(dict(viscous_term=-1.0))
    assert nsb200.solver_base.newton_converged(1e-11, 1.0, 1e-10, 1e-9)
    assert nsb200.solver_base.newton_converged(1e-3, 1e7, 1e-10, 1e-9)
    assert not nsb200.solver_base.newton_converged(1e-3, 1.0, 1e-10, 1e-9)


def test_ipcs_newton_forcing_and_update_guess_host_logic(built_library):
    """Host logic of the inexact first Newton update and of the extrapolated update guess (ipcs.py), on the recording device
    with scripted residual norms r0 = 1, r1 = 1e-4, r2 = 1e-13 per step: step 1 solves both updates to linear_rtol from a
    zero guess; from step 2 on the first update is solved to 0.001 x min(rho r0, r1 of the previous step) / r0 = 1e-7 from the
    previous step's update, the second exactly; a state restore drops both predictors."""
    import nsb200
    from recording_device import RecordingDevice
    m, mk = nsb200.hyper_cube(2, 3)
    ts = nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=0.01)
    dev = RecordingDevice(m, norm_script=(1.0, 1e-4, 1e-13))
    s = nsb200.IPCSSolver(m, mk, "standard", ts, device=dev)
    s.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.1))
    V = nsb200.VelocityBCType
    s.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3)] + [(V.constant, 4, (1.0, 0.0))])

    def step():
        dev.log.clear()
        ts.update_coefficients()
        s.solve()
        ts.advance_time()
        s.advance_time()
        return [(e[8], e[9], e[11]) for e in dev.log if e[0] == "gmres"]          # (rtol, atol, x0_is_zero)
    first = step()
    assert first == [(1e-12, 1e-12, True), (1e-12, 1e-12, True)]
    second = step()
    assert second[0][2] is False and second[1][2] is True                          # guess only for the first update
    assert abs(second[0][0] - 1e-7) < 1e-20 and abs(second[0][1] - 1e-7) < 1e-20 and second[1][:2] == (1e-12, 1e-12)
    assert s.newton_iterations == [2, 2]
    state = s.get_state()
    s.set_state(state)
    assert step() == first                                                         # predictors dropped by the restore
    # forcing off: every update exact, the guess stays
    s2 = nsb200.IPCSSolver(m, mk, "standard", nsb200.BDFTimeStepping(0.0, 1.0, order=2, desired_start_time_step=0.01),
                           device=RecordingDevice(m, norm_script=(1.0, 1e-4, 1e-13)), newton_forcing=0.0,
                           extrapolate_newton_update=0)
    s2.set_equation_coefficients(dict(convective_term=1.0, pressure_term=1.0, viscous_term=0.1))
    s2.set_boundary_conditions([(V.no_slip, w, None) for w in (1, 2, 3)] + [(V.constant, 4, (1.0, 0.0))])
    for _ in range(2):
        s2._time_stepping.update_coefficients()
        s2.solve()
        s2._time_stepping.advance_time()
        s2.advance_time()
    assert all(e[8] == 1e-12 and e[11] is True for e in s2.dev.log if e[0] == "gmres")
