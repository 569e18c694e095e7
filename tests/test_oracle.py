"""Pins of the CPU oracle (test infrastructure) against what CAN be pinned without FEniCS
("parity unpinned" against dolfin itself): exact element integrals, the size
formulas of App. C, dolfin's mesh / marker / last-wins BC conventions (App. A), the Poiseuille
steady state (reference tests/test_stationary_solvers.py:173-175), Jacobian = d(residual), and
the plain-C restatement against the numpy one."""
import numpy as np
import pytest

import oracle
from oracle import fem
from oracle.native import CAssembler


def spaces(dim, n, p1=None):
    p0 = (0.0,) * dim
    p1 = p1 or (1.0,) * dim
    n = (n,) * dim if isinstance(n, int) else n
    return oracle.dofmap.Spaces(oracle.mesh.hyper_rectangle(p0, p1, n))


def test_reference_element_integrals():
    # P2 mass on the unit triangle, UFC edge order (edge i opposite vertex i)
    M = np.array([[6, -1, -1, -4, 0, 0], [-1, 6, -1, 0, -4, 0], [-1, -1, 6, 0, 0, -4],
                  [-4, 0, 0, 32, 16, 16], [0, -4, 0, 16, 32, 16], [0, 0, -4, 16, 16, 32]]) / 360.0
    qp, qw = fem.simplex_quadrature(2, 4)
    phi, _ = fem.tabulate_p2(qp)
    assert np.allclose(np.einsum("q,qi,qj->ij", qw, phi, phi), M, rtol=0, atol=1e-15)
    for d in (2, 3):
        qp, qw = fem.simplex_quadrature(d, 6)
        assert abs(qw.sum() - (0.5 if d == 2 else 1.0 / 6.0)) < 1e-15
        phi, dphi = fem.tabulate_p2(qp)
        assert np.allclose(phi.sum(axis=1), 1.0, atol=1e-14) and np.allclose(dphi.sum(axis=1), 0.0, atol=1e-13)
        assert abs(np.einsum("q,qi,qj->", qw, phi, phi) - qw.sum()) < 1e-15      # tet: 1/6
        # nodal basis: phi_i(x_j) = delta_ij at vertices and edge midpoints
        assert np.allclose(fem.tabulate_p2(fem.p2_node_points(d))[0], np.eye(phi.shape[1]), atol=1e-15)
        # quadrature exact to degree 6: int x^a y^b (z^c) = a! b! c! / (a+b+c+d)!
        from math import factorial as f
        for e in ((6, 0, 0), (2, 3, 1), (1, 1, 4), (0, 5, 0)):
            e = e[:d]
            num = np.prod([f(k) for k in e]) / f(sum(e) + d)
            assert abs((qw * np.prod(qp ** np.array(e), axis=1)).sum() - num) < 1e-16 + 1e-13 * num


@pytest.mark.parametrize("n", [3, 5])
def test_size_formulas_appendix_c(n):
    s = spaces(2, n)
    assert s.mesh.num_cells == 2 * n * n and s.N0 == (n + 1) ** 2 and s.N1 == 3 * n * n + 2 * n
    assert s.n_nodes == (2 * n + 1) ** 2
    rp, _ = oracle.dofmap.sparsity_pattern(s.cell_nodes, s.cell_nodes, s.n_nodes, s.n_nodes)
    assert rp[-1] == 46 * n * n + 16 * n + 1
    rp, _ = oracle.dofmap.sparsity_pattern(s.cell_dofs_Q, s.cell_dofs_Q, s.N0, s.N0)
    assert rp[-1] == 7 * n * n + 6 * n + 1
    s = spaces(3, n)
    assert s.mesh.num_cells == 6 * n ** 3 and s.N0 == (n + 1) ** 3 and s.N1 == 7 * n ** 3 + 9 * n ** 2 + 3 * n
    nnz22 = 230 * n ** 3 + 138 * n ** 2 + 24 * n + 1
    nnz21 = 65 * n ** 3 + 57 * n ** 2 + 15 * n + 1
    nnz11 = 15 * n ** 3 + 21 * n ** 2 + 9 * n + 1
    rp, _ = oracle.dofmap.sparsity_pattern(s.cell_nodes, s.cell_nodes, s.n_nodes, s.n_nodes)
    assert rp[-1] == nnz22
    rp, _ = oracle.dofmap.sparsity_pattern(s.cell_dofs_Q, s.cell_dofs_Q, s.N0, s.N0)
    assert rp[-1] == nnz11
    rp, _ = oracle.dofmap.sparsity_pattern(s.cell_dofs_V, s.cell_dofs_V, s.ndofs_V, s.ndofs_V)
    assert rp[-1] == 9 * nnz22
    rp, col = oracle.dofmap.sparsity_pattern(s.cell_dofs_W, s.cell_dofs_W, s.ndofs_W, s.ndofs_W)
    assert rp[-1] == 9 * nnz22 + 6 * nnz21 + nnz11 and s.ndofs_W == 3 * (2 * n + 1) ** 3 + (n + 1) ** 3
    assert all(np.all(np.diff(col[rp[r]:rp[r + 1]]) > 0) for r in range(0, s.ndofs_W, 97))   # ascending columns


def test_mesh_conventions_appendix_a1():
    m = oracle.mesh.hyper_rectangle((0.0, 0.0), (2.0, 1.0), (3, 2))
    nx = 3
    assert np.allclose(m.coords[1 * (nx + 1) + 2], (2 * 2.0 / 3, 0.5))
    v0 = 1 * (nx + 1) + 1
    k = 1 * nx + 1
    assert list(m.cells[2 * k]) == [v0, v0 + 1, v0 + nx + 2] and list(m.cells[2 * k + 1]) == [v0, v0 + nx + 1, v0 + nx + 2]
    m = oracle.mesh.hyper_rectangle((0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (2, 2, 2))
    o = 9
    v = [0, 1, 3, 4, o, o + 1, o + 3, o + 4]       # v0..v7 of the first cube
    tets = [(v[0], v[1], v[3], v[7]), (v[0], v[1], v[7], v[5]), (v[0], v[5], v[7], v[4]),
            (v[0], v[3], v[2], v[7]), (v[0], v[6], v[4], v[7]), (v[0], v[2], v[6], v[7])]
    assert [list(c) for c in m.cells[:6]] == [sorted(t) for t in tets]
    assert np.all(np.diff(m.cells, axis=1) > 0)                       # ordered mesh
    # markers: left 1 right 2 bottom 3 top 4 back 5 front 6, every boundary facet exactly one id
    assert m.boundary_marker_ids() == {1, 2, 3, 4, 5, 6}
    x = m.coords[m.bfacets]
    for ax, lo, hi in ((0, 1, 2), (1, 3, 4), (2, 5, 6)):
        assert np.all(m.bfacet_marker[np.all(x[:, :, ax] == 0.0, axis=1)] == lo)
        assert np.all(m.bfacet_marker[np.all(x[:, :, ax] == 1.0, axis=1)] == hi)
    # edges numbered lexicographically by (min vertex, max vertex)
    key = m.edges[:, 0].astype(np.int64) * m.num_vertices + m.edges[:, 1]
    assert np.all(np.diff(key) > 0) and np.all(m.edges[:, 0] < m.edges[:, 1])


def test_dirichlet_last_wins_and_closure():
    s = spaces(2, 4)
    lid = [("V", w, None, (0.0, 0.0)) for w in (1, 2, 3)] + [("V", 4, None, (1.0, 0.0))]
    dofs, vals = oracle.bc.dirichlet_values(s, lid)
    x = s.dof_coords_V()[dofs]
    top_x = (dofs % 2 == 0) & (x[:, 1] == 1.0)
    assert np.all(vals[top_x] == 1.0) and np.all(vals[~top_x] == 0.0)   # corners belong to the lid (last in list)
    assert len(dofs) == 2 * 4 * 8                                       # 4 sides x 2n P2 nodes x 2 comps (closure)
    dofs2, vals2 = oracle.bc.dirichlet_values(s, lid[::-1])
    assert np.all(vals2[(x[:, 0] == 0.0) | (x[:, 0] == 1.0)] == 0.0)    # walls listed last now win the corners


@pytest.mark.parametrize("dim", [2, 3])
def test_global_identities_and_c_oracle(dim):
    s = spaces(dim, 3, (2.0, 1.0) if dim == 2 else (1.5, 1.0, 0.5))
    A1, A2 = oracle.forms.Assembler(s), CAssembler(s)
    vol = 2.0 if dim == 2 else 0.75
    assert abs(A1.volume() - vol) < 1e-14
    K, M1, M2 = A1.poisson_matrix(), A1.pressure_mass_matrix(), A1.velocity_mass_matrix()
    assert np.abs(K @ np.ones(s.ndofs_Q)).max() < 1e-13
    assert abs(M1.sum() - vol) < 1e-13 and abs(M2.sum() - dim * vol) < 1e-13
    rng = np.random.default_rng(0)
    u, u1, u2 = (rng.uniform(-1, 1, s.ndofs_V) for _ in range(3))
    p, q = rng.uniform(-1, 1, s.ndofs_Q), rng.uniform(-1, 1, s.ndofs_Q)
    al, k, cc, cp, cv = (1.5, -2.0, 0.5), 1e-2, 1.0, 1.0, 0.02

    def close(a, b):
        d = (a - b)
        err = np.abs(d.data).max() if hasattr(d, "data") and not isinstance(d, np.ndarray) else np.abs(d).max()
        ref = np.abs(a.data).max() if hasattr(a, "tocoo") else np.abs(a).max()
        assert err < 1e-13 * ref
    close(A1.ipcs_momentum_residual(u, u1, u2, p, al, k, cc, cp, cv), A2.ipcs_momentum_residual(u, u1, u2, p, al, k, cc, cp, cv))
    for newton in (True, False):
        close(A1.ipcs_momentum_jacobian(u, al[0], k, cc, cv, newton), A2.ipcs_momentum_jacobian(u, al[0], k, cc, cv, newton))
    close(K, A2.poisson_matrix())
    close(M1, A2.pressure_mass_matrix())
    close(M2, A2.velocity_mass_matrix())
    close(A1.poisson_rhs(p, u, al[0], k), A2.poisson_rhs(p, u, al[0], k))
    close(A1.correction_rhs(u, p, q, al[0], k), A2.correction_rhs(u, p, q, al[0], k))
    w, w1, w2 = (rng.uniform(-1, 1, s.ndofs_W) for _ in range(3))
    close(A1.mixed_residual(w, cc, cp, cv, bdf=(al, k, w1, w2)), A2.mixed_residual(w, cc, cp, cv, bdf=(al, k, w1, w2)))
    close(A1.mixed_residual(w, cc, cp, cv), A2.mixed_residual(w, cc, cp, cv))
    close(A1.mixed_jacobian(w, cc, cp, cv, 150.0, True), A2.mixed_jacobian(w, cc, cp, cv, 150.0, True))
    # Jacobian is the derivative of the residual (central differences; the form is quadratic in u)
    v = rng.uniform(-1, 1, s.ndofs_V)
    eps = 1e-4
    fd = (A2.ipcs_momentum_residual(u + eps * v, u1, u2, p, al, k, cc, cp, cv)
          - A2.ipcs_momentum_residual(u - eps * v, u1, u2, p, al, k, cc, cp, cv)) / (2 * eps)
    Jv = A2.ipcs_momentum_jacobian(u, al[0], k, cc, cv, True) @ v
    assert np.abs(fd - Jv).max() < 1e-8 * np.abs(Jv).max()
    # mixed Jacobian: symmetric Stokes part [[A, B], [B^T, 0]] when convection is off
    J = A1.mixed_jacobian(w, 0.0, cp, cv, 0.0, True)
    assert abs(J - J.T).max() < 1e-13 * abs(J).max()


def test_poiseuille_is_a_steady_state():
    """u = (6y(1-y), 0), p = -12/Re (x - L): zero residual on every non-Dirichlet row of the
    stationary form (P2-P1 reproduces it exactly; outflow is the natural condition with p = 0)."""
    Re, L = 10.0, 3.0
    s = spaces(2, (9, 4), (L, 1.0))
    asm = oracle.forms.Assembler(s)
    x = s.node_coords()
    w = np.zeros(s.ndofs_W)
    u = np.stack([6.0 * x[:, 1] * (1.0 - x[:, 1]), np.zeros(len(x))], axis=1).reshape(-1)
    w[s.V_to_W] = u
    w[s.Q_to_W] = -12.0 / Re * (s.mesh.coords[:, 0] - L)
    F = asm.mixed_residual(w, 1.0, 1.0, 1.0 / Re)
    dofs, _ = oracle.bc.dirichlet_values(s, [("W", 1, None, (0.0, 0.0)), ("W", 3, None, (0.0, 0.0)),
                                             ("W", 4, None, (0.0, 0.0))])
    free = np.ones(s.ndofs_W, dtype=bool)
    free[dofs] = False
    assert np.abs(F[free]).max() < 1e-13


def test_ipcs_oracle_lu_and_krylov_agree_and_newton_rule():
    om = oracle.mesh.hyper_rectangle((0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (3, 3, 3))
    vel = [(w, None, (0.0, 0.0, 0.0)) for w in (1, 2, 3, 5, 6)] + [(4, None, (1.0, 0.0, 0.0))]
    a = oracle.solvers.IPCSOracle(om, 100.0, vel, [], assembler="c")
    b = oracle.solvers.IPCSOracle(om, 100.0, vel, [], assembler="numpy")
    b.linear_solver = "krylov"
    for alpha, t in (((1.0, -1.0, 0.0), 0.01), ((1.5, -2.0, 0.5), 0.02)):
        for o in (a, b):
            o.solve(alpha, 0.01, t)
            o.advance_time()
    assert a.newton_iterations == b.newton_iterations
    assert a.newton_history[-1] < 1e-10 or a.newton_history[-1] / a.newton_history[0] < 1e-9
    assert np.abs(a.velocities[1] - b.velocities[1]).max() < 1e-8 * np.abs(a.velocities[1]).max()
    pa, pb = a.old_pressure - a.old_pressure.mean(), b.old_pressure - b.old_pressure.mean()
    assert np.abs(pa - pb).max() < 1e-7 * np.abs(pa).max()
    # first step is BDF-1 from rest, velocities[2] untouched (quirk Q5/Q6)
    assert np.all(a.velocities[2] == a.velocities[2]) and np.abs(a.velocities[1]).max() > 0.5


def test_stationary_oracle_cavity_converges():
    """the reference's own pass criterion (tests/test_stationary_solvers.py: solve_problem runs,
    final residual <= tol, source/ns_solver_base.py:986-988) on the demo cavity, small mesh."""
    om = oracle.mesh.hyper_cube(2, 6)
    vel = [(w, None, (0.0, 0.0)) for w in (1, 2, 3)] + [(4, None, (1.0, 0.0))]
    st = oracle.solvers.StationaryOracle(om, 10.0, vel, [])
    # singular in p (H4): pin one pressure dof like LU pivoting would
    st.bc_dofs = np.append(st.bc_dofs, st.sp.Q_to_W[0])
    st.bc_vals = np.append(st.bc_vals, 0.0)
    order = np.argsort(st.bc_dofs)
    st.bc_dofs, st.bc_vals = st.bc_dofs[order], st.bc_vals[order]
    r = st.solve()
    assert r <= 1e-10 and st.newton_iterations[1] <= 10


def test_c_krylov_matches_numpy_krylov():
    """the C/OpenMP Jacobi-GMRES / CG of the timed CPU arm (nsoracle.c) against their numpy statement"""
    from oracle import krylov as okr
    om = oracle.mesh.hyper_rectangle((0.0, 0.0, 0.0), (1.0, 1.0, 1.0), (3, 3, 3))
    asm = oracle.forms.Assembler(oracle.dofmap.Spaces(om))
    rng = np.random.default_rng(5)
    u = rng.uniform(-1, 1, asm.sp.ndofs_V)
    J = asm.ipcs_momentum_jacobian(u, 1.5, 1e-2, 1.0, 0.01)
    b = rng.uniform(-1, 1, asm.sp.ndofs_V)
    x0, n0 = okr.gmres(J, b, rtol=1e-12)
    x1, n1 = okr.gmres_c(J, b, rtol=1e-12)
    assert n0 == n1 and np.abs(x0 - x1).max() < 1e-10 * np.abs(x0).max()
    L = asm.poisson_matrix()
    bq = rng.uniform(-1, 1, asm.sp.ndofs_Q)
    y0, m0 = okr.pcg(L, bq, rtol=1e-12, nullspace=True)
    y1, m1 = okr.pcg_c(L, bq, rtol=1e-12, nullspace=True)
    assert abs(m0 - m1) <= 1 and np.abs(y0 - y1).max() < 1e-9 * np.abs(y0).max()


def test_lagrange_tables_of_the_product_against_oracle_elements():
    """navierstokes-with-fenics_b200/lagrange.py (exact rational tables, product-formula bases) against the oracle's
    independent construction (Vandermonde inversion + Gauss-Jacobi quadrature)"""
    import nsb200  # noqa: F401
    from nsb200 import lagrange as lg
    from oracle import projection as opj
    for d in (2, 3):
        for k in (1, 2, 3):
            el = opj.LagrangeElement(d, k)
            qp, qw = oracle.fem.simplex_quadrature(d, k + 2)
            # same space: the two bases span the same polynomials -> mass tensors agree after matching the nodes
            pts = lg.lattice_points(d, k)
            match = [int(np.argmin(np.abs(el.nodes - p).sum(axis=1))) for p in pts]
            psi = el.tabulate(qp)[:, match]
            for space, tab in (("P2", oracle.fem.tabulate_p2), ("P1", oracle.fem.tabulate_p1)):
                ref = np.einsum("q,qI,qJ->IJ", qw, tab(qp)[0], psi)
                assert np.abs(lg.cross_mass(d, space, k) - ref).max() < 1e-14
        X, W = lg.simplex_rule(d, 10)
        assert abs(W.sum() - (0.5 if d == 2 else 1.0 / 6.0)) < 1e-15
        Xf, Wf = lg.fiat_default_rule_degree4(d)
        assert abs(Wf.sum() - W.sum()) < 1e-12
        # degree-4 exactness of the FIAT default rule: x0^2 x1^2
        ex = np.einsum("q,q->", W, X[:, 0] ** 2 * X[:, 1] ** 2)
        assert abs(np.einsum("q,q->", Wf, Xf[:, 0] ** 2 * Xf[:, 1] ** 2) - ex) < 1e-12


def test_dg_postprocessing_known_answers():
    """oracle/projection.py vorticity_dg / pressure_gradient_dg (source/ns_problem.py:55-101) on fields whose curl /
    gradient is known in closed form: quadratic velocities (curl linear: reproduced exactly by DG_1), linear pressure"""
    import oracle
    from oracle import projection as P
    for dim, n in ((2, (3, 2)), (3, (2, 2, 2))):
        om = oracle.mesh.hyper_rectangle((0.0,) * dim, (1.5, 1.0, 0.5)[:dim], n)
        asm = oracle.forms.Assembler(oracle.dofmap.Spaces(om))
        x = asm.sp.node_coords()
        X = om.coords[om.cells]
        if dim == 2:
            u = np.stack([x[:, 1] ** 2, x[:, 0] * x[:, 1]], axis=1).reshape(-1)
            ref = -X[:, :, 1]                                                   # d_x (x y) - d_y (y^2)
        else:
            u = np.stack([x[:, 1] ** 2, x[:, 0] * x[:, 2], x[:, 0] * x[:, 1]], axis=1).reshape(-1)
            ref = np.stack([0.0 * X[:, :, 0], -X[:, :, 1], X[:, :, 2] - 2.0 * X[:, :, 1]], axis=2)
        assert np.abs(P.vorticity_dg(asm, u) - ref).max() < 1e-13
        p = 2.0 * x[:asm.sp.N0, 0] - 3.0 * x[:asm.sp.N0, 1]
        g = P.pressure_gradient_dg(asm, p)
        assert np.abs(g[:, 0] - 2.0).max() < 1e-13 and np.abs(g[:, 1] + 3.0).max() < 1e-13
