"""Oracle restatement of the dolfin calls of SURVEY.md 8f rows 1-3 -- TEST INFRASTRUCTURE.

* ``project(f, V|Q, degree)``   source/ns_solver_base.py:1151,1168 (App. A.7): the Expression is interpolated into
  the degree-k Lagrange element on every cell, ``int u w = int (I_k f) w`` assembled by quadrature, solved by LU
* ``mean_pressure_shift``       source/ns_solver_base.py:1190-1203
* ``errornorm``                 convergence_test/taylor_green_vortex.py:101-119 (dolfin default degree_rise = 3)
* ``cfl_number``                source/ns_problem.py:554-587: DG-k LocalSolver projection of k |u| dt / h, linf of the dofs
* ``vorticity_dg`` / ``pressure_gradient_dg``   source/ns_problem.py:55-101: dolfin.project onto DG_1 / [DG_0]^d, done
  here as a genuine cell-wise L2 projection by quadrature (the device code uses that it is the identity)

Written independently of navierstokes-with-fenics_b200/lagrange.py: Lagrange bases here come from inverting the
monomial Vandermonde matrix at an equispaced lattice, integrals from the Gauss-Jacobi rules of oracle/fem.py, the
local projections from per-cell dense solves.  Only the FIAT default-scheme quadrature points are shared constants.
"""
import itertools

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from . import fem


def _lattice(d, k):
    pts = [a for a in itertools.product(range(k + 1), repeat=d) if sum(a) <= k]
    return np.array(pts, dtype=np.float64) / k


def _monomials(d, k, X):
    exps = [a for a in itertools.product(range(k + 1), repeat=d) if sum(a) <= k]
    return np.stack([np.prod(X ** np.array(e)[None, :], axis=1) for e in exps], axis=1)


class LagrangeElement:
    """degree-k Lagrange element on the UFC simplex, nodes = equispaced lattice"""

    def __init__(self, d, k):
        self.d, self.k = d, k
        self.nodes = _lattice(d, k)
        self._coef = np.linalg.inv(_monomials(d, k, self.nodes))     # monomial coefficients of every basis function

    def tabulate(self, X):
        return _monomials(self.d, self.k, np.atleast_2d(X)) @ self._coef      # [npts, nk]


def _physical(mesh, ref_points):
    lam, _ = fem.barycentric(ref_points)
    return np.einsum("km,cmi->cki", lam, mesh.coords[mesh.cells])


def rhs_vector(asm, f, space, degree):
    """b_I = int (I_k f) phi_I dx in V (interleaved) / Q numbering: right-hand side of ``project`` and, with
    the factor -c_b, the body-force term of source/ns_solver_base.py:158-171"""
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    el = LagrangeElement(d, degree)
    ncomp = d if space == "V" else 1
    F = np.asarray(f(_physical(m, el.nodes).reshape(-1, d)), dtype=np.float64).reshape(m.num_cells, len(el.nodes), ncomp)
    qp, qw = fem.simplex_quadrature(d, degree + 2)
    psi = el.tabulate(qp)
    phi = fem.tabulate_p2(qp)[0] if space == "V" else fem.tabulate_p1(qp)[0]
    be = np.einsum("c,q,qI,qJ,cJk->cIk", asm.detJ, qw, phi, psi, F)
    if space == "V":
        b = np.zeros((s.n_nodes, d))
        np.add.at(b, s.cell_nodes, be)
        return b.reshape(-1)
    b = np.zeros(s.ndofs_Q)
    np.add.at(b, s.cell_dofs_Q, be[:, :, 0])
    return b


def traction_vector(asm, traction_bcs):
    """int_Gamma t . w ds (source/ns_solver_base.py:121-156), facet by facet: ``traction_bcs`` = [(marker, comp or
    None, value, degree)], value a callable x[n,d] -> [n, d] / [n] or a constant (tuple); V numbering"""
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    verts = np.zeros((d + 1, d))
    verts[1:, :] = np.eye(d)
    load = np.zeros((s.n_nodes, d))
    for marker, comp, value, degree in traction_bcs:
        el = LagrangeElement(d, degree)                       # the CELL's P_k interpolant, restricted to the facet
        qp, qw = fem.simplex_quadrature(d - 1, degree + 2)
        for f in np.nonzero(m.bfacet_marker == int(marker))[0]:
            c = int(m.bfacet_cell[f])
            cv = m.cells[c]
            loc = [int(np.nonzero(cv == v)[0][0]) for v in m.bfacets[f]]
            X = m.coords[cv]
            lam_f = np.concatenate([1.0 - qp.sum(axis=1, keepdims=True), qp], axis=1)
            ref = lam_f @ verts[loc]                           # facet quadrature points in the cell's reference frame
            E = X[loc][1:] - X[loc][0]
            meas = np.linalg.norm(E[0]) if d == 2 else np.linalg.norm(np.cross(E[0], E[1]))
            lamc, _ = fem.barycentric(el.nodes)
            xn = lamc @ X                                      # physical lattice nodes of the cell
            ncomp = d if comp is None else 1
            tn = np.asarray(value(xn), dtype=np.float64).reshape(len(xn), ncomp) if callable(value) else \
                np.broadcast_to(np.asarray(value, dtype=np.float64).reshape(1, ncomp), (len(xn), ncomp))
            tq = el.tabulate(ref) @ tn                         # [nq, ncomp]
            phi = fem.tabulate_p2(ref)[0]                      # cell P2 basis on the facet (zero for off-facet nodes)
            be = meas * np.einsum("q,qI,qk->Ik", qw, phi, tq)
            nodes = s.cell_nodes[c]
            if comp is None:
                np.add.at(load, nodes, be)
            else:
                np.add.at(load[:, comp], nodes, be[:, 0])
    return load.reshape(-1)


def project(asm, f, space, degree):
    """L2 projection of the cell-wise degree-`degree` interpolant of f(x[n,d]) onto V (interleaved) or Q"""
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    el = LagrangeElement(d, degree)
    ncomp = d if space == "V" else 1
    F = np.asarray(f(_physical(m, el.nodes).reshape(-1, d)), dtype=np.float64).reshape(m.num_cells, len(el.nodes), ncomp)
    qp, qw = fem.simplex_quadrature(d, degree + 2)
    psi = el.tabulate(qp)                                            # [nq, nk]
    phi = fem.tabulate_p2(qp)[0] if space == "V" else fem.tabulate_p1(qp)[0]
    fq = np.einsum("qJ,cJk->cqk", psi, F)
    be = np.einsum("c,q,qI,cqk->cIk", asm.detJ, qw, phi, fq)         # [nc, nloc, ncomp]
    if space == "V":
        b = np.zeros((s.n_nodes, d))
        np.add.at(b, s.cell_nodes, be)
        M = asm.velocity_mass_matrix()
        return spla.splu(M.tocsc()).solve(b.reshape(-1))
    b = np.zeros(s.ndofs_Q)
    np.add.at(b, s.cell_dofs_Q, be[:, :, 0])
    return spla.splu(asm.pressure_mass_matrix().tocsc()).solve(b)


def mean_pressure_shift(asm, p, target):
    """p - (int p / int 1 - target), re-projected onto Q (a no-op for a P1 function: kept literal)"""
    mean = asm.integral_Q(p) / asm.volume()
    M = asm.pressure_mass_matrix()
    return spla.splu(M.tocsc()).solve(M @ (p - (mean - target)))


def errornorm(asm, exact, uh, space, degree_rise=3):
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    k0, ncomp = (2, d) if space == "V" else (1, 1)
    el = LagrangeElement(d, k0 + degree_rise)
    Fe = np.asarray(exact(_physical(m, el.nodes).reshape(-1, d)), dtype=np.float64).reshape(m.num_cells, len(el.nodes), ncomp)
    basis_at_nodes = fem.tabulate_p2(el.nodes)[0] if space == "V" else fem.tabulate_p1(el.nodes)[0]
    if space == "V":
        Uc = uh.reshape(-1, d)[s.cell_nodes]                         # [nc, nl, d]
    else:
        Uc = uh[s.cell_dofs_Q][:, :, None]
    Fh = np.einsum("JL,cLk->cJk", basis_at_nodes, Uc)
    qp, qw = fem.simplex_quadrature(d, 2 * el.k)
    psi = el.tabulate(qp)
    eq = np.einsum("qJ,cJk->cqk", psi, Fe - Fh)
    return float(np.sqrt(np.einsum("c,q,cqk,cqk->", asm.detJ, qw, eq, eq)))


def fiat_degree4_rule(d):
    """FIAT "default" scheme, quadrature degree 4"""
    if d == 2:
        a, b = 0.09157621350977, 0.44594849091597
        X = np.array([[a, a], [a, 1 - 2 * a], [1 - 2 * a, a], [b, b], [b, 1 - 2 * b], [1 - 2 * b, b]])
        return X, np.array([0.10995174365532] * 3 + [0.22338158967801] * 3) / 2.0
    X, W = [[0.25, 0.25, 0.25]], [-74.0 / 5625.0]
    a, b = 1.0 / 14.0, 11.0 / 14.0
    for i in range(4):
        p = [a] * 4
        p[i] = b
        X.append(p[1:])
        W.append(343.0 / 45000.0)
    a, b = 0.3994035761667992, 0.1005964238332008
    for i, j in itertools.combinations(range(4), 2):
        p = [b] * 4
        p[i] = p[j] = a
        X.append(p[1:])
        W.append(56.0 / 2250.0)
    return np.array(X), np.array(W)


def cfl_number(asm, u, step_size, degree=2):
    """per cell: solve (int v w) c = int f w on DG-2 with f = degree |u| dt / h at the rule's points; return max |c|"""
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    qp, qw = fiat_degree4_rule(d)
    phi = fem.tabulate_p2(qp)[0]                                     # DG-2 basis = P2 nodal basis on the cell
    X = m.coords[m.cells]
    h = np.zeros(m.num_cells)
    for a, b in itertools.combinations(range(d + 1), 2):
        h = np.maximum(h, np.linalg.norm(X[:, a] - X[:, b], axis=1))
    Uc = u.reshape(-1, d)[s.cell_nodes]
    uq = np.einsum("qL,cLk->cqk", phi, Uc)
    f = degree * np.linalg.norm(uq, axis=2) * step_size / h[:, None]
    best = 0.0
    for c in range(m.num_cells):
        M = asm.detJ[c] * (phi.T @ (qw[:, None] * phi))
        rhs = asm.detJ[c] * (phi.T @ (qw * f[c]))
        best = max(best, float(np.abs(np.linalg.solve(M, rhs)).max()))
    return best


def _dg_project(asm, fq, qp, qw, degree):
    """cell-wise L2 projection onto DG-`degree` (dolfin.project onto a discontinuous space is block diagonal): fq
    [nc, nq, ncomp] are the integrand's values at the quadrature points; returns the dofs [nc, ndof, ncomp], dofs =
    the Lagrange nodes of the cell (vertices for degree 1, the single constant for degree 0)"""
    d = asm.sp.mesh.dim
    assert degree in (0, 1)
    psi = np.ones((len(qw), 1)) if degree == 0 else fem.tabulate_p1(qp)[0]     # DG-1 dofs in the cell's vertex order
    Mloc = psi.T @ (qw[:, None] * psi)
    rhs = np.einsum("q,qI,cqk->cIk", qw, psi, fq)               # detJ cancels between the two sides
    return np.linalg.solve(Mloc[None, :, :], rhs)


def vorticity_dg(asm, u):
    """``_compute_vorticity`` source/ns_problem.py:55-81: dolfin.project(curl(u), DG_1) -> [nc, d+1] (2D) / [nc, 4, 3]"""
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    qp, qw = fem.simplex_quadrature(d, 4)
    dphi = fem.tabulate_p2(qp)[1]                                # [nq, nl, d] reference gradients
    K, _ = fem.geometry(m.coords, m.cells)
    Uc = u.reshape(-1, d)[s.cell_nodes]                          # [nc, nl, d]
    # grad[c, q, b, j] = d_j u_b = sum_L U_Lb sum_k dphi[q, L, k] K[c, k, j]
    grad = np.einsum("cLb,qLk,ckj->cqbj", Uc, dphi, K)
    if d == 2:
        curl = (grad[:, :, 1, 0] - grad[:, :, 0, 1])[:, :, None]
    else:
        curl = np.stack([grad[:, :, 2, 1] - grad[:, :, 1, 2], grad[:, :, 0, 2] - grad[:, :, 2, 0],
                         grad[:, :, 1, 0] - grad[:, :, 0, 1]], axis=2)
    out = _dg_project(asm, curl, qp, qw, 1)
    return out[:, :, 0] if d == 2 else out


def pressure_gradient_dg(asm, p):
    """``_compute_pressure_gradient`` source/ns_problem.py:83-101: dolfin.project(grad(p), [DG_0]^d) -> [nc, d]"""
    s, m = asm.sp, asm.sp.mesh
    d = m.dim
    qp, qw = fem.simplex_quadrature(d, 2)
    dphi = fem.tabulate_p1(qp)[1]
    K, _ = fem.geometry(m.coords, m.cells)
    Pc = p[s.cell_dofs_Q]
    grad = np.einsum("cL,qLk,ckj->cqj", Pc, dphi, K)
    return _dg_project(asm, grad, qp, qw, 0)[:, 0, :]


__all__ = ["project", "mean_pressure_shift", "errornorm", "cfl_number", "vorticity_dg", "pressure_gradient_dg",
           "LagrangeElement", "sp"]
