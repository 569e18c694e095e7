"""Oracle restatement of the drag / lift post-processing of ``demo/dfg_benchmark.py:44-66``:
    traction = -p n + c_visc * 0.5 (grad u + grad u^T) n ,   F_i = assemble(-traction[i] * ds(id))
by Gauss quadrature on every marked exterior facet (degree 4: exact, the integrand is quadratic at
most along an affine facet).  Generic and slow on purpose: it shares nothing with the closed-form
CUDA kernel (csrc/functionals.cu).  Test infrastructure only (see ``oracle/__init__.py``).
"""
import numpy as np

from . import fem


def boundary_force(spaces, u, p, marker, c_visc):
    """u in V numbering (interleaved), p in Q numbering; returns F [dim]."""
    m = spaces.mesh
    d = m.dim
    qp, qw = fem.simplex_quadrature(d - 1, 4)            # on the reference facet
    verts = np.zeros((d + 1, d))
    verts[1:, :] = np.eye(d)
    cell_nodes = spaces.cell_nodes
    U = np.asarray(u).reshape(-1, d)
    F = np.zeros(d)
    sel = np.nonzero(m.bfacet_marker == int(marker))[0]
    for f in sel:
        c = int(m.bfacet_cell[f])
        cv = m.cells[c]
        fv = m.bfacets[f]
        loc = [int(np.nonzero(cv == v)[0][0]) for v in fv]            # local vertex numbers of the facet
        opp = [i for i in range(d + 1) if i not in loc][0]
        X = m.coords[cv]
        # reference points of the facet quadrature inside the cell
        lam_f = np.concatenate([1.0 - qp.sum(axis=1, keepdims=True), qp], axis=1)     # [nq, d]
        ref = lam_f @ verts[loc]                                                        # [nq, d]
        K, detJ = fem.geometry(m.coords, cv[None, :])
        K = K[0]
        # facet measure and outward unit normal from the physical facet vertices
        E = X[loc][1:] - X[loc][0]
        if d == 2:
            t = E[0]
            meas = np.linalg.norm(t)
            n = np.array([t[1], -t[0]]) / meas
            ref_meas = 1.0
        else:
            nn = np.cross(E[0], E[1])
            meas = 0.5 * np.linalg.norm(nn)
            n = nn / np.linalg.norm(nn)
            ref_meas = 0.5
        if np.dot(n, X[opp] - X[loc[0]]) > 0.0:
            n = -n
        phi2, dphi2 = fem.tabulate_p2(ref)
        phi1, _ = fem.tabulate_p1(ref)
        grad_phys = np.einsum("qLj,ji->qLi", dphi2, K)                  # d phi_L / d x_i
        Uc = U[cell_nodes[c]]                                          # [nloc, d]
        G = np.einsum("Lb,qLi->qbi", Uc, grad_phys)                    # grad u [q, b, i]
        sym = 0.5 * (G + np.swapaxes(G, 1, 2))
        pq = phi1 @ np.asarray(p)[cv]
        trac = -pq[:, None] * n[None, :] + c_visc * np.einsum("qbi,i->qb", sym, n)
        F += -(qw[:, None] * trac).sum(axis=0) * meas / ref_meas
    return F
