"""Boundary-traction load vector  int_Gamma t . w ds  (``SolverBase._add_boundary_tractions``,
source/ns_solver_base.py:121-156) on the host: the data live on O(boundary) facets and come from Python
``Expression`` callbacks like the Dirichlet values, so they are integrated here with numpy and
uploaded as one V-numbered vector; nothing of it is per-cell work.

The traction is interpolated into P_k on the facet (the restriction of dolfin's cell-wise P_k interpolant of an
``Expression(degree=k)``), the P2 test functions restricted to a facet are the P2 basis of the (d-1)-simplex; the
product is integrated exactly by a collapsed Gauss rule."""
import numpy as np

from . import lagrange
from .solver_base import Expression


def traction_load(spaces, traction_bcs):
    """``traction_bcs``: [(boundary id, component or None, value)], value a constant / tuple / Expression / callable
    x[n,d] -> values.  Returns the dense load vector in V numbering (interleaved)."""
    m, d = spaces.mesh, spaces.dim
    fd = d - 1
    load = np.zeros((spaces.n_nodes, d))
    edges = lagrange.UFC_EDGES[fd] if fd >= 2 else ((0, 1),)
    for marker, comp, value in traction_bcs:
        fv = m.bfacets[m.bfacet_marker == int(marker)].astype(np.int64)            # [nf, d]
        if len(fv) == 0:
            continue
        k = value.degree if isinstance(value, Expression) else (2 if callable(value) else 1)
        Xq, Wq = lagrange.simplex_rule(fd, k + 2) if fd >= 2 else _gauss_interval(k + 2)
        phi = lagrange.eval_p2(fd, Xq)                                              # [nq, nl_f]
        psi = lagrange.eval_lagrange(fd, k, Xq)                                     # [nq, nk]
        lam = lagrange._bary(lagrange.lattice_points(fd, k))                        # [nk, d]
        X = np.einsum("km,fmi->fki", lam, m.coords[fv])                             # lattice points of every facet
        ncomp = d if comp is None else 1
        if callable(value):
            T = np.asarray(value(X.reshape(-1, d)), dtype=np.float64).reshape(len(fv), lam.shape[0], ncomp)
        else:
            T = np.broadcast_to(np.asarray(value, dtype=np.float64).reshape(1, 1, ncomp), (len(fv), lam.shape[0], ncomp))
        E = m.coords[fv[:, 1:]] - m.coords[fv[:, :1]]
        if d == 2:
            meas = np.linalg.norm(E[:, 0], axis=1)                                  # |J_facet|, reference length 1
        else:
            meas = np.linalg.norm(np.cross(E[:, 0], E[:, 1]), axis=1)               # 2 * area, reference area 1/2
        be = np.einsum("f,q,qI,qK,fKc->fIc", meas, Wq, phi, psi, T)                 # [nf, nl_f, ncomp]
        if fd >= 2 or d == 2:
            enodes = np.stack([m.edge_index(fv[:, a], fv[:, b]) for a, b in edges], axis=1) + spaces.n_vertices
        nodes = np.concatenate([fv, enodes], axis=1)
        if spaces.node_perm is not None:
            nodes = spaces.node_perm[nodes]
        if comp is None:
            np.add.at(load, nodes, be)
        else:
            np.add.at(load[:, comp], nodes, be[:, :, 0])
    return load.reshape(-1)


def _gauss_interval(degree):
    g, w = np.polynomial.legendre.leggauss(degree // 2 + 1)
    return (0.5 * (g + 1.0))[:, None], 0.5 * w
