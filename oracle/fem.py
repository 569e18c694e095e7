"""Oracle finite-element building blocks: UFC reference simplices, P1/P2 Lagrange
bases, collapsed Gauss-Jacobi quadrature, affine geometry.

Test infrastructure only (see ``oracle/__init__.py``).
"""
import numpy as np
from scipy.special import roots_jacobi

from .mesh import UFC_EDGES


def simplex_quadrature(dim, degree):
    """Collapsed (Duffy) Gauss-Jacobi rule on the UFC reference simplex, exact for
    polynomials of total degree <= ``degree``.  Returns (points [nq,dim], weights [nq]),
    weights summing to 1/dim!."""
    n = degree // 2 + 1
    rules = []
    for k in range(dim):
        alpha = dim - 1 - k
        x, w = roots_jacobi(n, alpha, 0.0)
        x = 0.5 * (x + 1.0)
        w = w / 2.0 ** (alpha + 1)
        rules.append((x, w))
    grids = np.meshgrid(*[r[0] for r in rules], indexing="ij")
    wgrids = np.meshgrid(*[r[1] for r in rules], indexing="ij")
    xi = [g.ravel() for g in grids]
    w = np.ones_like(xi[0])
    for g in wgrids:
        w = w * g.ravel()
    pts = np.empty((len(w), dim))
    fac = np.ones_like(w)
    for k in range(dim):
        pts[:, k] = xi[k] * fac
        fac = fac * (1.0 - xi[k])
    return pts, w


def barycentric(points):
    """lambda [nq, d+1] and their constant reference gradients [d+1, d]."""
    points = np.atleast_2d(points)
    d = points.shape[1]
    lam = np.empty((points.shape[0], d + 1))
    lam[:, 0] = 1.0 - points.sum(axis=1)
    lam[:, 1:] = points
    dlam = np.zeros((d + 1, d))
    dlam[0, :] = -1.0
    dlam[1:, :] = np.eye(d)
    return lam, dlam


def tabulate_p1(points):
    """phi [nq, d+1], dphi [nq, d+1, d] (reference gradients)."""
    lam, dlam = barycentric(points)
    return lam, np.broadcast_to(dlam, (lam.shape[0],) + dlam.shape).copy()


def tabulate_p2(points):
    """P2 basis, dofs = vertices then edges in UFC order.
    phi [nq, nloc], dphi [nq, nloc, d]."""
    lam, dlam = barycentric(points)
    nq, nv = lam.shape
    d = nv - 1
    le = UFC_EDGES[d]
    nloc = nv + len(le)
    phi = np.empty((nq, nloc))
    dphi = np.empty((nq, nloc, d))
    for v in range(nv):
        phi[:, v] = lam[:, v] * (2.0 * lam[:, v] - 1.0)
        dphi[:, v, :] = (4.0 * lam[:, v] - 1.0)[:, None] * dlam[v][None, :]
    for e, (a, b) in enumerate(le):
        phi[:, nv + e] = 4.0 * lam[:, a] * lam[:, b]
        dphi[:, nv + e, :] = 4.0 * (lam[:, a][:, None] * dlam[b][None, :]
                                    + lam[:, b][:, None] * dlam[a][None, :])
    return phi, dphi


def p2_node_points(dim):
    """Reference coordinates of the P2 nodes (vertices, then edge midpoints)."""
    verts = np.zeros((dim + 1, dim))
    verts[1:, :] = np.eye(dim)
    le = UFC_EDGES[dim]
    mids = 0.5 * (verts[le[:, 0]] + verts[le[:, 1]])
    return np.concatenate([verts, mids])


def geometry(coords, cells):
    """Affine maps.  Returns K [nc,d,d] with K[c,j,i] = d xi_j / d x_i  and |det J| [nc]."""
    X = coords[cells]                                  # [nc, d+1, d]
    J = np.transpose(X[:, 1:, :] - X[:, :1, :], (0, 2, 1))   # J[c,i,j] = dx_i/dxi_j
    detJ = np.linalg.det(J)
    K = np.linalg.inv(J)
    return K, np.abs(detJ)
