"""Reference-element tables for the projection / diagnostics kernels (csrc/project.cu), host side, setup only.

* Lagrange bases of any degree k on the UFC simplex at the lattice nodes alpha / k (the nodal values of
  dolfin's per-cell interpolation of an ``Expression(degree=k)``)
* exact (rational) cross mass tensors  R_IJ = int phi_I psi^k_J  between the P1 / P2 test bases of the
  solver spaces (UFC order: vertices, then UFC edges) and P_k
* collapsed Gauss-Legendre rules on the simplex, exact to a requested degree (errornorm)
* the FIAT "default" degree-4 rules FFC picks for the CFL projection (``quadrature_degree = 2 * degree``,
  source/ns_problem.py:571) and the per-cell DG-2 LocalSolver projection matrix built from them
"""
import itertools
from fractions import Fraction
from functools import lru_cache
from math import factorial

import numpy as np

UFC_EDGES = {1: ((0, 1),), 2: ((1, 2), (0, 2), (0, 1)), 3: ((2, 3), (1, 3), (1, 2), (0, 3), (0, 2), (0, 1))}


# ---- multi-indices / lattice nodes ------------------------------------------------------------
@lru_cache(maxsize=None)
def lattice(d, k):
    """barycentric multi-indices alpha (|alpha| = k, d+1 entries) of the P_k lattice nodes; the node is
    x = alpha[1:] / k in reference coordinates.  Order: vertices first (so k = 1 is the P1 vertex order), then
    the rest lexicographically -- only internal consistency matters."""
    alls = [a for a in itertools.product(range(k + 1), repeat=d + 1) if sum(a) == k]
    verts = [tuple(k if j == i else 0 for j in range(d + 1)) for i in range(d + 1)]
    rest = sorted(a for a in alls if a not in verts)
    return tuple(verts + rest) if k > 0 else ((0,) * (d + 1),)


def lattice_points(d, k):
    a = np.array(lattice(d, k), dtype=np.float64)
    return a[:, 1:] / max(k, 1) if k > 0 else np.full((1, d), 1.0 / (d + 1))


def _bary(X):
    X = np.asarray(X, dtype=np.float64)
    return np.concatenate([1.0 - X.sum(axis=1, keepdims=True), X], axis=1)


def eval_lagrange(d, k, X):
    """[npts, nk] values of the P_k Lagrange basis (lattice order) at reference points X[npts, d]"""
    lam = _bary(X)
    nodes = lattice(d, k)
    out = np.ones((lam.shape[0], len(nodes)))
    if k == 0:
        return out
    for J, a in enumerate(nodes):
        for j in range(d + 1):
            for m in range(a[j]):
                out[:, J] *= (k * lam[:, j] - m) / (m + 1.0)
    return out


def eval_p2(d, X):
    """[npts, NL2] solver-space P2 basis (vertices, then UFC edges)"""
    lam = _bary(X)
    cols = [lam[:, i] * (2.0 * lam[:, i] - 1.0) for i in range(d + 1)]
    cols += [4.0 * lam[:, a] * lam[:, b] for a, b in UFC_EDGES[d]]
    return np.stack(cols, axis=1)


def eval_p1(d, X):
    return _bary(X)


# ---- exact integration of products of barycentric polynomials -----------------------------------
class _Poly:
    """polynomial in the barycentric coordinates: {exponent tuple (d+1): Fraction}"""

    def __init__(self, terms):
        self.t = {k: v for k, v in terms.items() if v != 0}

    def __mul__(self, o):
        t = {}
        for (k1, v1), (k2, v2) in itertools.product(self.t.items(), o.t.items()):
            k = tuple(a + b for a, b in zip(k1, k2))
            t[k] = t.get(k, 0) + v1 * v2
        return _Poly(t)

    def __add__(self, o):
        t = dict(self.t)
        for k, v in o.t.items():
            t[k] = t.get(k, 0) + v
        return _Poly(t)

    def scale(self, c):
        return _Poly({k: v * c for k, v in self.t.items()})

    def integrate(self, d):
        """int over the reference simplex: prod a_i! / (sum a_i + d)!  (Dirichlet integral)"""
        s = Fraction(0)
        for k, v in self.t.items():
            num = 1
            for a in k:
                num *= factorial(a)
            s += v * Fraction(num, factorial(sum(k) + d))
        return s


def _lam(d, i):
    e = [0] * (d + 1)
    e[i] = 1
    return _Poly({tuple(e): Fraction(1)})


def _one(d):
    return _Poly({(0,) * (d + 1): Fraction(1)})


def _lagrange_polys(d, k):
    out = []
    for a in lattice(d, k):
        p = _one(d)
        for j in range(d + 1):
            for m in range(a[j]):
                p = p * (_lam(d, j).scale(Fraction(k)) + _one(d).scale(Fraction(-m))).scale(Fraction(1, m + 1))
        out.append(p)
    return out


def _space_polys(d, space):
    lam = [_lam(d, i) for i in range(d + 1)]
    if space == "P1":
        return lam
    p2 = [l * (l.scale(2) + _one(d).scale(-1)) for l in lam]
    p2 += [(lam[a] * lam[b]).scale(4) for a, b in UFC_EDGES[d]]
    return p2


@lru_cache(maxsize=None)
def cross_mass(d, space, k):
    """R[I][J] = int_{ref} phi_I psi^k_J, exact, as float64 [nloc(space), nk]"""
    rows, cols = _space_polys(d, space), _lagrange_polys(d, k)
    return np.array([[float((r * c).integrate(d)) for c in cols] for r in rows])


# ---- quadrature ------------------------------------------------------------------------------
@lru_cache(maxsize=None)
def simplex_rule(d, degree):
    """collapsed (Duffy) Gauss-Legendre rule on the reference simplex, exact for polynomials of `degree`;
    weights sum to 1/d!"""
    n = (degree + d) // 2 + 1
    g, w = np.polynomial.legendre.leggauss(n)
    g, w = 0.5 * (g + 1.0), 0.5 * w
    if d == 2:
        a, b = np.meshgrid(g, g, indexing="ij")
        wa, wb = np.meshgrid(w, w, indexing="ij")
        X = np.stack([a.ravel(), (b * (1.0 - a)).ravel()], axis=1)
        W = (wa * wb * (1.0 - a)).ravel()
    else:
        a, b, c = np.meshgrid(g, g, g, indexing="ij")
        wa, wb, wc = np.meshgrid(w, w, w, indexing="ij")
        X = np.stack([a.ravel(), (b * (1.0 - a)).ravel(), (c * (1.0 - a) * (1.0 - b)).ravel()], axis=1)
        W = (wa * wb * wc * (1.0 - a) ** 2 * (1.0 - b)).ravel()
    return X, W


def fiat_default_rule_degree4(d):
    """the FIAT/FFC "default" scheme for quadrature degree 4: triangle 6 points
    (Strang-Fix / Dunavant), tetrahedron 11 points (Keast, one negative weight)"""
    if d == 2:
        a, b = 0.09157621350977, 0.44594849091597
        X = np.array([[a, a], [a, 1 - 2 * a], [1 - 2 * a, a], [b, b], [b, 1 - 2 * b], [1 - 2 * b, b]])
        W = np.array([0.10995174365532] * 3 + [0.22338158967801] * 3) / 2.0
        return X, W
    X = [[0.25, 0.25, 0.25]]
    W = [-74.0 / 5625.0]
    a, b = 1.0 / 14.0, 11.0 / 14.0
    for i in range(4):
        p = [a, a, a, a]
        p[i] = b
        X.append(p[1:])
        W.append(343.0 / 45000.0)
    a, b = 0.3994035761667992, 0.1005964238332008
    for i, j in itertools.combinations(range(4), 2):
        p = [b, b, b, b]
        p[i] = p[j] = a
        X.append(p[1:])
        W.append(56.0 / 2250.0)
    return np.array(X), np.array(W)


@lru_cache(maxsize=None)
def cfl_tables(d):
    """(B [nq][NL2], P [NL2][nq]) for nsb_cfl_max: P2 basis at the degree-4 default rule, and the DG-2 local L2
    projection as ONE matrix, nodal values = Phi_nodes (B^T W B)^-1 B^T W f_q  (dolfin LocalSolver with the same
    quadrature on both sides, source/ns_problem.py:571-580; Phi_nodes = identity for a nodal basis)"""
    X, W = fiat_default_rule_degree4(d)
    B = eval_p2(d, X)
    M = B.T @ (W[:, None] * B)
    P = np.linalg.solve(M, B.T * W[None, :])
    return np.ascontiguousarray(B), np.ascontiguousarray(P)
