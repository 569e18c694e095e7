"""Oracle Dirichlet boundary conditions: dolfin ``DirichletBC`` (method
"topological") semantics, SURVEY.md App. A.5 / A.6, as used by
``source/ns_solver_base.py:526-660`` and ``source/ns_ipcs_solver.py:45-64``.

A BC description here is a tuple ``(space, marker, comp, value)``:
  space  "V" | "Q" | "W"      which dof numbering the result is expressed in
  marker facet id
  comp   None (whole velocity vector), int c (velocity component c), "p" (pressure)
  value  callable g(x[npts,d]) -> [npts, d] (vector), [npts] (scalar), or a constant /
         tuple of constants.
Several BCs are merged in list order, later ones overriding earlier ones.

Test infrastructure only (see ``oracle/__init__.py``).
"""
import numpy as np


def _facet_nodes(spaces, marker):
    """Scalar P2 nodes (vertices + edges) and P1 vertices in the closure of all boundary
    facets carrying ``marker`` (in facet order)."""
    m = spaces.mesh
    sel = np.nonzero(m.bfacet_marker == marker)[0]
    fv = m.bfacets[sel].astype(np.int64)                      # [nf, d]
    verts = fv.ravel()
    d = m.dim
    pairs = [(i, j) for i in range(d) for j in range(i + 1, d)]
    if len(sel):
        e = np.concatenate([m.edge_index(fv[:, i], fv[:, j]) for i, j in pairs])
    else:
        e = np.zeros(0, dtype=np.int64)
    nodes = np.concatenate([verts, spaces.N0 + e])
    return nodes, verts


def _evaluate(value, x, ncomp):
    if callable(value):
        g = np.asarray(value(x), dtype=np.float64)
    else:
        g = np.broadcast_to(np.asarray(value, dtype=np.float64), (x.shape[0],) + ((ncomp,) if ncomp else ()))
    return g


def dirichlet_values(spaces, bcs):
    """Merge a list of BC tuples -> (dofs int64 ascending, values float64), last wins."""
    d = spaces.dim
    table = {}
    node_x = spaces.node_coords()
    for space, marker, comp, value in bcs:
        nodes, verts = _facet_nodes(spaces, marker)
        if comp == "p":
            x = spaces.mesh.coords[verts]
            g = _evaluate(value, x, 0)
            dofs = verts if space == "Q" else spaces.Q_to_W[verts]
            vals = g
        else:
            x = node_x[nodes]
            comps = range(d) if comp is None else [comp]
            g = _evaluate(value, x, d if comp is None else 0)
            dl, vl = [], []
            for c in comps:
                vd = d * nodes + c
                dl.append(vd if space == "V" else spaces.V_to_W[vd])
                vl.append(g[:, c] if comp is None else g)
            dofs, vals = np.concatenate(dl), np.concatenate(vl)
        for i, v in zip(dofs.tolist(), vals.tolist()):
            table[i] = v
    if not table:
        return np.zeros(0, dtype=np.int64), np.zeros(0)
    dofs = np.array(sorted(table), dtype=np.int64)
    return dofs, np.array([table[i] for i in dofs.tolist()])


def apply_ident(A, dofs):
    """``bc.apply(A)``: zero the BC rows (pattern kept), diagonal 1.0, columns untouched."""
    A = A.tocsr(copy=True)
    dofs = np.asarray(dofs, dtype=np.int64)
    if len(dofs) == 0:
        return A
    is_bc = np.zeros(A.shape[0], dtype=bool)
    is_bc[dofs] = True
    rows = np.repeat(np.arange(A.shape[0]), np.diff(A.indptr))
    in_bc_row = is_bc[rows]
    diag = in_bc_row & (A.indices == rows)
    assert diag.sum() == len(np.unique(dofs)), "a Dirichlet row has no diagonal entry in the pattern"
    A.data[in_bc_row] = 0.0
    A.data[diag] = 1.0
    return A


def apply_rhs_linear(b, dofs, vals):
    """``bc.apply(b)``: b[i] = g[i]."""
    b = b.copy()
    b[dofs] = vals
    return b


def apply_rhs_newton(b, x, dofs, vals):
    """``bc.apply(b, x)``: b[i] = x[i] - g[i]."""
    b = b.copy()
    b[dofs] = x[dofs] - vals
    return b
