"""Oracle dof maps and sparsity patterns.

Numbering contract ``"ufc"`` = dolfin 2019.1.0 with
``parameters["reorder_dofs_serial"] = False``:

* scalar P2 node numbering: vertex v -> v, edge e -> N0 + e
* collapsed velocity space V = [P2]^d (block size d): dof = d*node + comp
* collapsed pressure space Q = P1: dof = vertex
* mixed W = [P2]^d x P1 (block size 1): comp c vertex -> c(N0+N1)+v,
  comp c edge -> c(N0+N1)+N0+e, pressure -> d(N0+N1)+v;
  cell-local order [u_0: v0..vd,e0..| u_1 ... | p: v0..vd]

Test infrastructure only (see ``oracle/__init__.py``).
"""
import numpy as np


class Spaces:
    def __init__(self, mesh):
        self.mesh = mesh
        d = mesh.dim
        N0, N1 = mesh.num_vertices, mesh.num_edges
        self.dim = d
        self.N0, self.N1 = N0, N1
        self.n_nodes = N0 + N1
        # scalar P2 nodes per cell [nc, nloc2]
        self.cell_nodes = np.concatenate(
            [mesh.cells.astype(np.int64), N0 + mesh.cell_edges.astype(np.int64)], axis=1)
        self.nloc2 = self.cell_nodes.shape[1]
        self.nloc1 = d + 1
        # V: cell-local order [comp0 nodes | comp1 nodes | ...]
        self.cell_dofs_V = np.concatenate(
            [d * self.cell_nodes + c for c in range(d)], axis=1)
        self.ndofs_V = d * self.n_nodes
        self.cell_dofs_Q = mesh.cells.astype(np.int64)
        self.ndofs_Q = N0
        self.cell_dofs_W = np.concatenate(
            [c * self.n_nodes + self.cell_nodes for c in range(d)]
            + [d * self.n_nodes + mesh.cells.astype(np.int64)], axis=1)
        self.ndofs_W = d * self.n_nodes + N0
        # FunctionAssigner maps (source/ns_solver_base.py:424-476): W dof of each V / Q dof
        node = np.arange(self.n_nodes)
        self.V_to_W = np.empty(self.ndofs_V, dtype=np.int64)
        for c in range(d):
            self.V_to_W[d * node + c] = c * self.n_nodes + node
        self.Q_to_W = d * self.n_nodes + np.arange(N0, dtype=np.int64)

    # dof coordinates ---------------------------------------------------------------
    def node_coords(self):
        m = self.mesh
        mids = 0.5 * (m.coords[m.edges[:, 0]] + m.coords[m.edges[:, 1]])
        return np.concatenate([m.coords, mids])

    def dof_coords_V(self):
        return np.repeat(self.node_coords(), self.dim, axis=0)

    def dof_coords_Q(self):
        return self.mesh.coords.copy()


def sparsity_pattern(cell_rows, cell_cols, nrows, ncols):
    """Union over cells of rows x cols of the whole element; CSR with ascending
    int32 column indices (PETSc AIJ), no extra diagonal.  Returns (rowptr int64, col int32)."""
    nr, ncl = cell_rows.shape[1], cell_cols.shape[1]
    chunks = []
    step = max(1, 2_000_000 // (nr * ncl))
    for s in range(0, cell_rows.shape[0], step):
        r = cell_rows[s:s + step].astype(np.int64)
        c = cell_cols[s:s + step].astype(np.int64)
        key = (r[:, :, None] * ncols + c[:, None, :]).ravel()
        chunks.append(np.unique(key))
    keys = np.unique(np.concatenate(chunks))
    rows = keys // ncols
    col = (keys % ncols).astype(np.int32)
    rowptr = np.zeros(nrows + 1, dtype=np.int64)
    np.add.at(rowptr, rows + 1, 1)
    rowptr = np.cumsum(rowptr)
    return rowptr, col


def entry_positions(rowptr, col, ncols, cell_rows, cell_cols):
    """cell -> nnz map: pos[c,i,j] = index in the CSR value array of (rows[c,i], cols[c,j])."""
    nrows = len(rowptr) - 1
    rows_of_nnz = np.repeat(np.arange(nrows, dtype=np.int64), np.diff(rowptr))
    keys = rows_of_nnz * ncols + col.astype(np.int64)          # ascending
    want = cell_rows.astype(np.int64)[:, :, None] * ncols + cell_cols.astype(np.int64)[:, None, :]
    pos = np.searchsorted(keys, want.ravel())
    assert np.all(keys[pos] == want.ravel())
    return pos.reshape(want.shape)
