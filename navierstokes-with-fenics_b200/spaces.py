"""Taylor-Hood function spaces on the host: dof numbering, cell->dof maps, sub<->mixed maps and
Dirichlet dof/value tables.

Replaces what ``SolverBase._setup_function_spaces`` (source/ns_solver_base.py:501-524),
``_get_subspace(s)`` (:401-422), the ``FunctionAssigner`` factories (:424-476) and
``dolfin.DirichletBC`` (:546-660, source/ns_ipcs_solver.py:45-64) obtain from dolfin.
Numbering contract "ufc" (= dolfin with reorder_dofs_serial off):
  scalar P2 node: vertex v -> v, edge e -> N0 + e
  V = [P2]^d collapsed, block size d:   dof = d*node + comp
  Q = P1 collapsed:                     dof = vertex
  W = [P2]^d x P1 mixed, block size 1:  comp c node n -> c*(N0+N1) + n ; pressure -> d*(N0+N1) + v
  cell-local W order [u_0 | u_1 | (u_2) | p], P2 = vertices then UFC edges.
"""
import numpy as np

from .grid import SimplexMesh


class TaylorHoodSpaces:
    def __init__(self, mesh, node_perm=None):
        """``node_perm`` (partitioned meshes only, partition.py): natural P2 node number
        (vertex v / N0 + e) -> the owned-first local number used on the device."""
        assert isinstance(mesh, SimplexMesh)
        self.mesh = mesh
        self.node_perm = None if node_perm is None else np.ascontiguousarray(node_perm, dtype=np.int32)
        d = self.dim = mesh.dim
        N0, N1 = mesh.num_vertices, mesh.num_edges
        self.n_vertices, self.n_nodes = N0, N0 + N1
        self.nloc2 = (d + 1) + mesh.cell_edges.shape[1]
        self.nloc1 = d + 1
        self.cell_nodes = np.ascontiguousarray(
            np.concatenate([mesh.cells, mesh.cell_edges + np.int32(N0)], axis=1), dtype=np.int32)
        if self.node_perm is not None:
            assert self.node_perm.shape == (N0 + N1,)
            self.cell_nodes = np.ascontiguousarray(self.node_perm[self.cell_nodes])
        self.ndofs_V = d * self.n_nodes
        self.ndofs_Q = N0
        self.ndofs_W = d * self.n_nodes + N0
        if self.ndofs_W >= 2 ** 31:
            raise ValueError("mixed space exceeds int32 dof indices")
        self._cell_dofs_W = None

    @property
    def cell_dofs_W(self):
        if self._cell_dofs_W is None:
            d, nn = self.dim, np.int32(self.n_nodes)
            parts = [self.cell_nodes + np.int32(c) * nn for c in range(d)]
            parts.append(self.mesh.cells + np.int32(d) * nn)
            self._cell_dofs_W = np.ascontiguousarray(np.concatenate(parts, axis=1), dtype=np.int32)
        return self._cell_dofs_W

    @property
    def cell_dofs_V(self):
        d = self.dim
        return np.concatenate([d * self.cell_nodes + c for c in range(d)], axis=1).astype(np.int32)

    # FunctionAssigner index maps: W dof of every V / Q dof
    def V_to_W(self):
        d, nn = self.dim, self.n_nodes
        node = np.arange(nn, dtype=np.int64)
        out = np.empty(d * nn, dtype=np.int32)
        for c in range(d):
            out[d * node + c] = c * nn + node
        return out

    def Q_to_W(self):
        return (self.dim * self.n_nodes + np.arange(self.n_vertices, dtype=np.int64)).astype(np.int32)

    def node_coordinates(self):
        m = self.mesh
        mid = 0.5 * (m.coords[m.edges[:, 0]] + m.coords[m.edges[:, 1]])
        x = np.concatenate([m.coords, mid], axis=0)
        if self.node_perm is not None:
            out = np.empty_like(x)
            out[self.node_perm] = x
            return out
        return x

    # ---- Dirichlet tables ---------------------------------------------------------------------
    def _facet_closure(self, marker):
        """P2 nodes and P1 vertices in the closure of the facets carrying ``marker``
        (DirichletBC method "topological")."""
        m = self.mesh
        fv = m.bfacets[m.bfacet_marker == int(marker)].astype(np.int64)
        if fv.shape[0] == 0:
            e = np.zeros(0, dtype=np.int64)
            return e, e
        d = m.dim
        pairs = [(i, j) for i in range(d) for j in range(i + 1, d)]
        edges = np.concatenate([m.edge_index(fv[:, i], fv[:, j]) for i, j in pairs])
        verts = np.unique(fv.ravel())
        nodes = np.concatenate([verts, self.n_vertices + np.unique(edges)])
        if self.node_perm is not None:
            nodes = self.node_perm[nodes].astype(np.int64)
        return nodes, verts

    def dirichlet_table(self, bcs, space):
        """bcs: list of (marker, comp, value) with comp None (whole velocity vector), int
        (one velocity component) or "p" (pressure); value a constant / tuple / callable
        x[n,d] -> values.  space "V" | "Q" | "W" selects the numbering of the result.
        Returns (dofs int32 ascending, values f64); later entries override earlier ones."""
        d = self.dim
        n = {"V": self.ndofs_V, "Q": self.ndofs_Q, "W": self.ndofs_W}[space]
        vals = np.zeros(n)
        mask = np.zeros(n, dtype=bool)
        xnodes = None
        for marker, comp, value in bcs:
            nodes, verts = self._facet_closure(marker)
            if len(nodes) == 0:          # this rank's part of the mesh does not touch the boundary
                continue
            if comp == "p":
                if space == "V":
                    continue
                x = self.mesh.coords[verts]
                g = np.asarray(value(x), dtype=np.float64).reshape(-1) if callable(value) \
                    else np.full(len(verts), float(value))
                dofs = verts if space == "Q" else d * self.n_nodes + verts
                vals[dofs] = g
                mask[dofs] = True
                continue
            if space == "Q":
                continue
            if xnodes is None:
                xnodes = self.node_coordinates()
            x = xnodes[nodes]
            comps = list(range(d)) if comp is None else [int(comp)]
            if callable(value):
                g = np.asarray(value(x), dtype=np.float64)
                g = g.reshape(len(nodes), -1) if comp is None else g.reshape(len(nodes), 1)
            else:
                g = np.broadcast_to(np.asarray(value, dtype=np.float64).reshape(1, -1), (len(nodes), len(comps)))
            for k, c in enumerate(comps):
                dofs = d * nodes + c if space == "V" else c * self.n_nodes + nodes
                vals[dofs] = g[:, k]
                mask[dofs] = True
        dofs = np.nonzero(mask)[0].astype(np.int32)
        return dofs, vals[dofs]
