"""Oracle meshes: what dolfin's RectangleMesh / BoxMesh + the marker code of
``source/grid_generator.py:111-208`` produce.

Test infrastructure only (see ``oracle/__init__.py``).
"""
import numpy as np

# UFC local edge -> local vertex pairs
UFC_EDGES = {
    2: np.array([[1, 2], [0, 2], [0, 1]]),
    3: np.array([[2, 3], [1, 3], [1, 2], [0, 3], [0, 2], [0, 1]]),
}
# UFC local facet i = the facet opposite local vertex i
DOLFIN_EPS = 3.0e-16

# ``HyperCubeBoundaryMarkers`` (source/grid_generator.py:36-46)
LEFT, RIGHT, BOTTOM, TOP, BACK, FRONT, OPENING = 1, 2, 3, 4, 5, 6, 7


class Mesh:
    """Ordered simplicial mesh with UFC entity numbering.

    coords [nv,d] float64; cells [nc,d+1] int32 (ascending vertex index per cell);
    edges [ne,2] (numbered by lexicographic order of (min v, max v), App. A.1);
    cell_edges [nc, 3|6] in UFC local edge order;
    bfacets [nbf,d] boundary facets (sorted vertices), bfacet_cell, bfacet_marker.
    """

    def __init__(self, coords, cells):
        self.coords = np.ascontiguousarray(coords, dtype=np.float64)
        cells = np.sort(np.asarray(cells, dtype=np.int64), axis=1)
        self.cells = cells.astype(np.int32)
        self.dim = self.coords.shape[1]
        assert self.cells.shape[1] == self.dim + 1
        self._build_edges()
        self._build_boundary_facets()

    @property
    def num_vertices(self):
        return self.coords.shape[0]

    @property
    def num_cells(self):
        return self.cells.shape[0]

    @property
    def num_edges(self):
        return self.edges.shape[0]

    def _build_edges(self):
        le = UFC_EDGES[self.dim]
        c = self.cells.astype(np.int64)
        a = c[:, le[:, 0]]
        b = c[:, le[:, 1]]          # a < b because cells are ordered and le rows ascend
        nv = self.num_vertices
        key = (a * nv + b).ravel()
        ukey, inv = np.unique(key, return_inverse=True)   # sorted == lexicographic (min,max)
        self.edges = np.stack([ukey // nv, ukey % nv], axis=1).astype(np.int32)
        self.cell_edges = inv.reshape(c.shape[0], le.shape[0]).astype(np.int32)
        self._edge_keys = ukey

    def edge_index(self, a, b):
        """Global edge number(s) of the vertex pair(s) (a,b)."""
        a = np.asarray(a, dtype=np.int64)
        b = np.asarray(b, dtype=np.int64)
        lo, hi = np.minimum(a, b), np.maximum(a, b)
        k = lo * self.num_vertices + hi
        idx = np.searchsorted(self._edge_keys, k)
        assert np.all(self._edge_keys[idx] == k)
        return idx

    def _build_boundary_facets(self):
        d = self.dim
        c = self.cells.astype(np.int64)
        nc = c.shape[0]
        facs, owner = [], []
        for i in range(d + 1):           # facet opposite local vertex i
            keep = [j for j in range(d + 1) if j != i]
            facs.append(c[:, keep])
            owner.append(np.arange(nc))
        facs = np.concatenate(facs)
        owner = np.concatenate(owner)
        nv = self.num_vertices
        key = facs[:, 0]
        for j in range(1, d):
            key = key * nv + facs[:, j]
        ukey, first, counts = np.unique(key, return_index=True, return_counts=True)
        sel = first[counts == 1]
        order = np.argsort(ukey[counts == 1], kind="stable")
        sel = sel[order]
        self.bfacets = facs[sel].astype(np.int32)
        self.bfacet_cell = owner[sel].astype(np.int32)
        self.bfacet_marker = np.zeros(len(sel), dtype=np.int64)

    def mark_plane(self, axis, value, marker):
        """``CompiledSubDomain("near(x[axis], value) && on_boundary").mark(...)``:
        a facet is marked when all its vertices and its midpoint are inside."""
        x = self.coords[self.bfacets][:, :, axis]            # [nbf, d]
        mid = x.mean(axis=1)
        inside = np.all(np.abs(x - value) < DOLFIN_EPS, axis=1) & (np.abs(mid - value) < DOLFIN_EPS)
        self.bfacet_marker[inside] = marker

    def mark_where(self, predicate, marker):
        """Generic marker: predicate(x[npts,d]) -> bool[npts] on vertices and midpoint."""
        X = self.coords[self.bfacets]
        inside = np.ones(len(X), dtype=bool)
        for j in range(X.shape[1]):
            inside &= predicate(X[:, j, :])
        inside &= predicate(X.mean(axis=1))
        self.bfacet_marker[inside] = marker

    def boundary_marker_ids(self):
        """``extract_all_boundary_markers`` (source/auxiliary_methods.py:56-67)."""
        return set(int(m) for m in np.unique(self.bfacet_marker))


def rectangle_mesh(p0, p1, nx, ny):
    """dolfin ``RectangleMesh(p0,p1,nx,ny)`` with the default "right" diagonal."""
    a, c = float(p0[0]), float(p0[1])
    b, d = float(p1[0]), float(p1[1])
    ix = np.arange(nx + 1)
    iy = np.arange(ny + 1)
    x = a + ix * (b - a) / nx
    y = c + iy * (d - c) / ny
    X, Y = np.meshgrid(x, y, indexing="xy")           # row-major in iy
    coords = np.stack([X.ravel(), Y.ravel()], axis=1)
    jx, jy = np.meshgrid(np.arange(nx), np.arange(ny), indexing="xy")
    v0 = (jy * (nx + 1) + jx).ravel()
    v1 = v0 + 1
    v2 = v0 + (nx + 1)
    v3 = v1 + (nx + 1)
    cells = np.empty((2 * nx * ny, 3), dtype=np.int64)
    cells[0::2] = np.stack([v0, v1, v3], axis=1)
    cells[1::2] = np.stack([v0, v2, v3], axis=1)
    return Mesh(coords, cells)


def box_mesh(p0, p1, nx, ny, nz):
    """dolfin ``BoxMesh`` (six tetrahedra per cube around the v0-v7 diagonal)."""
    lo = [float(v) for v in p0]
    hi = [float(v) for v in p1]
    x = lo[0] + np.arange(nx + 1) * (hi[0] - lo[0]) / nx
    y = lo[1] + np.arange(ny + 1) * (hi[1] - lo[1]) / ny
    z = lo[2] + np.arange(nz + 1) * (hi[2] - lo[2]) / nz
    Z, Y, X = np.meshgrid(z, y, x, indexing="ij")
    coords = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    kz, ky, kx = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    v0 = (kz * (ny + 1) * (nx + 1) + ky * (nx + 1) + kx).ravel()
    v1 = v0 + 1
    v2 = v0 + (nx + 1)
    v3 = v1 + (nx + 1)
    off = (nx + 1) * (ny + 1)
    v4, v5, v6, v7 = v0 + off, v1 + off, v2 + off, v3 + off
    tets = [(v0, v1, v3, v7), (v0, v1, v7, v5), (v0, v5, v7, v4),
            (v0, v3, v2, v7), (v0, v6, v4, v7), (v0, v2, v6, v7)]
    cells = np.empty((6 * nx * ny * nz, 4), dtype=np.int64)
    for i, t in enumerate(tets):
        cells[i::6] = np.stack(t, axis=1)
    return Mesh(coords, cells)


def _mark_box(mesh, lo, hi):
    mesh.mark_plane(0, lo[0], LEFT)
    mesh.mark_plane(0, hi[0], RIGHT)
    mesh.mark_plane(1, lo[1], BOTTOM)
    mesh.mark_plane(1, hi[1], TOP)
    if mesh.dim == 3:
        mesh.mark_plane(2, lo[2], BACK)
        mesh.mark_plane(2, hi[2], FRONT)
    return mesh


def hyper_cube(dim, n_points=10):
    """``grid_generator.hyper_cube`` (source/grid_generator.py:111-151)."""
    if dim == 2:
        m = rectangle_mesh((0., 0.), (1., 1.), n_points, n_points)
    else:
        m = box_mesh((0., 0., 0.), (1., 1., 1.), n_points, n_points, n_points)
    return _mark_box(m, (0.,) * dim, (1.,) * dim)


def hyper_rectangle(first_point, second_point, n_points=10):
    """``grid_generator.hyper_rectangle`` (source/grid_generator.py:154-208)."""
    dim = len(first_point)
    if not isinstance(n_points, (tuple, list)):
        n_points = (n_points,) * dim
    if dim == 2:
        m = rectangle_mesh(first_point, second_point, *n_points)
    else:
        m = box_mesh(first_point, second_point, *n_points)
    return _mark_box(m, first_point, second_point)
