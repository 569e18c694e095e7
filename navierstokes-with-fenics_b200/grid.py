"""Host meshes and facet markers: the data ``source/grid_generator.py:111-208`` obtains from
dolfin's ``RectangleMesh`` / ``BoxMesh`` and ``CompiledSubDomain.mark``,
rebuilt with numpy because dolfin itself is replaced below the solver classes.  Same function
names and marker enum as the reference module.  Setup only; everything per-step is on the GPU.
"""
from enum import Enum

import numpy as np

DOLFIN_EPS = 3.0e-16


class HyperCubeBoundaryMarkers(Enum):
    """``source/grid_generator.py:36-46``"""
    left = 1
    right = 2
    bottom = 3
    top = 4
    back = 5
    front = 6
    opening = 7


_LOCAL_EDGES = {2: ((1, 2), (0, 2), (0, 1)),
                3: ((2, 3), (1, 3), (1, 2), (0, 3), (0, 2), (0, 1))}


class SimplexMesh:
    """Ordered simplicial mesh (UFC): ``coords`` f64[nv,d], ``cells`` i32[nc,d+1] ascending,
    ``edges`` i32[ne,2] numbered lexicographically by (min,max) vertex, ``cell_edges`` i32
    [nc,3|6] in UFC local order, exterior facets ``bfacets`` i32[nbf,d] with owning cell and
    marker (the ``MeshFunction("size_t", mesh, d-1)`` of the reference, boundary part)."""

    def __init__(self, coords, cells):
        self.coords = np.ascontiguousarray(coords, dtype=np.float64)
        self.cells = np.ascontiguousarray(np.sort(np.asarray(cells), axis=1), dtype=np.int32)
        self.dim = int(self.coords.shape[1])
        nv = np.int64(self.coords.shape[0])
        c = self.cells.astype(np.int64)
        # edges
        le = np.array(_LOCAL_EDGES[self.dim])
        keys = c[:, le[:, 0]] * nv + c[:, le[:, 1]]
        uk, inv = np.unique(keys.ravel(), return_inverse=True)
        self._edge_keys = uk
        self.edges = np.stack([uk // nv, uk % nv], axis=1).astype(np.int32)
        self.cell_edges = inv.reshape(keys.shape).astype(np.int32)
        # exterior facets = facets seen by exactly one cell
        d = self.dim
        fk, fv, fc = [], [], []
        for i in range(d + 1):
            keep = [j for j in range(d + 1) if j != i]
            f = c[:, keep]
            k = f[:, 0]
            for j in range(1, d):
                k = k * nv + f[:, j]
            fk.append(k)
            fv.append(f)
            fc.append(np.arange(c.shape[0], dtype=np.int64))
        fk, fv, fc = np.concatenate(fk), np.concatenate(fv), np.concatenate(fc)
        order = np.argsort(fk, kind="stable")
        sk = fk[order]
        single = np.ones(len(sk), dtype=bool)
        same = sk[1:] == sk[:-1]
        single[1:] &= ~same
        single[:-1] &= ~same
        sel = order[single]
        self.bfacets = fv[sel].astype(np.int32)
        self.bfacet_cell = fc[sel].astype(np.int32)
        self.bfacet_marker = np.zeros(len(sel), dtype=np.int64)

    num_vertices = property(lambda self: self.coords.shape[0])
    num_cells = property(lambda self: self.cells.shape[0])
    num_edges = property(lambda self: self.edges.shape[0])

    def edge_index(self, a, b):
        a, b = np.asarray(a, dtype=np.int64), np.asarray(b, dtype=np.int64)
        k = np.minimum(a, b) * np.int64(self.num_vertices) + np.maximum(a, b)
        idx = np.searchsorted(self._edge_keys, k)
        if not np.array_equal(self._edge_keys[idx], k):
            raise ValueError("vertex pair is not an edge of the mesh")
        return idx

    def mark_near(self, axis, value, marker):
        """``CompiledSubDomain("near(x[axis], value) && on_boundary").mark(markers, marker)``"""
        x = self.coords[:, axis][self.bfacets]
        inside = np.all(np.abs(x - value) < DOLFIN_EPS, axis=1) & (np.abs(x.mean(axis=1) - value) < DOLFIN_EPS)
        self.bfacet_marker[inside] = int(marker)

    def mark_where(self, inside_fn, marker):
        """generic ``SubDomain.inside`` marking: all facet vertices and the midpoint inside."""
        X = self.coords[self.bfacets]
        ok = inside_fn(X.mean(axis=1))
        for j in range(X.shape[1]):
            ok &= inside_fn(X[:, j, :])
        self.bfacet_marker[ok] = int(marker)

    def _unmarked(self):
        """all facet markers back to 0 (the dolfin-named compat layer lets the caller's SubDomains mark them)"""
        self.bfacet_marker[:] = 0
        return self

    def boundary_ids(self):
        """``extract_all_boundary_markers`` (source/auxiliary_methods.py:56-67)"""
        if getattr(self, "global_boundary_ids", None) is not None:   # local mesh of a partition
            return set(self.global_boundary_ids)
        return set(int(m) for m in np.unique(self.bfacet_marker))


def _structured(p0, p1, n):
    dim = len(p0)
    axes = [float(p0[i]) + np.arange(n[i] + 1) * (float(p1[i]) - float(p0[i])) / n[i] for i in range(dim)]
    if dim == 2:
        nx, ny = n
        Y, X = np.meshgrid(axes[1], axes[0], indexing="ij")
        coords = np.stack([X.ravel(), Y.ravel()], axis=1)
        iy, ix = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
        v0 = (iy * (nx + 1) + ix).ravel()
        v1, v2 = v0 + 1, v0 + nx + 1
        v3 = v2 + 1
        cells = np.empty((2 * nx * ny, 3), dtype=np.int64)
        cells[0::2] = np.stack([v0, v1, v3], axis=1)          # "right" diagonal
        cells[1::2] = np.stack([v0, v2, v3], axis=1)
    else:
        nx, ny, nz = n
        Z, Y, X = np.meshgrid(axes[2], axes[1], axes[0], indexing="ij")
        coords = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
        iz, iy, ix = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
        v0 = ((iz * (ny + 1) + iy) * (nx + 1) + ix).ravel()
        v1, v2 = v0 + 1, v0 + nx + 1
        v3 = v2 + 1
        o = (nx + 1) * (ny + 1)
        v4, v5, v6, v7 = v0 + o, v1 + o, v2 + o, v3 + o
        cells = np.empty((6 * nx * ny * nz, 4), dtype=np.int64)
        for i, t in enumerate(((v0, v1, v3, v7), (v0, v1, v7, v5), (v0, v5, v7, v4),
                               (v0, v3, v2, v7), (v0, v6, v4, v7), (v0, v2, v6, v7))):
            cells[i::6] = np.stack(t, axis=1)
    mesh = SimplexMesh(coords, cells)
    M = HyperCubeBoundaryMarkers
    ids = (M.left, M.right, M.bottom, M.top, M.back, M.front)
    for ax in range(dim):
        mesh.mark_near(ax, float(p0[ax]), ids[2 * ax].value)
        mesh.mark_near(ax, float(p1[ax]), ids[2 * ax + 1].value)
    return mesh


def hyper_cube(dim, n_points=10):
    """``grid_generator.hyper_cube`` (source/grid_generator.py:111-151): returns
    (mesh, boundary markers); the marker function lives on the mesh object."""
    assert dim in (2, 3)
    mesh = _structured((0.0,) * dim, (1.0,) * dim, (int(n_points),) * dim)
    return mesh, mesh.bfacet_marker


def hyper_rectangle(first_point, second_point, n_points=10):
    """``grid_generator.hyper_rectangle`` (source/grid_generator.py:154-208)"""
    dim = len(first_point)
    assert dim in (2, 3) and len(second_point) == dim
    if isinstance(n_points, int):
        n_points = (n_points,) * dim
    assert all(second_point[i] > first_point[i] for i in range(dim))
    mesh = _structured(tuple(first_point), tuple(second_point), tuple(int(v) for v in n_points))
    return mesh, mesh.bfacet_marker


# DFG 2D-2 geometry and physical ids of gmsh-collection/meshes/DFGBenchmark.geo:7-12,105-109
DFG_BOUNDARY_IDS = {"inlet": 100, "outlet": 101, "lower wall": 102, "upper wall": 103, "cylinder": 104}


def channel_with_cylinder(m=16, n_radial=None, n_downstream=None, grading=1.6):
    """``grid_generator.channel_with_cylinder`` (source/grid_generator.py:452-455) on a SYNTHETIC
    block-structured triangulation: the reference reads a gmsh mesh of DFGBenchmark.geo, and gmsh
    is not available here.  Same geometry (channel 22 x 4.1, cylinder
    centre (2, 2), radius 0.5) and the same physical ids.  An O-grid of 4 m angular x n_radial
    cells joins the cylinder to the box [0, 4] x [0, 4.1] (m segments per box side, straight rays,
    radial cells graded towards the cylinder), a rectangular block of n_downstream x m cells covers
    [4, 22]; every quadrilateral is cut along the diagonal that keeps the triangles best shaped.
    Returns (mesh, boundary markers, boundary marker map) like the reference."""
    xc, yc, rc, h, length, xb = 2.0, 2.0, 0.5, 4.1, 22.0, 4.0
    n_radial = n_radial or max(2, (3 * m) // 4)
    n_downstream = n_downstream or max(2, int(round((length - xb) / (h / m) / 1.5)))
    # box boundary points, counter-clockwise from the corner (xb, 0): right, top, left, bottom side
    t = np.arange(m) / m
    right = np.stack([np.full(m, xb), h * t], axis=1)
    top = np.stack([xb - xb * t, np.full(m, h)], axis=1)
    left = np.stack([np.zeros(m), h - h * t], axis=1)
    bottom = np.stack([xb * t, np.zeros(m)], axis=1)
    Q = np.concatenate([right, top, left, bottom])                   # [4m, 2]
    ang = np.arctan2(Q[:, 1] - yc, Q[:, 0] - xc)
    Cpt = np.stack([xc + rc * np.cos(ang), yc + rc * np.sin(ang)], axis=1)
    s = (np.arange(n_radial + 1) / n_radial) ** grading               # 0 at the cylinder, 1 at the box
    ring = (1.0 - s)[:, None, None] * Cpt[None, :, :] + s[:, None, None] * Q[None, :, :]   # [nr+1, 4m, 2]
    nth = 4 * m
    pts = [ring.reshape(-1, 2)]
    quads = []
    for r in range(n_radial):
        for k in range(nth):
            k1 = (k + 1) % nth
            quads.append((r * nth + k, r * nth + k1, (r + 1) * nth + k1, (r + 1) * nth + k))
    # downstream block [xb, length] x [0, h]; its left column coincides with the right side of the box
    xs = xb + (length - xb) * np.arange(n_downstream + 1) / n_downstream
    ys = h * np.arange(m + 1) / m
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    off = ring.shape[0] * ring.shape[1]
    pts.append(np.stack([X.ravel(), Y.ravel()], axis=1))
    for i in range(n_downstream):
        for j in range(m):
            a = off + i * (m + 1) + j
            quads.append((a, a + (m + 1), a + (m + 1) + 1, a + 1))
    P = np.concatenate(pts)
    # merge coincident points (box side x = xb is shared by the two blocks)
    key = np.round(P / 1e-9).astype(np.int64)
    _, first, inv = np.unique(key, axis=0, return_index=True, return_inverse=True)
    inv = inv.reshape(-1)
    coords = P[first]
    quads = inv[np.asarray(quads, dtype=np.int64)]
    # shorter diagonal
    d02 = np.linalg.norm(coords[quads[:, 0]] - coords[quads[:, 2]], axis=1)
    d13 = np.linalg.norm(coords[quads[:, 1]] - coords[quads[:, 3]], axis=1)
    use02 = d02 <= d13
    tri_a = np.where(use02[:, None], quads[:, [0, 1, 2]], quads[:, [0, 1, 3]])
    tri_b = np.where(use02[:, None], quads[:, [0, 2, 3]], quads[:, [1, 2, 3]])
    mesh = SimplexMesh(coords, np.concatenate([tri_a, tri_b]))
    ids = DFG_BOUNDARY_IDS
    mesh.mark_near(0, 0.0, ids["inlet"])
    mesh.mark_near(0, length, ids["outlet"])
    mesh.mark_near(1, 0.0, ids["lower wall"])
    mesh.mark_near(1, h, ids["upper wall"])
    mesh.mark_where(lambda x: np.abs(np.hypot(x[:, 0] - xc, x[:, 1] - yc) - rc) < 0.05 * rc, ids["cylinder"])
    return mesh, mesh.bfacet_marker, dict(ids)
