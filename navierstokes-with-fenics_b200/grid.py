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
