"""Shared helpers of the parity tests: seeded fields, small meshes, comparison metrics."""
import numpy as np
import scipy.sparse as sp


def meshes():
    """(name, product mesh, oracle mesh) for a 2D and a 3D anisotropic structured mesh."""
    import nsb200
    import oracle
    out = []
    for p0, p1, n in (((0.0, 0.0), (2.0, 1.0), (6, 4)), ((0.0, 0.0, 0.0), (1.5, 1.0, 0.5), (3, 2, 2))):
        m, _ = nsb200.hyper_rectangle(p0, p1, n)
        om = oracle.mesh.hyper_rectangle(p0, p1, n)
        out.append(("%dd" % len(p0), m, om))
    return out


def random_fields(spaces, seed=1234):
    rng = np.random.default_rng(seed)
    x = spaces.node_coordinates() if hasattr(spaces, "node_coordinates") else spaces.node_coords()
    d = x.shape[1]
    nn = x.shape[0]
    nv = spaces.n_vertices if hasattr(spaces, "n_vertices") else spaces.N0

    def vel(s):
        u = np.stack([np.sin(2.0 * x[:, (c + 1) % d] + s + c) + 0.3 * x[:, c] for c in range(d)], axis=1)
        return (u + rng.uniform(-1e-1, 1e-1, size=(nn, d))).reshape(-1)
    p = np.cos(3.0 * x[:nv, 0]) + rng.uniform(-1e-1, 1e-1, size=nv)
    return vel(0.0), vel(0.4), vel(0.9), p


def rel_max_err(a, b):
    a, b = np.asarray(a), np.asarray(b)
    scale = np.abs(b).max()
    return float(np.abs(a - b).max() / (scale if scale > 0 else 1.0))


def bsr_to_csr(rowptr, col, vals, bs):
    nb = len(rowptr) - 1
    return sp.bsr_matrix((vals.reshape(-1, bs, bs), col, rowptr), shape=(nb * bs, nb * bs)).tocsr()


def kron_csr(rowptr, col, vals, bs):
    n = len(rowptr) - 1
    A = sp.csr_matrix((vals, col, rowptr), shape=(n, n))
    return sp.kron(A, sp.identity(bs), format="csr")
