"""Oracle Krylov solvers (numpy) -- TEST INFRASTRUCTURE.

Algorithm-matched CPU counterpart of the CUDA Krylov path (Jacobi-preconditioned CG for the SPD
pressure-Poisson / mass systems, right-Jacobi-preconditioned restarted GMRES for the velocity
systems), used (a) as the timed CPU baseline at sizes where the reference's sparse direct LU
 is infeasible and (b) to cross-check the GPU Krylov results.  Stopping rule as in
include/nsb200.h: ||r|| <= max(rtol ||b||, atol).
"""
import ctypes as C

import numpy as np


class Operator:
    """CSR matvec through the OpenMP ``nso_spmv`` of oracle/c/nsoracle.c."""

    def __init__(self, A):
        from . import native
        self.lib = native.load()
        A = A.tocsr()
        A.sort_indices()
        self.n = A.shape[0]
        self.rowptr = np.ascontiguousarray(A.indptr, dtype=np.int64)
        self.col = np.ascontiguousarray(A.indices, dtype=np.int32)
        self.vals = np.ascontiguousarray(A.data, dtype=np.float64)
        self.diag = A.diagonal()

    def __call__(self, x, out=None):
        x = np.ascontiguousarray(x, dtype=np.float64)
        y = np.empty(self.n) if out is None else out
        self.lib.nso_spmv(C.c_int64(self.n), self.rowptr.ctypes.data, self.col.ctypes.data, self.vals.ctypes.data,
                          x.ctypes.data, y.ctypes.data)
        return y


def pcg(A, b, x0=None, rtol=1e-12, atol=0.0, maxit=10000, nullspace=False):
    op = A if isinstance(A, Operator) else Operator(A)
    dinv = 1.0 / op.diag
    b = np.array(b, dtype=np.float64)
    if nullspace:
        b -= b.mean()
    x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=np.float64)
    r = b - op(x)
    z = dinv * r
    p = z.copy()
    rz = r @ z
    tol = max(rtol * np.linalg.norm(b), atol)
    it = 0
    q = np.empty_like(b)
    while np.linalg.norm(r) > tol and it < maxit:
        op(p, out=q)
        alpha = rz / (p @ q)
        x += alpha * p
        r -= alpha * q
        z = dinv * r
        rz_new = r @ z
        p = z + (rz_new / rz) * p
        rz = rz_new
        it += 1
    if nullspace:
        x -= x.mean()
    return x, it


def gmres(A, b, x0=None, rtol=1e-12, atol=0.0, maxit=2000, restart=40):
    op = A if isinstance(A, Operator) else Operator(A)
    dinv = 1.0 / op.diag
    n = len(b)
    x = np.zeros(n) if x0 is None else np.array(x0, dtype=np.float64)
    tol = max(rtol * np.linalg.norm(b), atol)
    it = 0
    m = restart
    while True:
        r = b - op(x)
        beta = np.linalg.norm(r)
        if beta <= tol or it >= maxit:
            break
        V = np.empty((m + 1, n))
        H = np.zeros((m + 1, m))
        cs, sn, g = np.zeros(m), np.zeros(m), np.zeros(m + 1)
        V[0] = r / beta
        g[0] = beta
        j = 0
        while j < m and it < maxit:
            w = op(dinv * V[j])
            for _ in range(2):                      # CGS-2 like the CUDA path
                h = V[:j + 1] @ w
                w -= h @ V[:j + 1]
                H[:j + 1, j] += h
            H[j + 1, j] = np.linalg.norm(w)
            V[j + 1] = w / H[j + 1, j]
            for k in range(j):
                t = cs[k] * H[k, j] + sn[k] * H[k + 1, j]
                H[k + 1, j] = -sn[k] * H[k, j] + cs[k] * H[k + 1, j]
                H[k, j] = t
            dd = np.hypot(H[j, j], H[j + 1, j])
            cs[j], sn[j] = H[j, j] / dd, H[j + 1, j] / dd
            H[j, j], H[j + 1, j] = dd, 0.0
            g[j + 1] = -sn[j] * g[j]
            g[j] = cs[j] * g[j]
            it += 1
            j += 1
            if abs(g[j]) <= tol:
                break
        y = np.linalg.solve(np.triu(H[:j, :j]), g[:j])
        x += dinv * (y @ V[:j])
    return x, it


# ---- the same two solvers in C/OpenMP (oracle/c/nsoracle.c nso_pcg / nso_gmres): the timed CPU arm ----
def _csr_args(A):
    op = A if isinstance(A, Operator) else Operator(A)
    return op, (C.c_int64(op.n), op.rowptr.ctypes.data, op.col.ctypes.data, op.vals.ctypes.data)


def pcg_c(A, b, x0=None, rtol=1e-12, atol=0.0, maxit=10000, nullspace=False):
    op, args = _csr_args(A)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=np.float64)
    res = C.c_double()
    it = op.lib.nso_pcg(*args, b.ctypes.data, x.ctypes.data, rtol, atol, maxit, 1 if nullspace else 0, C.byref(res))
    if it < 0:
        raise RuntimeError("nso_pcg did not converge (%d iterations, residual %g)" % (-it - 1, res.value))
    return x, it


def gmres_c(A, b, x0=None, rtol=1e-12, atol=0.0, maxit=2000, restart=40):
    op, args = _csr_args(A)
    b = np.ascontiguousarray(b, dtype=np.float64)
    x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=np.float64)
    res = C.c_double()
    it = op.lib.nso_gmres(*args, b.ctypes.data, x.ctypes.data, rtol, atol, maxit, restart, C.byref(res))
    if it < 0:
        raise RuntimeError("nso_gmres did not converge (%d iterations, residual %g)" % (-it - 1, res.value))
    return x, it
