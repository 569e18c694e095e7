"""ctypes front end of the plain-C oracle (``oracle/c/nsoracle.c``) -- TEST INFRASTRUCTURE.

``CAssembler`` has the interface of ``oracle.forms.Assembler`` (same method names and
arguments) so that the oracle solvers can run on either.  The C code does what one dolfin
``assemble()`` does for the reference (cell loop, quadrature element tensors, PETSc-style CSR
insertion); it is checked against the numpy oracle in
``tests/test_oracle.py`` and is the timed CPU baseline of ``bench.py``.

Only ``tests/``, ``__graft_entry__`` and ``bench.py`` (cpu_baseline / --impl reference) may
import this module.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import scipy.sparse as sp

from . import fem
from .dofmap import sparsity_pattern

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "c", "nsoracle.c")
LIB = os.path.join(HERE, "libnsoracle.so")

VEC, MAT, PICARD, BDF, MIXED = 1, 2, 4, 8, 32
QUAD_DEGREE = 6


def build(force=False):
    """gcc -O3 -fopenmp -shared oracle/c/nsoracle.c -> oracle/libnsoracle.so (no -march: the
    library travels to the GPU box)."""
    if not force and os.path.exists(LIB) and os.path.getmtime(LIB) >= os.path.getmtime(SRC):
        return LIB
    cmd = ["gcc", "-O3", "-fopenmp", "-fPIC", "-shared", "-Wall", "-o", LIB, SRC, "-lm"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed:\n%s\n%s" % (" ".join(cmd), res.stderr[-4000:]))
    return LIB


class _Mesh(C.Structure):
    _fields_ = [("d", C.c_int32), ("nc", C.c_int32), ("nl2", C.c_int32), ("nq", C.c_int32),
                ("coords", C.c_void_p), ("cells", C.c_void_p), ("qw", C.c_void_p), ("phi2", C.c_void_p),
                ("dphi2", C.c_void_p), ("phi1", C.c_void_p), ("dphi1", C.c_void_p)]


class _Coef(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("inv_k", "a0", "a1", "a2", "c_conv", "c_p", "c_visc")]


_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        lib = C.CDLL(LIB)
        vp = C.c_void_p
        lib.nso_max_threads.restype = C.c_int
        lib.nso_set_threads.argtypes = [C.c_int]
        lib.nso_momentum.restype = C.c_int
        lib.nso_momentum.argtypes = [C.POINTER(_Mesh), vp, vp, C.c_int, C.POINTER(_Coef), vp, vp, vp, vp, vp, vp, vp, vp]
        lib.nso_scalar_matrix.restype = C.c_int
        lib.nso_scalar_matrix.argtypes = [C.POINTER(_Mesh), C.c_int, vp, vp, vp, vp]
        lib.nso_poisson_rhs.restype = C.c_int
        lib.nso_poisson_rhs.argtypes = [C.POINTER(_Mesh), vp, vp, vp, vp, C.c_double, vp]
        lib.nso_correction_rhs.restype = C.c_int
        lib.nso_correction_rhs.argtypes = [C.POINTER(_Mesh), vp, vp, vp, vp, vp, C.c_double, vp]
        lib.nso_spmv.restype = None
        lib.nso_spmv.argtypes = [C.c_int64, vp, vp, vp, vp, vp]
        lib.nso_pcg.restype = C.c_int
        lib.nso_pcg.argtypes = [C.c_int64, vp, vp, vp, vp, vp, C.c_double, C.c_double, C.c_int, C.c_int, vp]
        lib.nso_gmres.restype = C.c_int
        lib.nso_gmres.argtypes = [C.c_int64, vp, vp, vp, vp, vp, C.c_double, C.c_double, C.c_int, C.c_int, vp]
        _lib = lib
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class CAssembler:
    """Same interface as ``oracle.forms.Assembler``; arithmetic in ``nsoracle.c``."""

    def __init__(self, spaces, threads=None):
        self.lib = load()
        if threads:
            self.lib.nso_set_threads(int(threads))
        self.threads = int(threads) if threads else self.lib.nso_max_threads()
        self.sp = s = spaces
        m = s.mesh
        self.d = d = m.dim
        qp, qw = fem.simplex_quadrature(d, QUAD_DEGREE)
        phi2, dphi2 = fem.tabulate_p2(qp)
        phi1, dphi1 = fem.tabulate_p1(qp)
        self._keep = dict(coords=_f64(m.coords), cells=np.ascontiguousarray(m.cells, dtype=np.int32), qw=_f64(qw),
                          phi2=_f64(phi2), dphi2=_f64(dphi2), phi1=_f64(phi1), dphi1=_f64(dphi1[0]))
        k = self._keep
        self.mesh = _Mesh(d, m.num_cells, s.nloc2, len(qw), _p(k["coords"]), _p(k["cells"]), _p(k["qw"]),
                          _p(k["phi2"]), _p(k["dphi2"]), _p(k["phi1"]), _p(k["dphi1"]))
        i32 = lambda a: np.ascontiguousarray(a, dtype=np.int32)   # noqa: E731
        self.vdofs_V, self.pdofs_Q = i32(s.cell_dofs_V), i32(s.cell_dofs_Q)
        nv = d * s.nloc2
        self.vdofs_W, self.pdofs_W = i32(s.cell_dofs_W[:, :nv]), i32(s.cell_dofs_W[:, nv:])
        self._patterns = {}
        self.detJ = np.abs(np.linalg.det(np.transpose(m.coords[m.cells][:, 1:, :] - m.coords[m.cells][:, :1, :],
                                                      (0, 2, 1))))

    def pattern(self, name):
        """name in {"VV","QQ","WW"} -> (rowptr int64, col int32)."""
        if name not in self._patterns:
            s = self.sp
            cd, n = {"VV": (s.cell_dofs_V, s.ndofs_V), "QQ": (s.cell_dofs_Q, s.ndofs_Q),
                     "WW": (s.cell_dofs_W, s.ndofs_W)}[name]
            rowptr, col = sparsity_pattern(cd, cd, n, n)
            self._patterns[name] = (np.ascontiguousarray(rowptr, dtype=np.int64),
                                    np.ascontiguousarray(col, dtype=np.int32))
        return self._patterns[name]

    def _csr(self, name, vals):
        rowptr, col = self.pattern(name)
        n = len(rowptr) - 1
        return sp.csr_matrix((vals, col, rowptr), shape=(n, n))

    def _coef(self, alpha, k, c_conv, c_p, c_visc):
        return _Coef(1.0 / k if k else 0.0, alpha[0], alpha[1], alpha[2], c_conv, c_p, c_visc)

    # ---- IPCS forms -----------------------------------------------------------------------
    def ipcs_momentum_residual(self, ustar, un, unm1, pn, alpha, k, c_conv, c_p, c_visc):
        b = np.zeros(self.sp.ndofs_V)
        cf = self._coef(alpha, k, c_conv, c_p, c_visc)
        ustar, un, unm1, pn = _f64(ustar), _f64(un), _f64(unm1), _f64(pn)
        st = self.lib.nso_momentum(C.byref(self.mesh), _p(self.vdofs_V), _p(self.pdofs_Q), VEC | BDF, C.byref(cf),
                                   _p(ustar), _p(un), _p(unm1), _p(pn), _p(b), None, None, None)
        assert st == 0
        return b

    def ipcs_momentum_jacobian(self, ustar, alpha0, k, c_conv, c_visc, newton=True):
        rowptr, col = self.pattern("VV")
        vals = np.zeros(len(col))
        cf = self._coef((alpha0, 0.0, 0.0), k, c_conv, 0.0, c_visc)
        ustar = _f64(ustar)
        st = self.lib.nso_momentum(C.byref(self.mesh), _p(self.vdofs_V), _p(self.pdofs_Q),
                                   MAT | BDF | (0 if newton else PICARD), C.byref(cf), _p(ustar), _p(ustar),
                                   _p(ustar), None, None, _p(rowptr), _p(col), _p(vals))
        assert st == 0
        return self._csr("VV", vals)

    def _scalar(self, which, name, dofs):
        rowptr, col = self.pattern(name)
        vals = np.zeros(len(col))
        st = self.lib.nso_scalar_matrix(C.byref(self.mesh), which, _p(dofs), _p(rowptr), _p(col), _p(vals))
        assert st == 0
        return self._csr(name, vals)

    def poisson_matrix(self):
        return self._scalar(0, "QQ", self.pdofs_Q)

    def pressure_mass_matrix(self):
        return self._scalar(1, "QQ", self.pdofs_Q)

    def velocity_mass_matrix(self):
        return self._scalar(2, "VV", self.vdofs_V)

    def poisson_rhs(self, pn, ustar, alpha0, k):
        b = np.zeros(self.sp.ndofs_Q)
        pn, ustar = _f64(pn), _f64(ustar)
        self.lib.nso_poisson_rhs(C.byref(self.mesh), _p(self.vdofs_V), _p(self.pdofs_Q), _p(pn), _p(ustar),
                                 alpha0 / k, _p(b))
        return b

    def correction_rhs(self, ustar, pnew, pold, alpha0, k):
        b = np.zeros(self.sp.ndofs_V)
        ustar, pnew, pold = _f64(ustar), _f64(pnew), _f64(pold)
        self.lib.nso_correction_rhs(C.byref(self.mesh), _p(self.vdofs_V), _p(self.pdofs_Q), _p(ustar), _p(pnew),
                                    _p(pold), k / alpha0, _p(b))
        return b

    # ---- mixed forms ------------------------------------------------------------------------
    def mixed_residual(self, w0, c_conv, c_p, c_visc, bdf=None):
        b = np.zeros(self.sp.ndofs_W)
        w0 = _f64(w0)
        if bdf is not None:
            alpha, k, w1, w2 = bdf
            w1, w2 = _f64(w1), _f64(w2)
            cf, fl = self._coef(alpha, k, c_conv, c_p, c_visc), VEC | BDF | MIXED
        else:
            w1 = w2 = w0
            cf, fl = self._coef((0.0, 0.0, 0.0), 0.0, c_conv, c_p, c_visc), VEC | MIXED
        st = self.lib.nso_momentum(C.byref(self.mesh), _p(self.vdofs_W), _p(self.pdofs_W), fl, C.byref(cf),
                                   _p(w0), _p(w1), _p(w2), _p(w0), _p(b), None, None, None)
        assert st == 0
        return b

    def mixed_jacobian(self, w0, c_conv, c_p, c_visc, mass_coef=0.0, newton=True):
        rowptr, col = self.pattern("WW")
        vals = np.zeros(len(col))
        w0 = _f64(w0)
        cf = _Coef(1.0, mass_coef, 0.0, 0.0, c_conv, c_p, c_visc)     # a0 * inv_k = mass_coef
        fl = MAT | MIXED | (BDF if mass_coef != 0.0 else 0) | (0 if newton else PICARD)
        st = self.lib.nso_momentum(C.byref(self.mesh), _p(self.vdofs_W), _p(self.pdofs_W), fl, C.byref(cf),
                                   _p(w0), _p(w0), _p(w0), _p(w0), None, _p(rowptr), _p(col), _p(vals))
        assert st == 0
        return self._csr("WW", vals)

    # ---- functionals -------------------------------------------------------------------------
    def volume(self):
        return float(self.detJ.sum() * self._keep["qw"].sum())

    def integral_Q(self, p):
        w = np.asarray(self.pressure_mass_matrix().sum(axis=0)).ravel()
        return float(w @ p)
