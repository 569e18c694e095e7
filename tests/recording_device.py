"""A fake Device for CPU tests of the host logic above the C-ABI: same Python interface as nsb200.device.Device, torch
CPU tensors as "device" buffers, every call that would enter libnsb200 is RECORDED instead of executed (vector copies
are carried out so that data-dependent host logic still works).  Residual norms follow a script, so that the Newton
loops of two runs take the same path.  Test infrastructure only."""
import numpy as np
import torch

import nsb200._lib as L
from nsb200.device import Device, Pattern
from nsb200.spaces import TaylorHoodSpaces


class RecordingDevice(Device):
    def __init__(self, mesh, norm_script=(1.0, 1e-6, 1e-13)):
        self.lib = L.load()                     # host-side pattern / tile / schedule builders run for real
        self.torch_device = torch.device("cpu")
        self.ctx = None
        self.partition = None
        self.mesh = mesh
        self.spaces = sp = TaylorHoodSpaces(mesh)
        self.dim = sp.dim
        self.n_owned_nodes, self.n_owned_verts = sp.n_nodes, sp.n_vertices
        self.patterns, self.coef = {}, None
        self.log, self._names, self._script, self._k = [], {}, tuple(norm_script), 0
        self.halo_transport = "none"

    # ---- recording ---------------------------------------------------------------------------------
    def _t(self, t):
        if t is None:
            return None
        key = (t.data_ptr(), t.numel())
        if key not in self._names:
            self._names[key] = "t%d" % len(self._names)
        return self._names[key]

    def _rec(self, name, *args):
        out = []
        for a in args:
            if isinstance(a, torch.Tensor):
                out.append(self._t(a))
            elif isinstance(a, float):
                out.append(float("%.12e" % a))
            else:
                out.append(a)
        self.log.append((name,) + tuple(out))

    @property
    def stream(self):
        return None

    def check(self, status):
        assert status == 0

    def pattern(self, name):
        if name not in self.patterns:
            sp = self.spaces
            if name == "P2P2":
                pat = Pattern(self, sp.n_nodes, sp.cell_nodes, sp.cell_nodes)
            elif name == "P1P1":
                pat = Pattern(self, sp.n_vertices, self.mesh.cells, self.mesh.cells)
            else:
                pat = Pattern(self, sp.ndofs_W, sp.cell_dofs_W, sp.cell_dofs_W)
            self.patterns[name] = pat
        return self.patterns[name]

    def matrix(self, pattern, vals, kind, bs, bc_mask=None, vals32=None, diag32=None, compress=False):
        self.pattern(pattern)
        return super().matrix(pattern, vals, kind, bs, bc_mask, vals32, diag32, compress=False)

    # ---- everything that would enter libnsb200 -------------------------------------------------------
    def set_coefficients(self, k, alpha, c_conv=1.0, c_p=1.0, c_visc=1.0, c_body=1.0):
        self.coef = (k, alpha[0], alpha[1], alpha[2], c_conv, c_p, c_visc, c_body)
        self._rec("set_coefficients", *[float(v) for v in self.coef[:7]])

    def set_forms(self, conv=0, visc=0, coriolis=None, generic=False):
        self._rec("set_forms", conv, visc)

    def assemble_ipcs_momentum(self, flags, ustar, un=None, unm1=None, pn=None, b=None, A=None):
        self._rec("assemble_ipcs_momentum", flags, ustar, un, unm1, pn, b, A)

    def assemble_poisson(self, flags, pn=None, ustar=None, b=None, A=None):
        self._rec("assemble_poisson", flags, pn, ustar, b, A)

    def assemble_correction(self, flags, ustar=None, pnew=None, pold=None, b=None, A=None):
        self._rec("assemble_correction", flags, ustar, pnew, pold, b, A)

    def assemble_mixed(self, flags, w0, w1=None, w2=None, b=None, A=None):
        self._rec("assemble_mixed", flags, w0, w1, w2, b, A)

    def set_mixed_bc(self, mask, g):
        self._mixed_bc = (mask, g)
        self._rec("set_mixed_bc", mask, g)

    def direct_solve(self, pattern, vals, b, x, reorder=1):
        self._rec("direct_solve", pattern, vals, b, x, reorder)
        x.zero_()

    def assemble_p1_mass(self, flags, w=None, A=None):
        self._rec("assemble_p1_mass", flags, w, A)

    def bc_apply_matrix(self, pattern, vals, kind, bs, bc_dofs):
        self._rec("bc_apply_matrix", pattern, vals, kind, bs, bc_dofs.numel())

    def bc_apply_vector(self, mode, b, bc_dofs, bc_vals, x=None):
        self._rec("bc_apply_vector", mode, b, bc_dofs.numel(), float(bc_vals.sum()) if bc_vals.numel() else 0.0, x)

    def inv_diagonal(self, A, out):
        self._rec("inv_diagonal", out)

    def lambda_max(self, A, inv_diag, iters=12, work=None, warm_start=False, lowp=False):
        self._rec("lambda_max", iters, bool(warm_start))
        return 2.0

    def rhs_basis(self, n, lmax=10):
        return dict(n=n, lmax=lmax, l=0)

    def rhs_guess(self, basis, b, x):
        self._rec("rhs_guess", basis["l"], b, x)
        return basis["l"]

    def rhs_update(self, basis, A, x):
        basis["l"] = min(basis["l"] + 1, basis["lmax"])
        self._rec("rhs_update", basis["l"], x)
        return basis["l"]

    def set_block_scalar_out(self, out):
        pass

    def block_scalar_rows_f32(self, vals, bs, block_rows, out):
        self._rec("block_scalar_f32", vals, bs, out)

    def block_scalar_f32(self, vals, bs, out):
        self._rec("block_scalar_f32", vals, bs, out)

    def to_f32(self, x, y):
        self._rec("to_f32", x, y)

    def cg(self, A, M, b, x, rtol=1e-12, atol=0.0, maxit=10000, nullspace=False, work=None):
        M = self._pc(M)
        self._rec("cg", A.kind, A.bs, M.kind, M.degree, b, x, rtol, bool(nullspace))
        return 5, 0.0

    def gmres(self, A, M, b, x, rtol=1e-12, atol=0.0, maxit=2000, restart=40, work=None, x0_is_zero=False):
        M = self._pc(M)
        self._rec("gmres", A.kind, A.bs, M.kind, M.degree, M.lowp, b, x, rtol, atol, restart, bool(x0_is_zero))
        return 7, 0.0

    def spmv(self, A, x, y):
        self._rec("spmv", x, y)

    def halo_exchange(self, which, bs, x):
        pass

    def halo_status(self):
        pass

    def copy(self, x, y):
        self._rec("copy", x, y)
        y.copy_(x)

    def axpy(self, a, x, y):
        self._rec("axpy", float(a), x, y)
        y.add_(x, alpha=float(a))

    def scale(self, a, x):
        self._rec("scale", float(a), x)
        x.mul_(float(a))

    def shift(self, x, s):
        self._rec("shift", x, float(s))
        x.sub_(float(s))

    def scatter(self, idx, x, y):
        self._rec("scatter", x, y)
        y[idx.long()] = x

    def gather(self, idx, x, y):
        self._rec("gather", x, y)
        y.copy_(x[idx.long()])

    def ipcs_advance(self, u0, u1, u2, p, p_old):
        self._rec("copy", u1, u2)
        self._rec("copy", u0, u1)
        self._rec("copy", p, p_old)
        u2.copy_(u1)
        u1.copy_(u0)
        p_old.copy_(p)

    def dot(self, x, y):
        return float(x @ y)

    def sum(self, x):
        return float(x.sum())

    def norm(self, x):
        v = self._script[self._k % len(self._script)]
        self._k += 1
        self._rec("norm", x)
        return v

    def launch_count(self):
        return len(self.log)

    def close(self):
        pass

    def __del__(self):
        pass


def numpy_of(t):
    return np.asarray(t)
