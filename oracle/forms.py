"""Oracle weak forms: element tensors by quadrature + dolfin-style global assembly.

Index conventions (UFL ``grad(u)[i,j] = d u_i / d x_j``,
``dot(grad(u),u)_i = sum_j d_j u_i u_j``):  I,J scalar P2 basis, b test component,
a trial component, M,N P1 basis.

Forms restated:
* IPCS tentative velocity F / J        source/ns_ipcs_solver.py:106-147, :19-33
* IPCS pressure Poisson lhs / rhs      source/ns_ipcs_solver.py:149-171
* IPCS velocity correction lhs / rhs   source/ns_ipcs_solver.py:173-196
* monolithic BDF F / J                 source/ns_bdf_solver.py:36-100, :19-34
* stationary F / J_picard / J_newton   source/ns_solver_base.py:894-949
with the term builders source/ns_solver_base.py:370-399 (convective, standard
form; divergence), :478-499 (Picard), :662-673 (viscous, reduced form).

Test infrastructure only (see ``oracle/__init__.py``).
"""
import numpy as np
import scipy.sparse as sp

from . import fem
from .dofmap import sparsity_pattern, entry_positions

QUAD_DEGREE = 6      # >= 5 = FFC's estimate for the momentum forms (App. A.8); all integrands are polynomials
CHUNK = 4096


class Assembler:
    """Holds mesh geometry, tabulations, patterns and cell->nnz maps for one mesh."""

    def __init__(self, spaces):
        self.sp = spaces
        m = spaces.mesh
        self.d = d = m.dim
        self.K, self.detJ = fem.geometry(m.coords, m.cells)
        self.qp, self.qw = fem.simplex_quadrature(d, QUAD_DEGREE)
        self.phi2, self.dphi2 = fem.tabulate_p2(self.qp)
        self.phi1, self.dphi1 = fem.tabulate_p1(self.qp)
        self._patterns = {}
        # weak-form variants (source/ns_solver_base.py:78-115): restated at the quadrature points, term by term
        self.conv_form = "standard"      # | "rotational" | "divergence" | "skew_symmetric"   (:370-390, :478-499)
        self.visc_form = "reduced"       # | "traction"                                        (:662-673)
        self.coriolis = None             # 2 c_Omega Omega as a 3-vector (2D: only [2]), :173-191

    # ---- patterns ---------------------------------------------------------------------------
    def pattern(self, name):
        """name in {"VV","QQ","WW"} -> (rowptr, col, pos[c,i,j])."""
        if name not in self._patterns:
            s = self.sp
            cd, n = {"VV": (s.cell_dofs_V, s.ndofs_V), "QQ": (s.cell_dofs_Q, s.ndofs_Q),
                     "WW": (s.cell_dofs_W, s.ndofs_W)}[name]
            rowptr, col = sparsity_pattern(cd, cd, n, n)
            pos = entry_positions(rowptr, col, n, cd, cd)
            self._patterns[name] = (rowptr, col, pos)
        return self._patterns[name]

    def _scatter_matrix(self, name, Ae_iter):
        rowptr, col, pos = self.pattern(name)
        vals = np.zeros(len(col))
        for s, Ae in Ae_iter:
            p = pos[s:s + Ae.shape[0]].ravel()
            vals += np.bincount(p, weights=Ae.ravel(), minlength=len(col))
        n = len(rowptr) - 1
        return sp.csr_matrix((vals, col, rowptr), shape=(n, n))

    def _scatter_vector(self, cell_dofs, n, be_iter):
        b = np.zeros(n)
        for s, be in be_iter:
            b += np.bincount(cell_dofs[s:s + be.shape[0]].ravel(), weights=be.ravel(), minlength=n)
        return b

    def _chunks(self):
        nc = self.sp.mesh.num_cells
        for s in range(0, nc, CHUNK):
            yield s, min(nc, s + CHUNK)

    # ---- per-chunk quadrature data ------------------------------------------------------------
    def _geo(self, s, e):
        K = self.K[s:e]
        wdet = self.detJ[s:e, None] * self.qw[None, :]                    # [c,q]
        G2 = np.einsum("cji,qIj->cqIi", K, self.dphi2)                    # physical grad P2
        G1 = np.einsum("cji,Mj->cMi", K, self.dphi1[0])                   # physical grad P1 (const)
        return wdet, G2, G1

    def _cell_vel(self, u, s, e):
        """V-space vector -> U[c,I,a]."""
        d = self.d
        nodes = self.sp.cell_nodes[s:e]
        return u.reshape(-1, d)[nodes]                                     # [c,I,a]

    def _cell_vel_W(self, w, s, e):
        d, nl = self.d, self.sp.nloc2
        dofs = self.sp.cell_dofs_W[s:e, :d * nl]
        return np.transpose(w[dofs].reshape(-1, d, nl), (0, 2, 1))         # [c,I,a]

    def _cell_p_W(self, w, s, e):
        d, nl = self.d, self.sp.nloc2
        return w[self.sp.cell_dofs_W[s:e, d * nl:]]

    # ---- element tensors ------------------------------------------------------------------------
    @staticmethod
    def _pad3(x, axis):
        """embed a 2-vector index into 3D (zero z component)"""
        if x.shape[axis] == 3:
            return x
        pad = [(0, 0)] * x.ndim
        pad[axis] = (0, 1)
        return np.pad(x, pad)

    _EPS = np.zeros((3, 3, 3))
    for _i, _j, _k in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
        _EPS[_i, _j, _k], _EPS[_i, _k, _j] = 1.0, -1.0

    def _cross_matrix(self):
        """X[b,k] with (Omega x u)_b = X[b,k] u_k, restricted to the mesh dimension"""
        om = np.asarray(self.coriolis, dtype=np.float64)
        X = np.einsum("bik,i->bk", self._EPS, om)
        return X[:self.d, :self.d]

    def _momentum_residual(self, wdet, G2, G1, U, Uacc, P, c_conv, c_p, c_visc, inv_k):
        """F[c,I,b] of  (acc.w)/k + c_c conv(u).w - c_p div(w) p + c_nu visc(u,w)  [+ 2 c_Om (Om x u).w].
        Uacc = sum_i alpha_i U_i (or None for the stationary form); P[c,M] pressure dofs."""
        phi2, phi1 = self.phi2, self.phi1
        d = self.d
        uq = np.einsum("qI,cIa->cqa", phi2, U)
        gu = np.einsum("cqIj,cIb->cqbj", G2, U)                            # d_j u_b
        conv = np.einsum("cqbj,cqj->cqb", gu, uq)                          # dot(grad(u), u)
        if self.conv_form == "rotational":                                 # cross(curl(u), u); 2D: (-curl u u_y, curl u u_x)
            gu3, u3 = self._pad3(self._pad3(gu, 2), 3), self._pad3(uq, 2)
            curl = np.einsum("imn,cqnm->cqi", self._EPS, gu3)
            conv = np.einsum("bik,cqi,cqk->cqb", self._EPS, curl, u3)[:, :, :d]
        elif self.conv_form == "divergence":
            conv = conv + 0.5 * np.einsum("cqjj,cqb->cqb", gu, uq)
        F = c_conv * np.einsum("cq,cqb,qI->cIb", wdet, conv, phi2)
        if self.conv_form == "skew_symmetric":                            # 1/2 [conv.w - dot(dot(grad(w), u), u)]
            F = 0.5 * F - 0.5 * c_conv * np.einsum("cq,cqIj,cqj,cqb->cIb", wdet, G2, uq, uq)
        pq = np.einsum("qM,cM->cq", phi1, P)
        F -= c_p * np.einsum("cq,cq,cqIb->cIb", wdet, pq, G2)
        if self.visc_form == "traction":                                   # (grad u + grad u^T) : 1/2 (grad w + grad w^T)
            F += c_visc * np.einsum("cq,cqbj,cqIj->cIb", wdet, gu + np.swapaxes(gu, 2, 3), G2)
        else:
            F += c_visc * np.einsum("cq,cqbj,cqIj->cIb", wdet, gu, G2)
        if self.coriolis is not None:
            F += np.einsum("cq,bk,cqk,qI->cIb", wdet, self._cross_matrix(), uq, phi2)
        if Uacc is not None:
            aq = np.einsum("qI,cIa->cqa", phi2, Uacc)
            F += inv_k * np.einsum("cq,cqb,qI->cIb", wdet, aq, phi2)
        return F

    def _momentum_jacobian(self, wdet, G2, U, c_conv, c_visc, mass_coef, newton=True):
        """A[c,I,b,J,a]: derivative of the residual w.r.t. the coefficient of phi_J e_a (newton) or the reference's
        Picard linearisation with the convecting velocity frozen (source/ns_solver_base.py:478-499)."""
        phi2 = self.phi2
        d = self.d
        nl = phi2.shape[1]
        eye = np.eye(d)
        uq = np.einsum("qI,cIa->cqa", phi2, U)
        gu = np.einsum("cqIj,cIb->cqbj", G2, U)                            # d_j u_b
        adv = np.einsum("cqJj,cqj->cqJ", G2, uq)                           # u . grad phi_J
        mI_adv = np.einsum("cq,qI,cqJ->cIJ", wdet, phi2, adv)              # int (u.grad phi_J) phi_I
        mIJ_gu = np.einsum("cq,qI,qJ,cqba->cIbJa", wdet, phi2, phi2, gu)   # int phi_J (d_a u_b) phi_I
        A = np.zeros((U.shape[0], nl, d, nl, d))
        form = self.conv_form
        if form == "standard":
            A += np.einsum("cIJ,ba->cIbJa", mI_adv, eye)
            if newton:
                A += mIJ_gu
        elif form == "divergence":
            divu = np.einsum("cqjj->cq", gu)
            A += np.einsum("cIJ,ba->cIbJa", mI_adv + 0.5 * np.einsum("cq,cq,qI,qJ->cIJ", wdet, divu, phi2, phi2), eye)
            if newton:
                A += mIJ_gu + 0.5 * np.einsum("cq,cqJa,cqb,qI->cIbJa", wdet, G2, uq, phi2)
        elif form == "skew_symmetric":
            A += 0.5 * np.einsum("cIJ,ba->cIbJa", mI_adv - np.swapaxes(mI_adv, 1, 2), eye)
            if newton:
                A += 0.5 * mIJ_gu - 0.5 * np.einsum("cq,cqIa,qJ,cqb->cIbJa", wdet, G2, phi2, uq)
        elif form == "rotational":
            gu3, u3 = self._pad3(self._pad3(gu, 2), 3), self._pad3(uq, 2)
            curl = np.einsum("imn,cqnm->cqi", self._EPS, gu3)
            # cross(curl(u), v):  eps[b,i,a] curl_i phi_J
            A += np.einsum("cq,bia,cqi,qI,qJ->cIbJa", wdet, self._EPS, curl, phi2, phi2)[:, :, :d, :, :d]
            if newton:
                # cross(curl(v), u), v = phi_J e_a:  curl(v)_i = eps[i,m,a] d_m phi_J
                G3 = self._pad3(G2, 3)
                A += np.einsum("cq,bik,ima,cqJm,cqk,qI->cIbJa", wdet, self._EPS, self._EPS, G3, u3, phi2)[:, :, :d, :, :d]
        else:
            raise ValueError(form)
        A *= c_conv
        S = c_visc * np.einsum("cq,cqIj,cqJj->cIJ", wdet, G2, G2)
        if mass_coef != 0.0:
            S += mass_coef * np.einsum("cq,qI,qJ->cIJ", wdet, phi2, phi2)
        A += np.einsum("cIJ,ba->cIbJa", S, eye)
        if self.visc_form == "traction":
            A += c_visc * np.einsum("cq,cqIa,cqJb->cIbJa", wdet, G2, G2)
        if self.coriolis is not None:
            A += np.einsum("cq,ba,qI,qJ->cIbJa", wdet, self._cross_matrix(), phi2, phi2)
        return A

    def _div_block(self, wdet, G2, c_p):
        """B[c,I,b,M] = -c_p int d_b phi_I psi_M."""
        return -c_p * np.einsum("cq,cqIb,qM->cIbM", wdet, G2, self.phi1)

    @staticmethod
    def _to_V_local(T):
        """[c,I,b,...] -> V cell-local order index b*nloc+I."""
        c, nl, d = T.shape[:3]
        return np.swapaxes(T, 1, 2).reshape((c, d * nl) + T.shape[3:])

    # ---- IPCS forms --------------------------------------------------------------------------------
    def ipcs_momentum_residual(self, ustar, un, unm1, pn, alpha, k, c_conv, c_p, c_visc):
        s_ = self.sp

        def it():
            for s, e in self._chunks():
                wdet, G2, G1 = self._geo(s, e)
                U = self._cell_vel(ustar, s, e)
                Uacc = (alpha[0] * U + alpha[1] * self._cell_vel(un, s, e)
                        + alpha[2] * self._cell_vel(unm1, s, e))
                P = pn[s_.cell_dofs_Q[s:e]]
                F = self._momentum_residual(wdet, G2, G1, U, Uacc, P, c_conv, c_p, c_visc, 1.0 / k)
                yield s, self._to_V_local(F)
        return self._scatter_vector(s_.cell_dofs_V, s_.ndofs_V, it())

    def ipcs_momentum_jacobian(self, ustar, alpha0, k, c_conv, c_visc, newton=True):
        def it():
            for s, e in self._chunks():
                wdet, G2, _ = self._geo(s, e)
                U = self._cell_vel(ustar, s, e)
                A = self._momentum_jacobian(wdet, G2, U, c_conv, c_visc, alpha0 / k, newton)
                A = self._to_V_local(A)                                   # [c, bI, J, a]
                A = np.swapaxes(A, 2, 3).reshape(A.shape[0], A.shape[1], -1)   # [c, bI, aJ]
                yield s, A
        return self._scatter_matrix("VV", it())

    def poisson_matrix(self):
        def it():
            for s, e in self._chunks():
                wdet, _, G1 = self._geo(s, e)
                vol = wdet.sum(axis=1)
                yield s, np.einsum("c,cMi,cNi->cMN", vol, G1, G1)
        return self._scatter_matrix("QQ", it())

    def poisson_rhs(self, pn, ustar, alpha0, k):
        """int grad(p^n).grad(q) - (alpha0/k) div(u*) q."""
        s_ = self.sp

        def it():
            for s, e in self._chunks():
                wdet, G2, G1 = self._geo(s, e)
                U = self._cell_vel(ustar, s, e)
                P = pn[s_.cell_dofs_Q[s:e]]
                gp = np.einsum("cMi,cM->ci", G1, P)
                divu = np.einsum("cqIb,cIb->cq", G2, U)
                vol = wdet.sum(axis=1)
                b = np.einsum("c,ci,cMi->cM", vol, gp, G1)
                b -= (alpha0 / k) * np.einsum("cq,cq,qM->cM", wdet, divu, self.phi1)
                yield s, b
        return self._scatter_vector(s_.cell_dofs_Q, s_.ndofs_Q, it())

    def velocity_mass_matrix(self):
        d = self.d

        def it():
            for s, e in self._chunks():
                wdet, _, _ = self._geo(s, e)
                Ms = np.einsum("cq,qI,qJ->cIJ", wdet, self.phi2, self.phi2)
                nl = Ms.shape[1]
                A = np.zeros((Ms.shape[0], d, nl, d, nl))
                for a in range(d):
                    A[:, a, :, a, :] = Ms
                yield s, A.reshape(Ms.shape[0], d * nl, d * nl)
        return self._scatter_matrix("VV", it())

    def correction_rhs(self, ustar, pnew, pold, alpha0, k):
        """int u*.w - (k/alpha0) grad(p - p_old).w."""
        s_ = self.sp

        def it():
            for s, e in self._chunks():
                wdet, _, G1 = self._geo(s, e)
                U = self._cell_vel(ustar, s, e)
                dP = (pnew - pold)[s_.cell_dofs_Q[s:e]]
                gp = np.einsum("cMi,cM->ci", G1, dP)
                uq = np.einsum("qI,cIa->cqa", self.phi2, U)
                f = uq - (k / alpha0) * gp[:, None, :]
                F = np.einsum("cq,cqb,qI->cIb", wdet, f, self.phi2)
                yield s, self._to_V_local(F)
        return self._scatter_vector(s_.cell_dofs_V, s_.ndofs_V, it())

    # ---- mixed Taylor-Hood forms ------------------------------------------------------------------------
    def mixed_residual(self, w0, c_conv, c_p, c_visc, bdf=None):
        """F(w,q) = -c_p div(u) q + momentum(w).  bdf = (alpha, k, w1, w2) adds the BDF mass term."""
        s_ = self.sp
        d, nl = self.d, s_.nloc2

        def it():
            for s, e in self._chunks():
                wdet, G2, G1 = self._geo(s, e)
                U = self._cell_vel_W(w0, s, e)
                P = self._cell_p_W(w0, s, e)
                Uacc, inv_k = None, 0.0
                if bdf is not None:
                    alpha, k, w1, w2 = bdf
                    Uacc = (alpha[0] * U + alpha[1] * self._cell_vel_W(w1, s, e)
                            + alpha[2] * self._cell_vel_W(w2, s, e))
                    inv_k = 1.0 / k
                F = self._momentum_residual(wdet, G2, G1, U, Uacc, P, c_conv, c_p, c_visc, inv_k)
                divu = np.einsum("cqIb,cIb->cq", G2, U)
                Fp = -c_p * np.einsum("cq,cq,qM->cM", wdet, divu, self.phi1)
                yield s, np.concatenate([self._to_V_local(F), Fp], axis=1)
        return self._scatter_vector(s_.cell_dofs_W, s_.ndofs_W, it())

    def mixed_jacobian(self, w0, c_conv, c_p, c_visc, mass_coef=0.0, newton=True):
        s_ = self.sp
        d, nl = self.d, s_.nloc2
        nv = d * nl
        nw = nv + d + 1

        def it():
            for s, e in self._chunks():
                wdet, G2, _ = self._geo(s, e)
                U = self._cell_vel_W(w0, s, e)
                A = self._momentum_jacobian(wdet, G2, U, c_conv, c_visc, mass_coef, newton)
                A = self._to_V_local(A)
                A = np.swapaxes(A, 2, 3).reshape(A.shape[0], nv, nv)
                B = self._to_V_local(self._div_block(wdet, G2, c_p))       # [c, bI, M]
                Ae = np.zeros((A.shape[0], nw, nw))
                Ae[:, :nv, :nv] = A
                Ae[:, :nv, nv:] = B
                Ae[:, nv:, :nv] = np.swapaxes(B, 1, 2)
                yield s, Ae
        return self._scatter_matrix("WW", it())

    # ---- functionals / projections ------------------------------------------------------------------------
    def pressure_mass_matrix(self):
        def it():
            for s, e in self._chunks():
                wdet, _, _ = self._geo(s, e)
                yield s, np.einsum("cq,qM,qN->cMN", wdet, self.phi1, self.phi1)
        return self._scatter_matrix("QQ", it())

    def volume(self):
        return float(self.detJ.sum() * self.qw.sum())

    def integral_Q(self, p):
        """int p dx for p in Q (source/ns_solver_base.py:1194)."""
        P = p[self.sp.cell_dofs_Q]
        return float(np.einsum("c,q,qM,cM->", self.detJ, self.qw, self.phi1, P))
