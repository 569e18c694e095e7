// assemble_forms.cuh -- one velocity row (test function phi_I e_b) of the momentum element tensors for EVERY
// weak-form variant of the reference, in closed form from the exact reference tensors:
//   convective term  source/ns_solver_base.py:370-390 (residual / Newton), :478-499 (Picard):
//       standard        (grad u . u) . w
//       rotational      (curl u x u) . w            = sum_k (d_k u_b - d_b u_k) u_k   (2D and 3D alike)
//       divergence      standard + 1/2 div(u) u . w
//       skew-symmetric  1/2 [ (grad u . u) . w - (grad w . u) . u ]
//   viscous term     :662-673   reduced  grad u : grad w   |   traction  (grad u + grad u^T) : sym grad w
//   Coriolis term    :173-191   2 c_Omega (Omega x u) . w   (2D: Omega = Omega_z)
// Every factor that is a derivative of a P2 function is P1 and therefore given exactly by its vertex values, so
// all integrands reduce to the tensors T_IJM = int phi_I phi_J psi_M, D2 and M2 of tables.h (App. B.2).
// Used by the generic one-thread-per-row kernels of assemble.cu (IPCS fallback path and the mixed kernel); the
// tuned IPCS kernels (assemble_jac.cu, assemble_res.cu) cover the default standard / reduced combination.
#pragma once

struct nsb_forms {
    int32_t conv;        // 0 standard | 1 rotational | 2 divergence | 3 skew-symmetric
    int32_t visc;        // 0 reduced | 1 traction
    double cor[3];       // 2 c_Omega Omega (2D: cor[2] = 2 c_Omega Omega_z); all zero: no Coriolis term
};
#define NSB_CONV_STANDARD 0
#define NSB_CONV_ROTATIONAL 1
#define NSB_CONV_DIVERGENCE 2
#define NSB_CONV_SKEW 3

template <int D> struct Tab;

template <int D>
struct CellCtx {
    static constexpr int NL = D == 2 ? 6 : 10, NV = D + 1;
    Geo<D> g;
    double U[NL][D];
    double G[D][D];            // detJ K K^T
    double gM[NV][D][D];       // gM[M][b][a] = d_a u_b (x_M)
    double dv[NV];             // div u (x_M)
};

template <int D>
__device__ __forceinline__ void cell_context(CellCtx<D>& c) {
    metric<D>(c.g, c.G);
    vertex_gradients<D>(c.g, c.U, c.gM);
#pragma unroll
    for (int M = 0; M <= D; ++M) {
        double s = 0.0;
#pragma unroll
        for (int a = 0; a < D; ++a) s += c.gM[M][a][a];
        c.dv[M] = s;
    }
}

// physical gradient of phi_J at vertex M
template <int D>
__device__ __forceinline__ void grad_phi_vertex(const Geo<D>& g, int J, int M, double (&out)[D]) {
    using TB = Tab<D>;
#pragma unroll
    for (int a = 0; a < D; ++a) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < D; ++k) s += g.K[k][a] * TB::DV(J, M, k);
        out[a] = s;
    }
}

// Row (b, I): Ae[a*NL + J] = dF_(b,I) / dU_(a,J) (want_mat) and the residual contribution Fe of the mass / BDF,
// viscous, convective and Coriolis terms (no pressure term).  acc_b[J] = (a0 u + a1 u^n + a2 u^{n-1})_b at node J
// (nullptr: no BDF term).  newton = 0: Picard linearisation (convecting velocity frozen).
template <int D>
__device__ void velocity_row(const CellCtx<D>& c, int I, int b, const double* acc_b, double mass_coef, double inv_k,
                             const nsb_coef& cf, const nsb_forms& fm, bool newton, bool want_mat, double* Ae, double& Fe) {
    using TB = Tab<D>;
    constexpr int NL = CellCtx<D>::NL, NV = D + 1;
    const double dj = c.g.detJ;
    // W[M][j] = sum_L T(I,L,M) U_L[j]
    double W[NV][D];
#pragma unroll
    for (int M = 0; M < NV; ++M)
#pragma unroll
        for (int j = 0; j < D; ++j) W[M][j] = 0.0;
    for (int L = 0; L < NL; ++L)
#pragma unroll
        for (int M = 0; M < NV; ++M) {
            const double t = TB::T(I, L, M);
#pragma unroll
            for (int j = 0; j < D; ++j) W[M][j] += t * c.U[L][j];
        }
    double GI[NV][D];           // grad phi_I at the vertices (skew form)
    if (fm.conv == NSB_CONV_SKEW) {
#pragma unroll
        for (int M = 0; M < NV; ++M) grad_phi_vertex<D>(c.g, I, M, GI[M]);
    }
    double F = 0.0;
    // ---- convective residual -----------------------------------------------------------------------
    {
        double s = 0.0;
#pragma unroll
        for (int M = 0; M < NV; ++M) {
#pragma unroll
            for (int j = 0; j < D; ++j) {
                double coef = c.gM[M][b][j];
                if (fm.conv == NSB_CONV_ROTATIONAL) coef -= c.gM[M][j][b];
                s += coef * W[M][j];
            }
            if (fm.conv == NSB_CONV_DIVERGENCE) s += 0.5 * c.dv[M] * W[M][b];
        }
        if (fm.conv == NSB_CONV_SKEW) {
            // -1/2 sum_M sum_j GI[M][j] Q[M][j][b],  Q[M][j][b] = sum_L sum_N U_Lj U_Nb T(L,N,M)
            double q = 0.0;
            for (int L = 0; L < NL; ++L)
                for (int N = 0; N < NL; ++N) {
                    const double unb = c.U[N][b];
#pragma unroll
                    for (int M = 0; M < NV; ++M) {
                        double gu = 0.0;
#pragma unroll
                        for (int j = 0; j < D; ++j) gu += GI[M][j] * c.U[L][j];
                        q += TB::T(L, N, M) * gu * unb;
                    }
                }
            s = 0.5 * s - 0.5 * q;
        }
        F += cf.c_conv * dj * s;
    }
    // ---- Coriolis residual: (cor x u)_b phi_I -------------------------------------------------------
    const bool coriolis = fm.cor[0] != 0.0 || fm.cor[1] != 0.0 || fm.cor[2] != 0.0;
    // cross-product matrix X[b][k]: (cor x u)_b = sum_k X[b][k] u_k
    double X[D][D];
#pragma unroll
    for (int i = 0; i < D; ++i)
#pragma unroll
        for (int k = 0; k < D; ++k) X[i][k] = 0.0;
    if (coriolis) {
        if (D == 2) {
            X[0][1] = -fm.cor[2]; X[1][0] = fm.cor[2];
        } else {
            X[0][D - 2] = -fm.cor[2]; X[0][D - 1] = fm.cor[1];
            X[D - 2][0] = fm.cor[2];  X[D - 2][D - 1] = -fm.cor[0];
            X[D - 1][0] = -fm.cor[1]; X[D - 1][D - 2] = fm.cor[0];
        }
    }
    for (int J = 0; J < NL; ++J) {
        const double mm = dj * TB::M2(I, J);
        // viscous: reduced part (scalar) and the traction extra E[a][b] = detJ sum_jk K[j][a] K[k][b] D2(I,J,j,k)
        double vis = 0.0;
#pragma unroll
        for (int j = 0; j < D; ++j)
#pragma unroll
            for (int k = 0; k < D; ++k) vis += c.G[j][k] * TB::D2(I, J, j, k);
        vis *= cf.c_visc;
        double E[D];            // E[a] = c_visc * int d_a phi_I d_b phi_J   (row b fixed)
#pragma unroll
        for (int a = 0; a < D; ++a) E[a] = 0.0;
        if (fm.visc == 1) {
#pragma unroll
            for (int a = 0; a < D; ++a) {
                double s = 0.0;
#pragma unroll
                for (int j = 0; j < D; ++j)
#pragma unroll
                    for (int k = 0; k < D; ++k) s += c.g.K[j][a] * c.g.K[k][b] * TB::D2(I, J, j, k);
                E[a] = cf.c_visc * dj * s;
            }
        }
        // residual: mass / BDF, viscous, Coriolis
        F += (acc_b ? inv_k * mm * acc_b[J] : 0.0) + vis * c.U[J][b];
#pragma unroll
        for (int a = 0; a < D; ++a) F += E[a] * c.U[J][a] + (coriolis ? X[b][a] * mm * c.U[J][a] : 0.0);
        if (!want_mat) continue;
        // ---- Jacobian row entries (a, J) ---------------------------------------------------------------
        double GJ[NV][D];
#pragma unroll
        for (int M = 0; M < NV; ++M) grad_phi_vertex<D>(c.g, J, M, GJ[M]);
        double cv = 0.0;        // int (u . grad phi_J) phi_I / detJ
        double tdv = 0.0;       // sum_M div u (x_M) T(I,J,M)
#pragma unroll
        for (int M = 0; M < NV; ++M) {
#pragma unroll
            for (int j = 0; j < D; ++j) cv += GJ[M][j] * W[M][j];
            tdv += c.dv[M] * TB::T(I, J, M);
        }
        double cvT = 0.0;       // int (u . grad phi_I) phi_J / detJ   (skew form)
        double WJ[NV][D];       // sum_N T(J,N,M) U_N[c]               (skew form)
        if (fm.conv == NSB_CONV_SKEW) {
#pragma unroll
            for (int M = 0; M < NV; ++M)
#pragma unroll
                for (int j = 0; j < D; ++j) WJ[M][j] = 0.0;
            for (int N = 0; N < NL; ++N)
#pragma unroll
                for (int M = 0; M < NV; ++M) {
                    const double t = TB::T(J, N, M);
#pragma unroll
                    for (int j = 0; j < D; ++j) WJ[M][j] += t * c.U[N][j];
                }
#pragma unroll
            for (int M = 0; M < NV; ++M)
#pragma unroll
                for (int j = 0; j < D; ++j) cvT += GI[M][j] * WJ[M][j];
        }
#pragma unroll
        for (int a = 0; a < D; ++a) {
            double e = 0.0;
            // scalar (delta_ab) part of the convective linearisation
            double sc = 0.0, blk = 0.0;
            switch (fm.conv) {
                case NSB_CONV_STANDARD:
                    sc = cv;
                    if (newton) {
#pragma unroll
                        for (int M = 0; M < NV; ++M) blk += TB::T(I, J, M) * c.gM[M][b][a];
                    }
                    break;
                case NSB_CONV_ROTATIONAL:
                    // Picard: curl(u_frozen) x v ; Newton adds curl(v) x u = delta_ab cv - sum_M GJ[M][b] W[M][a]
#pragma unroll
                    for (int M = 0; M < NV; ++M) blk += TB::T(I, J, M) * (c.gM[M][b][a] - c.gM[M][a][b]);
                    if (newton) {
                        sc = cv;
#pragma unroll
                        for (int M = 0; M < NV; ++M) blk -= GJ[M][b] * W[M][a];
                    }
                    break;
                case NSB_CONV_DIVERGENCE:
                    sc = cv + 0.5 * tdv;
                    if (newton) {
#pragma unroll
                        for (int M = 0; M < NV; ++M) blk += TB::T(I, J, M) * c.gM[M][b][a] + 0.5 * GJ[M][a] * W[M][b];
                    }
                    break;
                default:    // skew-symmetric
                    sc = 0.5 * (cv - cvT);
                    if (newton) {
#pragma unroll
                        for (int M = 0; M < NV; ++M) blk += 0.5 * TB::T(I, J, M) * c.gM[M][b][a] - 0.5 * GI[M][a] * WJ[M][b];
                    }
                    break;
            }
            e = cf.c_conv * dj * ((a == b ? sc : 0.0) + blk);
            if (a == b) e += mass_coef * mm + vis;
            e += E[a];
            if (coriolis) e += X[b][a] * mm;
            Ae[a * NL + J] = e;
        }
    }
    Fe = F;
}
