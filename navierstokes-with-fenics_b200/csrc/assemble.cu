// assemble.cu -- atomic-free, write-once "row-gather" finite-element assembly kernels.
//
// One thread owns one matrix row (a scalar P2 node / P1 vertex / mixed dof).  It walks the
// cells incident to that row in ascending cell order (the summation order of dolfin's cell
// loop), evaluates the row of every element tensor in closed form from
// exact reference tensors (tables.h, no quadrature), and adds it into its own row of the
// CSR/BSR value array through the uint8 slot map.  No atomics, no colouring, every value is
// written by exactly one thread, results are bit-reproducible run to run.
//
// Forms (index notation of SURVEY.md App. B.2):
//   K1/K2  IPCS tentative velocity F / dF/du*     source/ns_ipcs_solver.py:106-147
//   K3/K4  pressure Poisson a / L                 source/ns_ipcs_solver.py:149-171
//   K5/K6  velocity correction a / L              source/ns_ipcs_solver.py:173-196
//   K7/K8  mixed Taylor-Hood F / J                source/ns_bdf_solver.py:68-88,
//                                                 source/ns_solver_base.py:913-936
#include "common.cuh"
#include "tables.h"
#include "fe_geometry.cuh"

// ---- table access by dimension -------------------------------------------------------------
template <int D> struct Tab;
template <> struct Tab<2> {
    static constexpr int NL2 = 6, NL1 = 3;
    static constexpr double INV_FACT = 0.5;
    __device__ static __forceinline__ double M2(int i, int j) { return M2_2D[i][j]; }
    __device__ static __forceinline__ double m2(int i) { return m2_2D[i]; }
    __device__ static __forceinline__ double M1(int i, int j) { return M1_2D[i][j]; }
    __device__ static __forceinline__ double m1(int i) { return m1_2D[i]; }
    __device__ static __forceinline__ double D2(int i, int j, int a, int b) { return D2_2D[i][j][a][b]; }
    __device__ static __forceinline__ double DP(int i, int a, int m) { return DP_2D[i][a][m]; }
    __device__ static __forceinline__ double T(int i, int j, int m) { return T_2D[i][j][m]; }
    __device__ static __forceinline__ double DV(int j, int m, int k) { return DV_2D[j][m][k]; }
};
template <> struct Tab<3> {
    static constexpr int NL2 = 10, NL1 = 4;
    static constexpr double INV_FACT = 1.0 / 6.0;
    __device__ static __forceinline__ double M2(int i, int j) { return M2_3D[i][j]; }
    __device__ static __forceinline__ double m2(int i) { return m2_3D[i]; }
    __device__ static __forceinline__ double M1(int i, int j) { return M1_3D[i][j]; }
    __device__ static __forceinline__ double m1(int i) { return m1_3D[i]; }
    __device__ static __forceinline__ double D2(int i, int j, int a, int b) { return D2_3D[i][j][a][b]; }
    __device__ static __forceinline__ double DP(int i, int a, int m) { return DP_3D[i][a][m]; }
    __device__ static __forceinline__ double T(int i, int j, int m) { return T_3D[i][j][m]; }
    __device__ static __forceinline__ double DV(int j, int m, int k) { return DV_3D[j][m][k]; }
};

// velocity gradient at the cell vertices: gM[M][b][a] = d_a u_b (x_M)   (P1 in x, exact)
template <int D>
__device__ __forceinline__ void vertex_gradients(const Geo<D>& g, const double (&U)[Tab<D>::NL2][D],
                                                 double (&gM)[D + 1][D][D]) {
    using TB = Tab<D>;
#pragma unroll
    for (int M = 0; M <= D; ++M) {
        double ref[D][D];   // ref[b][k] = sum_L U_Lb dhat_k phi_L(x_M)
#pragma unroll
        for (int b = 0; b < D; ++b)
#pragma unroll
            for (int k = 0; k < D; ++k) ref[b][k] = 0.0;
        for (int L = 0; L < TB::NL2; ++L)
#pragma unroll
            for (int k = 0; k < D; ++k) {
                const double d = TB::DV(L, M, k);
#pragma unroll
                for (int b = 0; b < D; ++b) ref[b][k] += U[L][b] * d;
            }
#pragma unroll
        for (int b = 0; b < D; ++b)
#pragma unroll
            for (int a = 0; a < D; ++a) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < D; ++k) s += ref[b][k] * g.K[k][a];
                gM[M][b][a] = s;
            }
    }
}

#include "assemble_forms.cuh"

// ===========================================================================================
// K2  IPCS momentum Jacobian, BSR-D rows
//   A[(I,b),(J,a)] = delta_ab S_IJ + c_c detJ sum_M T_IJM gM[b][a]
//   S_IJ = (a0/k) detJ M2_IJ + c_nu sum_jk G_jk D2_IJjk + c_c sum_M y_IM . DV_JM
//   y_IM[k] = sum_L T_ILM z_L[k],  z_L[k] = detJ sum_i K[k][i] U_L[i]
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_momentum_jacobian(int32_t nrows, const int32_t* __restrict__ rowptr,
                    const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
                    const uint8_t* __restrict__ pair_iloc, const uint8_t* __restrict__ pair_slots,
                    const double* __restrict__ coords, const int32_t* __restrict__ cell_nodes,
                    const int32_t* __restrict__ cell_verts,
                    const double* __restrict__ ustar, double mass_coef, nsb_coef cf, nsb_forms fm,
                    int newton, double* __restrict__ vals) {
    using TB = Tab<D>;
    constexpr int NL = TB::NL2, DD = D * D;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    const int64_t base = (int64_t)rowptr[row] * DD;
    const int64_t nval = (int64_t)(rowptr[row + 1] - rowptr[row]) * DD;
    for (int64_t s = 0; s < nval; ++s) vals[base + s] = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int I = pair_iloc[p];
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
        CellCtx<D> cc;
        cell_geometry(coords, vtx, cc.g);
        for (int L = 0; L < NL; ++L) {
            const int64_t nd = cell_nodes[(int64_t)c * NL + L];
#pragma unroll
            for (int b = 0; b < D; ++b) cc.U[L][b] = ustar[(int64_t)D * nd + b];
        }
        cell_context<D>(cc);
        for (int b = 0; b < D; ++b) {
            double Ae[D * NL], Fe;
            velocity_row<D>(cc, I, b, nullptr, mass_coef, 0.0, cf, fm, newton != 0, true, Ae, Fe);
            for (int J = 0; J < NL; ++J) {
                double* dst = vals + base + (int64_t)pair_slots[(int64_t)p * NL + J] * DD;
#pragma unroll
                for (int a = 0; a < D; ++a) dst[b * D + a] += Ae[a * NL + J];
            }
        }
    }
}

// ===========================================================================================
// K1  IPCS momentum residual, one thread per P2 node (D components)
//   F_Ib = (1/k) detJ sum_L M2_IL acc_Lb + c_nu sum_L A_IL U_Lb
//        + c_c detJ sum_LM T_ILM sum_j U_Lj gM[b][j] - c_p detJ sum_M P_M sum_j K[j][b] DP_IjM
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_momentum_residual(int32_t nrows, const int32_t* __restrict__ pair_ptr,
                    const int32_t* __restrict__ pair_cell, const uint8_t* __restrict__ pair_iloc,
                    const double* __restrict__ coords, const int32_t* __restrict__ cell_nodes,
                    const int32_t* __restrict__ cell_verts,
                    const double* __restrict__ ustar, const double* __restrict__ un,
                    const double* __restrict__ unm1, const double* __restrict__ pn, nsb_coef cf, nsb_forms fm,
                    double* __restrict__ b_out) {
    using TB = Tab<D>;
    constexpr int NL = TB::NL2;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    double F[D];
#pragma unroll
    for (int b = 0; b < D; ++b) F[b] = 0.0;
    const double inv_k = 1.0 / cf.k;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int I = pair_iloc[p];
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
        CellCtx<D> cc;
        cell_geometry(coords, vtx, cc.g);
        double acc[D][NL];
        for (int L = 0; L < NL; ++L) {
            const int64_t nd = cell_nodes[(int64_t)c * NL + L];
#pragma unroll
            for (int b = 0; b < D; ++b) {
                const int64_t dof = (int64_t)D * nd + b;
                const double u = ustar[dof];
                cc.U[L][b] = u;
                acc[b][L] = cf.a0 * u + cf.a1 * un[dof] + cf.a2 * unm1[dof];
            }
        }
        cell_context<D>(cc);
        for (int b = 0; b < D; ++b) {
            double Fe;
            velocity_row<D>(cc, I, b, acc[b], 0.0, inv_k, cf, fm, true, false, nullptr, Fe);
            // pressure: - c_p int div(w) p^n
#pragma unroll
            for (int M = 0; M <= D; ++M) {
                double s = 0.0;
#pragma unroll
                for (int j = 0; j < D; ++j) s += cc.g.K[j][b] * TB::DP(I, j, M);
                Fe -= cf.c_p * cc.g.detJ * pn[vtx[M]] * s;
            }
            F[b] += Fe;
        }
    }
#pragma unroll
    for (int b = 0; b < D; ++b) b_out[(int64_t)D * row + b] = F[b];
}

// ===========================================================================================
// K3  P1 stiffness (pressure Poisson lhs):  L_MN = (detJ/d!) grad lam_M . grad lam_N
// K5' P1 mass:                              detJ M1_MN      (mode = 1)
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_p1_matrix(int32_t nrows, int mode, const int32_t* __restrict__ rowptr,
            const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
            const uint8_t* __restrict__ pair_iloc, const uint8_t* __restrict__ pair_slots,
            const double* __restrict__ coords, const int32_t* __restrict__ cell_verts,
            double* __restrict__ vals) {
    using TB = Tab<D>;
    constexpr int NV = D + 1;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    const int32_t base = rowptr[row];
    for (int32_t s = base; s < rowptr[row + 1]; ++s) vals[s] = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int I = pair_iloc[p];
        int32_t v[NV];
#pragma unroll
        for (int M = 0; M < NV; ++M) v[M] = cell_verts[(int64_t)c * NV + M];
        Geo<D> g;
        cell_geometry(coords, v, g);
        for (int N = 0; N < NV; ++N) {
            double e;
            if (mode == 0) {
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < D; ++i) s += grad_lambda<D>(g, I, i) * grad_lambda<D>(g, N, i);
                e = s * g.detJ * TB::INV_FACT;
            } else {
                e = g.detJ * TB::M1(I, N);
            }
            vals[base + pair_slots[(int64_t)p * NV + N]] += e;
        }
    }
}

// lumped P1 weights w_v = int psi_v dx  (functionals int p dx = w.p, int 1 dx = sum w)
template <int D>
__global__ void __launch_bounds__(128)
k_p1_weights(int32_t nrows, const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
             const uint8_t* __restrict__ pair_iloc, const double* __restrict__ coords,
             const int32_t* __restrict__ cell_verts, double* __restrict__ w) {
    using TB = Tab<D>;
    constexpr int NV = D + 1;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    double s = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        int32_t v[NV];
#pragma unroll
        for (int M = 0; M < NV; ++M) v[M] = cell_verts[(int64_t)c * NV + M];
        Geo<D> g;
        cell_geometry(coords, v, g);
        s += g.detJ * TB::m1(pair_iloc[p]);
    }
    w[row] = s;
}

// ===========================================================================================
// K4  Poisson rhs:  b_M = (detJ/d!) grad p^n . grad lam_M - (a0/k) detJ sum_Lb U_Lb sum_j K[j][b] DP_LjM
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_poisson_rhs(int32_t nrows, const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
              const uint8_t* __restrict__ pair_iloc, const double* __restrict__ coords,
              const int32_t* __restrict__ cell_nodes, const int32_t* __restrict__ cell_verts,
              const double* __restrict__ pn,
              const double* __restrict__ ustar, double a0_over_k, double* __restrict__ b_out) {
    using TB = Tab<D>;
    constexpr int NL = TB::NL2, NV = D + 1;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    double acc = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int Mr = pair_iloc[p];
        int32_t nodes[NL];
#pragma unroll
        for (int L = 0; L < NL; ++L) nodes[L] = cell_nodes[(int64_t)c * NL + L];
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
        Geo<D> g;
        cell_geometry(coords, vtx, g);
        double e = 0.0;
#pragma unroll
        for (int i = 0; i < D; ++i) {
            double gp = 0.0;
#pragma unroll
            for (int M = 0; M < NV; ++M) gp += pn[vtx[M]] * grad_lambda<D>(g, M, i);
            e += gp * grad_lambda<D>(g, Mr, i);
        }
        e *= g.detJ * TB::INV_FACT;
        double dv = 0.0;
        for (int L = 0; L < NL; ++L)
#pragma unroll
            for (int b = 0; b < D; ++b) {
                double s = 0.0;
#pragma unroll
                for (int j = 0; j < D; ++j) s += g.K[j][b] * TB::DP(L, j, Mr);
                dv += ustar[(int64_t)D * nodes[L] + b] * s;
            }
        acc += e - a0_over_k * g.detJ * dv;
    }
    b_out[row] = acc;
}

// ===========================================================================================
// K5  scalar P2 mass on the P2xP2 pattern:  detJ M2_IJ
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_p2_mass(int32_t nrows, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ pair_ptr,
          const int32_t* __restrict__ pair_cell, const uint8_t* __restrict__ pair_iloc,
          const uint8_t* __restrict__ pair_slots, const double* __restrict__ coords,
          const int32_t* __restrict__ cell_nodes, const int32_t* __restrict__ cell_verts,
          double* __restrict__ vals) {
    using TB = Tab<D>;
    constexpr int NL = TB::NL2;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    const int32_t base = rowptr[row];
    for (int32_t s = base; s < rowptr[row + 1]; ++s) vals[s] = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int I = pair_iloc[p];
        int32_t v[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) v[M] = cell_verts[(int64_t)c * (D + 1) + M];
        Geo<D> g;
        cell_geometry(coords, v, g);
        for (int J = 0; J < NL; ++J)
            vals[base + pair_slots[(int64_t)p * NL + J]] += g.detJ * TB::M2(I, J);
    }
}

// ===========================================================================================
// K6  correction rhs:  F_Ib = detJ sum_L M2_IL U_Lb - (k/a0) detJ m2_I (grad(p - p_old))_b
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_correction_rhs(int32_t nrows, const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
                 const uint8_t* __restrict__ pair_iloc, const double* __restrict__ coords,
                 const int32_t* __restrict__ cell_nodes, const int32_t* __restrict__ cell_verts,
                 const double* __restrict__ ustar, const double* __restrict__ pnew, const double* __restrict__ pold, double k_over_a0,
                 double* __restrict__ b_out) {
    using TB = Tab<D>;
    constexpr int NL = TB::NL2, NV = D + 1;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    double F[D];
#pragma unroll
    for (int b = 0; b < D; ++b) F[b] = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int I = pair_iloc[p];
        int32_t nodes[NL];
#pragma unroll
        for (int L = 0; L < NL; ++L) nodes[L] = cell_nodes[(int64_t)c * NL + L];
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
        Geo<D> g;
        cell_geometry(coords, vtx, g);
        double Fe[D];
#pragma unroll
        for (int b = 0; b < D; ++b) {
            double gp = 0.0;
#pragma unroll
            for (int M = 0; M < NV; ++M) gp += (pnew[vtx[M]] - pold[vtx[M]]) * grad_lambda<D>(g, M, b);
            Fe[b] = -k_over_a0 * TB::m2(I) * gp;
        }
        for (int L = 0; L < NL; ++L) {
            const double m = TB::M2(I, L);
#pragma unroll
            for (int b = 0; b < D; ++b) Fe[b] += m * ustar[(int64_t)D * nodes[L] + b];
        }
#pragma unroll
        for (int b = 0; b < D; ++b) F[b] += g.detJ * Fe[b];
    }
#pragma unroll
    for (int b = 0; b < D; ++b) b_out[(int64_t)D * row + b] = F[b];
}

// ===========================================================================================
// K7/K8  mixed Taylor-Hood residual / Jacobian rows on W (one thread per mixed dof).
//   local order [u_0: NL2 | u_1 | (u_2) | p: D+1]  
//   velocity row (b,I):  momentum residual as K1 (BDF term optional) and Jacobian as K2,
//                        B[(b,I),M] = -c_p detJ sum_j K[j][b] DP_IjM
//   pressure row M:      Fp_M = -c_p detJ sum_Lb U_Lb sum_j K[j][b] DP_LjM ; row = B^T
//   symbc: dolfin SystemAssembler symmetric local elimination (App. A.5): per cell, BC column j
//          with v_j = x_j - g_j: b_e -= A_e[:,j] v_j; zero row/col j; A_e[j,j]=1; b_e[j]=v_j.
// ===========================================================================================
template <int D>
__global__ void __launch_bounds__(128)
k_mixed_rows(int32_t nrows, int do_vec, int do_mat, int newton, int bdf, int symbc,
             const int32_t* __restrict__ rowptr, const int32_t* __restrict__ pair_ptr,
             const int32_t* __restrict__ pair_cell, const uint8_t* __restrict__ pair_iloc,
             const uint8_t* __restrict__ pair_slots, const double* __restrict__ coords,
             const int32_t* __restrict__ cell_verts, const int32_t* __restrict__ cell_dofs,
             const double* __restrict__ w0, const double* __restrict__ w1, const double* __restrict__ w2,
             const uint8_t* __restrict__ bc_mask, const double* __restrict__ bc_g, nsb_coef cf, nsb_forms fm,
             double* __restrict__ b_out, double* __restrict__ vals) {
    using TB = Tab<D>;
    constexpr int NL = TB::NL2, NV = D + 1, NW = D * NL + NV;
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    const int32_t base = rowptr[row];
    if (do_mat)
        for (int32_t s = base; s < rowptr[row + 1]; ++s) vals[s] = 0.0;
    double Facc = 0.0;
    const double inv_k = bdf ? 1.0 / cf.k : 0.0;
    const double mass_coef = bdf ? cf.a0 / cf.k : 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const int il = pair_iloc[p];
        const int32_t* dofs = cell_dofs + (int64_t)c * NW;
        int32_t v[NV];
#pragma unroll
        for (int M = 0; M < NV; ++M) v[M] = cell_verts[(int64_t)c * NV + M];
        Geo<D> g;
        cell_geometry(coords, v, g);
        CellCtx<D> cc;
        cc.g = g;
        double (&U)[NL][D] = cc.U;
#pragma unroll
        for (int L = 0; L < NL; ++L)
#pragma unroll
            for (int b = 0; b < D; ++b) U[L][b] = w0[dofs[b * NL + L]];
        double Ae[NW];   // element-matrix row
        double Fe = 0.0;
        if (il < D * NL) {
            const int br = il / NL, I = il % NL;
            cell_context<D>(cc);
            double acc[NL];
            if (bdf) {
                for (int J = 0; J < NL; ++J) {
                    const int32_t dof = dofs[br * NL + J];
                    acc[J] = cf.a0 * U[J][br] + cf.a1 * w1[dof] + cf.a2 * w2[dof];
                }
            }
            velocity_row<D>(cc, I, br, bdf ? acc : nullptr, mass_coef, inv_k, cf, fm, newton != 0, do_mat || symbc, Ae, Fe);
#pragma unroll
            for (int M = 0; M < NV; ++M) {
                double s = 0.0;
#pragma unroll
                for (int j = 0; j < D; ++j) s += g.K[j][br] * TB::DP(I, j, M);
                const double bb = -cf.c_p * g.detJ * s;
                Ae[D * NL + M] = bb;
                Fe += bb * w0[dofs[D * NL + M]];
            }
        } else {
            const int Mr = il - D * NL;
            for (int L = 0; L < NL; ++L)
#pragma unroll
                for (int b = 0; b < D; ++b) {
                    double s = 0.0;
#pragma unroll
                    for (int j = 0; j < D; ++j) s += g.K[j][b] * TB::DP(L, j, Mr);
                    const double bb = -cf.c_p * g.detJ * s;
                    Ae[b * NL + L] = bb;
                    Fe += bb * U[L][b];
                }
#pragma unroll
            for (int M = 0; M < NV; ++M) Ae[D * NL + M] = 0.0;
        }
        if (symbc) {
            if (bc_mask[row]) {
                for (int j = 0; j < NW; ++j) Ae[j] = 0.0;
                Ae[il] = 1.0;
                Fe = w0[row] - bc_g[row];
            } else {
                for (int j = 0; j < NW; ++j) {
                    const int32_t dj = dofs[j];
                    if (bc_mask[dj]) {
                        Fe -= Ae[j] * (w0[dj] - bc_g[dj]);
                        Ae[j] = 0.0;
                    }
                }
            }
        }
        Facc += Fe;
        if (do_mat)
            for (int j = 0; j < NW; ++j) vals[base + pair_slots[(int64_t)p * NW + j]] += Ae[j];
    }
    if (do_vec) b_out[row] = Facc;
}

// ===========================================================================================
// C-ABI
// ===========================================================================================
#define NSB_DIM_DISPATCH(ctx, expr2, expr3)                                   \
    do {                                                                      \
        if ((ctx)->dim == 2) { expr2; } else if ((ctx)->dim == 3) { expr3; }  \
        else return nsb_fail(ctx, "mesh not set%s%s");                        \
    } while (0)

extern "C" int nsb_assemble_ipcs_momentum(nsb_ctx* ctx, int32_t flags, const double* ustar,
                                          const double* un, const double* unm1, const double* pn,
                                          double* b_out, double* A_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P2P2];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr, "P2xP2 pattern not set");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 128, B = nsb_blocks(P.nrows, T);
    const nsb_coef cf = ctx->coef;
    nsb_forms fm;
    fm.conv = ctx->form_conv; fm.visc = ctx->form_visc;
    for (int i = 0; i < 3; ++i) fm.cor[i] = ctx->coriolis[i];
    // the tuned kernels cover the default weak form (standard convection, reduced viscous term, no Coriolis term);
    // every other variant of the reference runs through the generic row kernels (assemble_forms.cuh)
    const bool tuned = fm.conv == 0 && fm.visc == 0 && fm.cor[0] == 0.0 && fm.cor[1] == 0.0 && fm.cor[2] == 0.0 &&
                       !ctx->force_generic_assembly;
    if (flags & NSB_VEC) {
        NSB_REQUIRE(ctx, ustar && un && unm1 && pn && b_out, "null vector argument");
        NSB_TIMED(ctx, NSB_T_ASM_RES, st);
        const int tp = tuned ? nsb_launch_residual_cells(ctx, ustar, un, unm1, pn, b_out, st) : 1;
        if (tp < 0) return -1;
        if (tp == 1) {
        NSB_DIM_DISPATCH(ctx,
            (k_momentum_residual<2><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                                     ctx->cell_nodes, ctx->cell_verts, ustar, un, unm1, pn, cf, fm, b_out)),
            (k_momentum_residual<3><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                                     ctx->cell_nodes, ctx->cell_verts, ustar, un, unm1, pn, cf, fm, b_out)));
        NSB_LAUNCH_CHECK(ctx, "k_momentum_residual");
        }
    }
    if (flags & NSB_MAT) {
        NSB_REQUIRE(ctx, ustar && A_out, "null matrix argument");
        const int newton = (flags & NSB_PICARD) ? 0 : 1;
        const double mc = cf.a0 / cf.k;
        NSB_TIMED(ctx, NSB_T_ASM_JAC, st);
        const int rg = tuned ? nsb_launch_jacobian_rowgroups(ctx, ustar, mc, newton, A_out, ctx->jac_block_scalar_out, st) : 1;
        if (rg < 0) return -1;
        if (rg == 0) return 0;
        NSB_DIM_DISPATCH(ctx,
            (k_momentum_jacobian<2><<<B, T, 0, st>>>(P.nrows, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
                                                     P.pair_slots, ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, mc,
                                                     cf, fm, newton, A_out)),
            (k_momentum_jacobian<3><<<B, T, 0, st>>>(P.nrows, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
                                                     P.pair_slots, ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, mc,
                                                     cf, fm, newton, A_out)));
        NSB_LAUNCH_CHECK(ctx, "k_momentum_jacobian");
        // the generic kernels do not carry the fused block-scalar output: same contract through the stand-alone kernel
        if (ctx->jac_block_scalar_out &&
            nsb_bsr_block_scalar_f32(ctx, ctx->jac_block_scalar_nnz, ctx->dim, A_out, ctx->jac_block_scalar_out, stream))
            return -1;
    }
    return 0;
}

extern "C" int nsb_assemble_poisson(nsb_ctx* ctx, int32_t flags, const double* pn, const double* ustar,
                                    double* b_out, double* A_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P1P1];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr, "P1xP1 pattern not set");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 128, B = nsb_blocks(P.nrows, T);
    if (flags & NSB_MAT) {
        NSB_REQUIRE(ctx, A_out, "null matrix argument");
        NSB_TIMED(ctx, NSB_T_ASM_CONST, st);
        NSB_DIM_DISPATCH(ctx,
            (k_p1_matrix<2><<<B, T, 0, st>>>(P.nrows, 0, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
                                             P.pair_slots, ctx->coords, ctx->cell_verts, A_out)),
            (k_p1_matrix<3><<<B, T, 0, st>>>(P.nrows, 0, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
                                             P.pair_slots, ctx->coords, ctx->cell_verts, A_out)));
        NSB_LAUNCH_CHECK(ctx, "k_p1_matrix");
    }
    if (flags & NSB_VEC) {
        NSB_REQUIRE(ctx, pn && ustar && b_out, "null vector argument");
        const double aok = ctx->coef.a0 / ctx->coef.k;
        NSB_TIMED(ctx, NSB_T_ASM_RHS, st);
        NSB_DIM_DISPATCH(ctx,
            (k_poisson_rhs<2><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                               ctx->cell_nodes, ctx->cell_verts, pn, ustar, aok, b_out)),
            (k_poisson_rhs<3><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                               ctx->cell_nodes, ctx->cell_verts, pn, ustar, aok, b_out)));
        NSB_LAUNCH_CHECK(ctx, "k_poisson_rhs");
    }
    return 0;
}

extern "C" int nsb_assemble_correction(nsb_ctx* ctx, int32_t flags, const double* ustar,
                                       const double* pnew, const double* pold, double* b_out,
                                       double* A_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P2P2];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr, "P2xP2 pattern not set");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 128, B = nsb_blocks(P.nrows, T);
    if (flags & NSB_MAT) {
        NSB_REQUIRE(ctx, A_out, "null matrix argument");
        NSB_TIMED(ctx, NSB_T_ASM_CONST, st);
        NSB_DIM_DISPATCH(ctx,
            (k_p2_mass<2><<<B, T, 0, st>>>(P.nrows, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc, P.pair_slots,
                                           ctx->coords, ctx->cell_nodes, ctx->cell_verts, A_out)),
            (k_p2_mass<3><<<B, T, 0, st>>>(P.nrows, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc, P.pair_slots,
                                           ctx->coords, ctx->cell_nodes, ctx->cell_verts, A_out)));
        NSB_LAUNCH_CHECK(ctx, "k_p2_mass");
    }
    if (flags & NSB_VEC) {
        NSB_REQUIRE(ctx, ustar && pnew && pold && b_out, "null vector argument");
        const double koa = ctx->coef.k / ctx->coef.a0;
        NSB_TIMED(ctx, NSB_T_ASM_RHS, st);
        NSB_DIM_DISPATCH(ctx,
            (k_correction_rhs<2><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                                  ctx->cell_nodes, ctx->cell_verts, ustar, pnew, pold, koa, b_out)),
            (k_correction_rhs<3><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                                  ctx->cell_nodes, ctx->cell_verts, ustar, pnew, pold, koa, b_out)));
        NSB_LAUNCH_CHECK(ctx, "k_correction_rhs");
    }
    return 0;
}

extern "C" int nsb_assemble_p1_mass(nsb_ctx* ctx, int32_t flags, double* w_out, double* A_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P1P1];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr, "P1xP1 pattern not set");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 128, B = nsb_blocks(P.nrows, T);
    if (flags & NSB_MAT) {
        NSB_REQUIRE(ctx, A_out, "null matrix argument");
        NSB_TIMED(ctx, NSB_T_ASM_CONST, st);
        NSB_DIM_DISPATCH(ctx,
            (k_p1_matrix<2><<<B, T, 0, st>>>(P.nrows, 1, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
                                             P.pair_slots, ctx->coords, ctx->cell_verts, A_out)),
            (k_p1_matrix<3><<<B, T, 0, st>>>(P.nrows, 1, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
                                             P.pair_slots, ctx->coords, ctx->cell_verts, A_out)));
        NSB_LAUNCH_CHECK(ctx, "k_p1_matrix(mass)");
    }
    if (flags & NSB_VEC) {
        NSB_REQUIRE(ctx, w_out, "null vector argument");
        NSB_TIMED(ctx, NSB_T_ASM_CONST, st);
        NSB_DIM_DISPATCH(ctx,
            (k_p1_weights<2><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                              ctx->cell_verts, w_out)),
            (k_p1_weights<3><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords,
                                              ctx->cell_verts, w_out)));
        NSB_LAUNCH_CHECK(ctx, "k_p1_weights");
    }
    return 0;
}

extern "C" int nsb_assemble_mixed(nsb_ctx* ctx, int32_t flags, const double* w0, const double* w1,
                                  const double* w2, double* b_out, double* A_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_WW];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr && ctx->cell_dofs_w, "mixed pattern / dofmap not set");
    const int do_vec = (flags & NSB_VEC) ? 1 : 0, do_mat = (flags & NSB_MAT) ? 1 : 0;
    const int bdf = (flags & NSB_BDF) ? 1 : 0, symbc = (flags & NSB_SYMBC) ? 1 : 0;
    const int newton = (flags & NSB_PICARD) ? 0 : 1;
    NSB_REQUIRE(ctx, w0 && (!do_vec || b_out) && (!do_mat || A_out), "null argument");
    NSB_REQUIRE(ctx, !bdf || (w1 && w2), "BDF form needs the two history vectors");
    NSB_REQUIRE(ctx, !symbc || (ctx->w_bc_mask && ctx->w_bc_g), "symmetric BC elimination needs nsb_set_mixed_bc");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 128, B = nsb_blocks(P.nrows, T);
    nsb_forms fm;
    fm.conv = ctx->form_conv; fm.visc = ctx->form_visc;
    for (int i = 0; i < 3; ++i) fm.cor[i] = ctx->coriolis[i];
    nsb_timed_region mixed_region(ctx, do_mat ? NSB_T_ASM_JAC : NSB_T_ASM_RES, st);
    NSB_DIM_DISPATCH(ctx,
        (k_mixed_rows<2><<<B, T, 0, st>>>(P.nrows, do_vec, do_mat, newton, bdf, symbc, P.rowptr, P.pair_ptr,
                                          P.pair_cell, P.pair_iloc, P.pair_slots, ctx->coords, ctx->cell_verts,
                                          ctx->cell_dofs_w, w0, w1, w2, ctx->w_bc_mask, ctx->w_bc_g, ctx->coef, fm,
                                          b_out, A_out)),
        (k_mixed_rows<3><<<B, T, 0, st>>>(P.nrows, do_vec, do_mat, newton, bdf, symbc, P.rowptr, P.pair_ptr,
                                          P.pair_cell, P.pair_iloc, P.pair_slots, ctx->coords, ctx->cell_verts,
                                          ctx->cell_dofs_w, w0, w1, w2, ctx->w_bc_mask, ctx->w_bc_g, ctx->coef, fm,
                                          b_out, A_out)));
    NSB_LAUNCH_CHECK(ctx, "k_mixed_rows");
    return 0;
}

// weak-form variants (source/ns_solver_base.py:78-115 constructor arguments form_convective_term / form_viscous_term,
// :173-191 Coriolis term): conv 0 standard | 1 rotational | 2 divergence | 3 skew_symmetric; visc 0 reduced | 1 traction;
// coriolis3 = 2 c_Omega Omega (2D: only [2]); generic != 0 forces the generic row kernels for the default form too
extern "C" int nsb_set_forms(nsb_ctx* ctx, int32_t conv, int32_t visc, const double* coriolis3, int32_t generic) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, conv >= 0 && conv <= 3 && (visc == 0 || visc == 1), "unknown weak-form variant");
    ctx->form_conv = conv;
    ctx->form_visc = visc;
    for (int i = 0; i < 3; ++i) ctx->coriolis[i] = coriolis3 ? coriolis3[i] : 0.0;
    ctx->force_generic_assembly = generic != 0;
    return 0;
}
