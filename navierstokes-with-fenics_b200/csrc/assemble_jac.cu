// assemble_jac.cu -- K2, the IPCS tentative-velocity Jacobian (source/ns_ipcs_solver.py:136,
// J = derivative(F, u*)), as a warp-cooperative, atomic-free, write-once row-gather.
//
//   A[(I,b),(J,a)] = delta_ab S_IJ + c_c detJ sum_M T_IJM  d_a u_b(x_M)
//   S_IJ = (a0/k) detJ M2_IJ + c_nu sum_jk G_jk D2_IJjk + c_c sum_M y_IM . DV_JM
//   y_IM[k] = sum_L T_ILM z_L[k],  z_L[k] = detJ sum_i K[k][i] U_L[i]        
//
// Work unit = a "row group": consecutive rows (scalar P2 nodes) with at most 32 (row, cell)
// incidence pairs (nsb_row_groups_build).  One warp per group:
//   phase A  lane = pair: gather the cell, geometry, u*; evaluate in closed form the pair's
//            scalar row S_I. (NL values) and the vertex velocity gradients (the Newton term),
//            and park them in the warp's shared staging buffer.  Loops over the column index J
//            and the vertex index M are fully unrolled against constexpr tables, so the zeros
//            of the reference tensors cost nothing; only the row index I is a run-time index
//            (tables in shared memory, no constant-cache serialisation).
//   phase B  lane = (J, b): walk the pairs in order (ascending cell order inside a row = the
//            summation order of dolfin's cell loop), form the d entries of block row b of
//            block J and add them into the shared copy of the current matrix row; when the row
//            changes the finished row is written out with fully coalesced stores.
// Every value is written to global memory exactly once, by one warp, in a fixed order:
// bit-reproducible, no atomics, no colouring, no read-modify-write traffic on HBM.
#define NSB_TABLE static __device__ constexpr
#include "tables.h"
#include "common.cuh"
#include "fe_geometry.cuh"

namespace {

template <int D> struct CT;
template <> struct CT<2> {
    static constexpr int NL = 6;
    static __device__ constexpr double M2(int i, int j) { return M2_2D[i][j]; }
    static __device__ constexpr double D2(int i, int j, int a, int b) { return D2_2D[i][j][a][b]; }
    static __device__ constexpr double T(int i, int j, int m) { return T_2D[i][j][m]; }
    static __device__ constexpr double DV(int j, int m, int k) { return DV_2D[j][m][k]; }
};
template <> struct CT<3> {
    static constexpr int NL = 10;
    static __device__ constexpr double M2(int i, int j) { return M2_3D[i][j]; }
    static __device__ constexpr double D2(int i, int j, int a, int b) { return D2_3D[i][j][a][b]; }
    static __device__ constexpr double T(int i, int j, int m) { return T_3D[i][j][m]; }
    static __device__ constexpr double DV(int j, int m, int k) { return DV_3D[j][m][k]; }
};

template <int D>
struct JacCfg {
    static constexpr int NL = CT<D>::NL, NV = D + 1, DD = D * D, NS = D * (D + 1) / 2;
    static constexpr int WARPS = 4;
    static constexpr int MAXROW = (D == 3) ? 96 : 64;            // blocks in one row handled in shared memory
    static constexpr int STAGE = NL + NV * DD + ((NL + NV * DD) % 2 == 0 ? 1 : 0);   // odd stride: conflict-free
    // shared tables indexed by the run-time row index I
    static constexpr int TAB_T = NL * NL * NV, TAB_M = NL * NL, TAB_D = NL * NL * NS;
    static constexpr int TAB = TAB_T + TAB_M + TAB_D;
    static constexpr int PER_WARP = MAXROW * DD + 32 * STAGE;    // doubles
    static constexpr size_t SMEM = sizeof(double) * (TAB + WARPS * PER_WARP) + WARPS * 32 * (NL + 8);
};

// symmetric index of (j,k), j <= k
__device__ constexpr int sym_index(int D, int j, int k) { return j * D - j * (j - 1) / 2 + (k - j); }

}  // namespace

template <int D>
__global__ void __launch_bounds__(JacCfg<D>::WARPS * 32)
k_jacobian_rowgroups(int32_t ngroups, const int32_t* __restrict__ group_row, const int32_t* __restrict__ rowptr,
                     const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_row,
                     const int32_t* __restrict__ pair_cell, const uint8_t* __restrict__ pair_iloc,
                     const uint8_t* __restrict__ pair_slots, const double* __restrict__ coords,
                     const int32_t* __restrict__ cell_nodes, const double* __restrict__ ustar, double mass_coef,
                     double c_conv, double c_visc, int newton, double* __restrict__ vals) {
    using C = JacCfg<D>;
    using TB = CT<D>;
    constexpr int NL = C::NL, NV = C::NV, DD = C::DD, NS = C::NS, STAGE = C::STAGE;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sT = reinterpret_cast<double*>(smem_raw);
    double* sM2 = sT + C::TAB_T;
    double* sD2 = sM2 + C::TAB_M;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double* rowbuf = sD2 + C::TAB_D + (size_t)w * C::PER_WARP;
    double* stage = rowbuf + C::MAXROW * DD;
    uint8_t* sslot = reinterpret_cast<uint8_t*>(sT + C::TAB + C::WARPS * C::PER_WARP) + w * 32 * (NL + 8);

    // tables indexed by the run-time row index I -> shared memory
    for (int e = threadIdx.x; e < NL * NL; e += blockDim.x) {
        const int i = e / NL, j = e % NL;
        sM2[e] = TB::M2(i, j);
        for (int m = 0; m < NV; ++m) sT[e * NV + m] = TB::T(i, j, m);
        for (int a = 0; a < D; ++a)
            for (int b = a; b < D; ++b)
                sD2[e * NS + sym_index(D, a, b)] = TB::D2(i, j, a, b) + (a < b ? TB::D2(i, j, b, a) : 0.0);
    }
    __syncthreads();

    for (int32_t grp = blockIdx.x * C::WARPS + w; grp < ngroups; grp += gridDim.x * C::WARPS) {
        const int32_t row_lo = group_row[grp], row_hi = group_row[grp + 1];
        const int32_t p_lo = pair_ptr[row_lo], p_hi = pair_ptr[row_hi];
        int32_t cur = -1, cur_rp0 = 0, cur_len = 0;
        for (int32_t chunk = p_lo; chunk < p_hi; chunk += 32) {
            const int npc = min(32, p_hi - chunk);
            const bool active = lane < npc;
            const int32_t P = chunk + lane;
            int32_t my_row = -1, my_rp0 = 0, my_len = 0;
            int I = 0;
            // ---------------------------------------------------------------- phase A: lane = pair
            if (active) {
                my_row = pair_row[P];
                my_rp0 = rowptr[my_row];
                my_len = rowptr[my_row + 1] - my_rp0;
                const int32_t c = pair_cell[P];
                I = pair_iloc[P];
                int32_t nodes[NL];
#pragma unroll
                for (int L = 0; L < NL; ++L) nodes[L] = cell_nodes[(int64_t)c * NL + L];
#pragma unroll
                for (int J = 0; J < NL; ++J) sslot[lane * (NL + 8) + J] = pair_slots[(int64_t)P * NL + J];
                Geo<D> g;
                cell_geometry(coords, nodes, g);
                double U[NL][D];
#pragma unroll
                for (int L = 0; L < NL; ++L)
#pragma unroll
                    for (int b = 0; b < D; ++b) U[L][b] = ustar[(int64_t)D * nodes[L] + b];
                double* st = stage + lane * STAGE;
                if (newton) {
                    const double f = c_conv * g.detJ;
#pragma unroll
                    for (int M = 0; M < NV; ++M) {
                        double ref[D][D];   // ref[b][k] = sum_L U_Lb dhat_k phi_L(x_M)
#pragma unroll
                        for (int b = 0; b < D; ++b)
#pragma unroll
                            for (int k = 0; k < D; ++k) ref[b][k] = 0.0;
#pragma unroll
                        for (int L = 0; L < NL; ++L)
#pragma unroll
                            for (int k = 0; k < D; ++k)
                                if (TB::DV(L, M, k) != 0.0) {
#pragma unroll
                                    for (int b = 0; b < D; ++b) ref[b][k] += U[L][b] * TB::DV(L, M, k);
                                }
#pragma unroll
                        for (int b = 0; b < D; ++b)
#pragma unroll
                            for (int a = 0; a < D; ++a) {
                                double s = 0.0;
#pragma unroll
                                for (int k = 0; k < D; ++k) s += ref[b][k] * g.K[k][a];
                                st[NL + M * DD + b * D + a] = f * s;
                            }
                    }
                }
                // y[M][k] = sum_L T(I,L,M) z_L[k]
                double y[NV][D];
#pragma unroll
                for (int M = 0; M < NV; ++M)
#pragma unroll
                    for (int k = 0; k < D; ++k) y[M][k] = 0.0;
                const double* tI = sT + I * NL * NV;
#pragma unroll
                for (int L = 0; L < NL; ++L) {
                    double z[D];
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        double s = 0.0;
#pragma unroll
                        for (int i = 0; i < D; ++i) s += g.K[k][i] * U[L][i];
                        z[k] = s * g.detJ;
                    }
#pragma unroll
                    for (int M = 0; M < NV; ++M) {
                        const double t = tI[L * NV + M];
#pragma unroll
                        for (int k = 0; k < D; ++k) y[M][k] += t * z[k];
                    }
                }
                double G[D][D];
                metric<D>(g, G);
                const double mcd = mass_coef * g.detJ;
                const double* mI = sM2 + I * NL;
                const double* dI = sD2 + I * NL * NS;
#pragma unroll
                for (int J = 0; J < NL; ++J) {
                    double vis = 0.0;
#pragma unroll
                    for (int a = 0; a < D; ++a)
#pragma unroll
                        for (int b = a; b < D; ++b) vis += G[a][b] * dI[J * NS + sym_index(D, a, b)];
                    double cv = 0.0;
#pragma unroll
                    for (int M = 0; M < NV; ++M)
#pragma unroll
                        for (int k = 0; k < D; ++k)
                            if (TB::DV(J, M, k) != 0.0) cv += y[M][k] * TB::DV(J, M, k);
                    st[J] = mcd * mI[J] + c_visc * vis + c_conv * cv;
                }
            }
            __syncwarp();
            // ---------------------------------------------------------------- phase B: lane = (J, b)
            const int J = lane / D, b = lane - J * D;
            for (int q = 0; q < npc; ++q) {
                const int32_t rq = __shfl_sync(FULL, my_row, q);
                if (rq != cur) {
                    if (cur >= 0) {      // finished row: coalesced write-out, exactly once
                        double* dst = vals + (int64_t)cur_rp0 * DD;
                        for (int e = lane; e < cur_len * DD; e += 32) dst[e] = rowbuf[e];
                    }
                    cur = rq;
                    cur_rp0 = __shfl_sync(FULL, my_rp0, q);
                    cur_len = __shfl_sync(FULL, my_len, q);
                    for (int e = lane; e < cur_len * DD; e += 32) rowbuf[e] = 0.0;
                    __syncwarp();
                }
                const int Iq = __shfl_sync(FULL, I, q);
                if (lane < NL * D) {
                    const double* st = stage + q * STAGE;
                    double v[D];
                    const double S = st[J];
#pragma unroll
                    for (int a = 0; a < D; ++a) v[a] = (a == b) ? S : 0.0;
                    if (newton) {
                        const double* t = sT + (Iq * NL + J) * NV;
#pragma unroll
                        for (int M = 0; M < NV; ++M) {
                            const double tm = t[M];
#pragma unroll
                            for (int a = 0; a < D; ++a) v[a] += tm * st[NL + M * DD + b * D + a];
                        }
                    }
                    double* dst = rowbuf + (int)sslot[q * (NL + 8) + J] * DD + b * D;
#pragma unroll
                    for (int a = 0; a < D; ++a) dst[a] += v[a];
                }
                __syncwarp();
            }
        }
        if (cur >= 0) {
            double* dst = vals + (int64_t)cur_rp0 * DD;
            for (int e = lane; e < cur_len * DD; e += 32) dst[e] = rowbuf[e];
        }
        __syncwarp();
    }
}

// launcher used by nsb_assemble_ipcs_momentum (assemble.cu); returns 1 when the row-group path
// does not apply (groups not registered / a row longer than the shared row buffer) so that the
// caller keeps the generic one-thread-per-row kernel.
template <int D>
static int launch_rowgroups(nsb_ctx* ctx, const nsb_pattern_dev& P, const double* ustar, double mass_coef,
                            int newton, double* A_out, cudaStream_t st) {
    using C = JacCfg<D>;
    if (!P.group_row || !P.pair_row || P.ngroups <= 0 || P.max_row_len > C::MAXROW) return 1;
    static bool configured = false;
    if (!configured) {
        NSB_CUDA(ctx, cudaFuncSetAttribute(k_jacobian_rowgroups<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)C::SMEM));
        NSB_CUDA(ctx, cudaFuncSetAttribute(k_jacobian_rowgroups<D>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        configured = true;
    }
    int per_sm = 0;
    NSB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_jacobian_rowgroups<D>, C::WARPS * 32, C::SMEM));
    if (per_sm < 1) return 1;
    int64_t blocks = (int64_t)ctx->sm_count * per_sm;
    const int64_t need = ((int64_t)P.ngroups + C::WARPS - 1) / C::WARPS;
    if (blocks > need) blocks = need;
    k_jacobian_rowgroups<D><<<(int)blocks, C::WARPS * 32, C::SMEM, st>>>(
        P.ngroups, P.group_row, P.rowptr, P.pair_ptr, P.pair_row, P.pair_cell, P.pair_iloc, P.pair_slots, ctx->coords,
        ctx->cell_nodes, ustar, mass_coef, ctx->coef.c_conv, ctx->coef.c_visc, newton, A_out);
    NSB_LAUNCH_CHECK(ctx, "k_jacobian_rowgroups");
    return 0;
}

int nsb_launch_jacobian_rowgroups(nsb_ctx* ctx, const double* ustar, double mass_coef, int newton, double* A_out,
                                  cudaStream_t st) {
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P2P2];
    if (ctx->dim == 2) return launch_rowgroups<2>(ctx, P, ustar, mass_coef, newton, A_out, st);
    if (ctx->dim == 3) return launch_rowgroups<3>(ctx, P, ustar, mass_coef, newton, A_out, st);
    return 1;
}
