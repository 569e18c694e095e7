// assemble_jac.cu -- K2, the IPCS tentative-velocity Jacobian (source/ns_ipcs_solver.py:136,
// J = derivative(F, u*)), as a warp-cooperative, atomic-free, write-once row-gather.
//
//   A[(I,b),(J,a)] = delta_ab S_IJ + c_c detJ sum_M T_IJM  d_a u_b(x_M)
//   S_IJ = (a0/k) detJ M2_IJ + c_nu sum_jk G_jk D2_IJjk + c_c sum_M y_IM . DV_JM
//   y_IM[k] = sum_L T_ILM z_L[k],  z_L[k] = detJ sum_i K[k][i] U_L[i]        
//
// Work unit = a "row group": consecutive rows (scalar P2 nodes) with at most 32 (row, cell)
// incidence pairs and MAXB entries (nsb_sched_build).  One warp per group, lane = pair:
//   phase A  gather the cell, geometry, u*; evaluate in closed form the pair's scalar row S_I.
//            (NL values, parked in shared memory) and the scaled vertex velocity gradients
//            g[b][a][M] = c_c detJ d_a u_b(x_M) (kept in REGISTERS).  Loops over the column index
//            J and the vertex index M are fully unrolled against constexpr tables, so the zeros
//            of the reference tensors cost nothing; only the row index I is a run-time index
//            (tables in shared memory: no constant-cache serialisation).
//   phase B  scheduled scatter: in round t the lane forms ONE d x d block (local column
//            J = sched[t][lane]) as delta_ab S_J + sum_M T_IJM g[b][a][M] and adds it to the
//            group's rows held in shared memory.  The host-built schedule is a proper edge
//            colouring of the pair/entry graph: within a round all lanes hit distinct entries,
//            so the shared-memory read-modify-writes need no atomics and the summation order
//            is fixed; it prefers rounds without bank conflicts, and entries with many contributions
//            (diagonal of a vertex row) are spread over shadow slots that are folded at the end.
//            Finally the group's rows -- one contiguous range of the BSR value array -- are written
//            with fully coalesced stores.
// Every value is written to global memory exactly once, by one warp, in a fixed order:
// bit-reproducible, no atomics, no colouring of cells, no read-modify-write traffic on HBM.
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
    static constexpr int MAXB = 160;                              // entries of one row group held in shared memory
    static constexpr int SSTR = NL + 1 - (NL & 1);                // odd stride of the per-lane S rows: conflict-free
    // shared tables indexed by the run-time row index I (T padded to 4 columns for 128-bit loads)
    static constexpr int TAB_T = NL * NL * 4, TAB_M = NL * NL, TAB_D = NL * NL * NS;
    static constexpr int TAB = TAB_T + TAB_M + TAB_D;
    static constexpr int SROUNDS = 32;                            // rounds of one chunk staged in shared memory (64 B each)
    static constexpr int PER_WARP = MAXB * DD + 32 * SSTR + ((MAXB * DD + 32 * SSTR) & 1) + SROUNDS * 8;   // doubles
    static constexpr size_t SMEM = sizeof(double) * (TAB + WARPS * PER_WARP);
};

// symmetric index of (j,k), j <= k
__device__ constexpr int sym_index(int D, int j, int k) { return j * D - j * (j - 1) / 2 + (k - j); }

}  // namespace

template <int D>
__global__ void __launch_bounds__(JacCfg<D>::WARPS * 32, 3)
k_jacobian_rowgroups(int32_t ngroups, const int32_t* __restrict__ group_row, const int32_t* __restrict__ chunk_first,
                     const int32_t* __restrict__ round_ptr, const uint16_t* __restrict__ sched,
                     const int32_t* __restrict__ fold_ptr, const uint32_t* __restrict__ fold,
                     const int32_t* __restrict__ rowptr, const int32_t* __restrict__ pair_ptr,
                     const int32_t* __restrict__ pair_cell, const uint8_t* __restrict__ pair_iloc,
                     const double* __restrict__ coords, const int32_t* __restrict__ cell_nodes,
                     const int32_t* __restrict__ cell_verts, const double* __restrict__ ustar, double mass_coef, double c_conv, double c_visc, int newton,
                     double* __restrict__ vals, float* __restrict__ scal32) {
    using C = JacCfg<D>;
    using TB = CT<D>;
    constexpr int NL = C::NL, NV = C::NV, DD = C::DD, NS = C::NS, SSTR = C::SSTR;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double* sT = reinterpret_cast<double*>(smem_raw);
    double* sM2 = sT + C::TAB_T;
    double* sD2 = sM2 + C::TAB_M;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    double* rowbuf = sD2 + C::TAB_D + (size_t)w * C::PER_WARP;
    double* sS = rowbuf + C::MAXB * DD;
    // the chunk's schedule, copied asynchronously while phase A runs (16-byte aligned: every part above is even)
    uint16_t* ssched = reinterpret_cast<uint16_t*>(rowbuf + C::PER_WARP - C::SROUNDS * 8);

    // tables indexed by the run-time row index I -> shared memory
    for (int e = threadIdx.x; e < NL * NL; e += blockDim.x) {
        const int i = e / NL, j = e % NL;
        sM2[e] = TB::M2(i, j);
        for (int m = 0; m < 4; ++m) sT[e * 4 + m] = m < NV ? TB::T(i, j, m) : 0.0;
        for (int a = 0; a < D; ++a)
            for (int b = a; b < D; ++b)
                sD2[e * NS + sym_index(D, a, b)] = TB::D2(i, j, a, b) + (a < b ? TB::D2(i, j, b, a) : 0.0);
    }
    __syncthreads();

    for (int32_t grp = blockIdx.x * C::WARPS + w; grp < ngroups; grp += gridDim.x * C::WARPS) {
        const int32_t row_lo = group_row[grp], row_hi = group_row[grp + 1];
        const int32_t e_lo = rowptr[row_lo], nval = (rowptr[row_hi] - e_lo) * DD;
        const int32_t p_lo = pair_ptr[row_lo], p_hi = pair_ptr[row_hi];
        const int32_t c_lo = chunk_first[grp], c_hi = chunk_first[grp + 1];
        const int32_t f_lo = fold_ptr[grp], f_hi = fold_ptr[grp + 1];
        for (int e = lane; e < nval + (f_hi - f_lo) * DD; e += 32) rowbuf[e] = 0.0;   // real entries, then shadow slots
        for (int32_t ch = c_lo; ch < c_hi; ++ch) {
            const int32_t q0 = p_lo + 32 * (ch - c_lo);
            const int npc = min(32, p_hi - q0);
            const int32_t r_lo = round_ptr[ch], r_hi = round_ptr[ch + 1];
            const int nstaged = min(r_hi - r_lo, C::SROUNDS);
            {   // rounds x 64 B -> 16-byte granules, lane-strided: coalesced, no register staging
                const uint16_t* src = sched + (int64_t)r_lo * 32;
                for (int gidx = lane; gidx < nstaged * 4; gidx += 32) {
                    const unsigned dst = (unsigned)__cvta_generic_to_shared(ssched + gidx * 8);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src + gidx * 8) : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
            }
            int I = 0;
            double g[DD][4];      // g[b*D+a][M] = c_c detJ d_a u_b (x_M), registers
#pragma unroll
            for (int i = 0; i < DD; ++i)
#pragma unroll
                for (int M = 0; M < 4; ++M) g[i][M] = 0.0;
            // ---------------------------------------------------------------- phase A: lane = pair
            if (lane < npc) {
                const int32_t P = q0 + lane;
                const int32_t c = pair_cell[P];
                I = pair_iloc[P];
                int32_t nodes[NL];
#pragma unroll
                for (int L = 0; L < NL; ++L) nodes[L] = cell_nodes[(int64_t)c * NL + L];
                int32_t vtx[D + 1];
#pragma unroll
                for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
                Geo<D> geo;
                cell_geometry(coords, vtx, geo);
                double U[NL][D];
#pragma unroll
                for (int L = 0; L < NL; ++L)
#pragma unroll
                    for (int b = 0; b < D; ++b) U[L][b] = ustar[(int64_t)D * nodes[L] + b];
                if (newton) {
                    const double f = c_conv * geo.detJ;
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
                                for (int k = 0; k < D; ++k) s += ref[b][k] * geo.K[k][a];
                                g[b * D + a][M] = f * s;
                            }
                    }
                }
                // y[M][k] = sum_L T(I,L,M) z_L[k]
                double y[NV][D];
#pragma unroll
                for (int M = 0; M < NV; ++M)
#pragma unroll
                    for (int k = 0; k < D; ++k) y[M][k] = 0.0;
                const double* tI = sT + I * NL * 4;
#pragma unroll
                for (int L = 0; L < NL; ++L) {
                    double z[D];
#pragma unroll
                    for (int k = 0; k < D; ++k) {
                        double s = 0.0;
#pragma unroll
                        for (int i = 0; i < D; ++i) s += geo.K[k][i] * U[L][i];
                        z[k] = s * geo.detJ;
                    }
#pragma unroll
                    for (int M = 0; M < NV; ++M) {
                        const double t = tI[L * 4 + M];
#pragma unroll
                        for (int k = 0; k < D; ++k) y[M][k] += t * z[k];
                    }
                }
                double G[D][D];
                metric<D>(geo, G);
                const double mcd = mass_coef * geo.detJ;
                const double* mI = sM2 + I * NL;
                const double* dI = sD2 + I * NL * NS;
                double* st = sS + lane * SSTR;
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
            asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            // ---------------------------------------------------------------- phase B: scheduled scatter
            const double* myS = sS + lane * SSTR;
            const double* myT = sT + I * NL * 4;
            for (int32_t t = r_lo; t < r_hi; ++t) {
                const unsigned e = t - r_lo < nstaged ? ssched[(t - r_lo) * 32 + lane] : sched[(int64_t)t * 32 + lane];
                if (e != 0xFFFFu) {
                    const int J = e & 15u;
                    double* dst = rowbuf + (e >> 4) * DD;
                    const double S = myS[J];
                    const double2 t01 = *reinterpret_cast<const double2*>(myT + J * 4);
                    const double2 t23 = *reinterpret_cast<const double2*>(myT + J * 4 + 2);
#pragma unroll
                    for (int b = 0; b < D; ++b)
#pragma unroll
                        for (int a = 0; a < D; ++a) {
                            const double* gi = g[b * D + a];
                            double v = (t01.x * gi[0] + t01.y * gi[1]) + (t23.x * gi[2] + t23.y * gi[3]);
                            if (a == b) v += S;
                            dst[b * D + a] += v;
                        }
                }
                __syncwarp();
            }
        }
        // shadow copies of the hot entries -> their real entry, in list order (copies of one entry are consecutive)
        for (int32_t f = f_lo; f < f_hi; ++f) {
            const uint32_t it = fold[f];
            if (lane < DD) rowbuf[(it >> 16) * DD + lane] += rowbuf[(it & 0xFFFFu) * DD + lane];
            __syncwarp();
        }
        // the group's rows: one contiguous range of the value array, written once, coalesced
        double* out = vals + (int64_t)e_lo * DD;
        for (int e = lane; e < nval; e += 32) out[e] = rowbuf[e];
        // block-scalar copy for the single-precision preconditioner (nsb_matrix.diag32), while the blocks are still
        // in shared memory: saves the 6.5 GB re-read of nsb_bsr_block_scalar_f32
        if (scal32 != nullptr)
            for (int e = lane; e < nval / DD; e += 32) {
                double sd = 0.0;
#pragma unroll
                for (int i = 0; i < D; ++i) sd += rowbuf[e * DD + i * (D + 1)];
                scal32[e_lo + e] = (float)(sd / D);
            }
        __syncwarp();
    }
}
// launcher used by nsb_assemble_ipcs_momentum (assemble.cu); returns 1 when the row-group path
// does not apply (scatter schedule not registered or a row longer than the shared row buffer) so that the
// caller keeps the generic one-thread-per-row kernel.
template <int D>
static int launch_rowgroups(nsb_ctx* ctx, const nsb_pattern_dev& P, const double* ustar, double mass_coef,
                            int newton, double* A_out, float* S32_out, cudaStream_t st) {
    using C = JacCfg<D>;
    if (!P.group_row || !P.chunk_first || !P.round_ptr || !P.sched || !P.fold_ptr || P.ngroups <= 0 || P.group_max_entries > C::MAXB)
        return 1;
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
        P.ngroups, P.group_row, P.chunk_first, P.round_ptr, P.sched, P.fold_ptr, P.fold, P.rowptr, P.pair_ptr, P.pair_cell, P.pair_iloc,
        ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, mass_coef, ctx->coef.c_conv, ctx->coef.c_visc, newton, A_out, S32_out);
    NSB_LAUNCH_CHECK(ctx, "k_jacobian_rowgroups");
    return 0;
}

int nsb_launch_jacobian_rowgroups(nsb_ctx* ctx, const double* ustar, double mass_coef, int newton, double* A_out,
                                  float* S32_out, cudaStream_t st) {
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P2P2];
    if (ctx->dim == 2) return launch_rowgroups<2>(ctx, P, ustar, mass_coef, newton, A_out, S32_out, st);
    if (ctx->dim == 3) return launch_rowgroups<3>(ctx, P, ustar, mass_coef, newton, A_out, S32_out, st);
    return 1;
}
