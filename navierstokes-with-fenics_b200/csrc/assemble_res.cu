// assemble_res.cu -- K1, the IPCS tentative-velocity residual F(u*) (source/ns_ipcs_solver.py:126-130),
// as a two-phase atomic-free assembly:
//   phase 1  one thread per CELL evaluates the whole element vector once (closed form, exact
//            reference tensors) and stores it in a scratch array [ncells][NL][D];
//   phase 2  one thread per ROW (scalar P2 node) sums its incident cells' entries in ascending
//            cell order (dolfin's summation order) and writes b once.
// Every element vector is computed once (the one-thread-per-row kernel of assemble.cu recomputes
// the cell setup for each of the NL rows of a cell), the scatter is a deterministic gather, and
// the scratch traffic (2 x 8 NL D ncells bytes) is small against the compute it saves.
// All threads of a warp run the same (I, L, M) at the same time, so the run-time indexed
// tables stay in __constant__ memory (pure broadcasts).
//
//   F_Ib = (1/k) detJ sum_L M2_IL acc_Lb + c_nu sum_L A_IL U_Lb
//        + c_c detJ sum_M sum_j gM[M][b][j] w_I[M][j] - c_p detJ sum_j K[j][b] sum_M P_M DP_IjM
//   acc = a0 u* + a1 u^n + a2 u^{n-1},  A_IL = sum_jk G_jk D2_ILjk,  w_I[M][j] = sum_L T_ILM U_Lj,
//   gM[M][b][j] = d_j u_b (x_M)   (the gradient of a P2 field is P1: exact from its vertex values)
#include "common.cuh"
#include "tables.h"
#include "fe_geometry.cuh"

namespace {

template <int D> struct RT;
template <> struct RT<2> {
    static constexpr int NL = 6;
    __device__ static __forceinline__ double M2(int i, int j) { return M2_2D[i][j]; }
    __device__ static __forceinline__ double D2(int i, int j, int a, int b) { return D2_2D[i][j][a][b]; }
    __device__ static __forceinline__ double DP(int i, int a, int m) { return DP_2D[i][a][m]; }
    __device__ static __forceinline__ double T(int i, int j, int m) { return T_2D[i][j][m]; }
    __device__ static __forceinline__ double DV(int j, int m, int k) { return DV_2D[j][m][k]; }
};
template <> struct RT<3> {
    static constexpr int NL = 10;
    __device__ static __forceinline__ double M2(int i, int j) { return M2_3D[i][j]; }
    __device__ static __forceinline__ double D2(int i, int j, int a, int b) { return D2_3D[i][j][a][b]; }
    __device__ static __forceinline__ double DP(int i, int a, int m) { return DP_3D[i][a][m]; }
    __device__ static __forceinline__ double T(int i, int j, int m) { return T_3D[i][j][m]; }
    __device__ static __forceinline__ double DV(int j, int m, int k) { return DV_3D[j][m][k]; }
};

template <int D>
__global__ void __launch_bounds__(128)
k_residual_cells(int32_t ncells, const double* __restrict__ coords, const int32_t* __restrict__ cell_nodes,
                 const int32_t* __restrict__ cell_verts,
                 const double* __restrict__ ustar, const double* __restrict__ un, const double* __restrict__ unm1,
                 const double* __restrict__ pn, nsb_coef cf, double* __restrict__ scratch) {
    using TB = RT<D>;
    constexpr int NL = TB::NL, NV = D + 1;
    const int32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= ncells) return;
    int32_t nodes[NL];
#pragma unroll
    for (int L = 0; L < NL; ++L) nodes[L] = cell_nodes[(int64_t)c * NL + L];
    int32_t vtx[D + 1];
#pragma unroll
    for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
    Geo<D> g;
    cell_geometry(coords, vtx, g);
    double U[NL][D], A[NL][D];
#pragma unroll
    for (int L = 0; L < NL; ++L)
#pragma unroll
        for (int b = 0; b < D; ++b) {
            const int64_t dof = (int64_t)D * nodes[L] + b;
            const double u = ustar[dof];
            U[L][b] = u;
            A[L][b] = cf.a0 * u + cf.a1 * un[dof] + cf.a2 * unm1[dof];
        }
    double P[NV];
#pragma unroll
    for (int M = 0; M < NV; ++M) P[M] = pn[vtx[M]];
    double G[D][D];
    metric<D>(g, G);
    // vertex velocity gradients gM[M][b][j] = d_j u_b (x_M)
    double gM[NV][D][D];
#pragma unroll
    for (int M = 0; M < NV; ++M) {
        double ref[D][D];
#pragma unroll
        for (int b = 0; b < D; ++b)
#pragma unroll
            for (int k = 0; k < D; ++k) ref[b][k] = 0.0;
#pragma unroll
        for (int L = 0; L < NL; ++L)
#pragma unroll
            for (int k = 0; k < D; ++k) {
                const double d = TB::DV(L, M, k);
#pragma unroll
                for (int b = 0; b < D; ++b) ref[b][k] += U[L][b] * d;
            }
#pragma unroll
        for (int b = 0; b < D; ++b)
#pragma unroll
            for (int j = 0; j < D; ++j) {
                double s = 0.0;
#pragma unroll
                for (int k = 0; k < D; ++k) s += ref[b][k] * g.K[k][j];
                gM[M][b][j] = s;
            }
    }
    const double mk = g.detJ / cf.k, cc = cf.c_conv * g.detJ, cp = cf.c_p * g.detJ;
    double* out = scratch + (int64_t)c * NL * D;
#pragma unroll 1
    for (int I = 0; I < NL; ++I) {
        double F[D], w[NV][D];
#pragma unroll
        for (int b = 0; b < D; ++b) F[b] = 0.0;
#pragma unroll
        for (int M = 0; M < NV; ++M)
#pragma unroll
            for (int j = 0; j < D; ++j) w[M][j] = 0.0;
#pragma unroll
        for (int L = 0; L < NL; ++L) {
            double a = 0.0;
#pragma unroll
            for (int j = 0; j < D; ++j)
#pragma unroll
                for (int k = 0; k < D; ++k) a += G[j][k] * TB::D2(I, L, j, k);
            a *= cf.c_visc;
            const double m = mk * TB::M2(I, L);
#pragma unroll
            for (int b = 0; b < D; ++b) F[b] += m * A[L][b] + a * U[L][b];
#pragma unroll
            for (int M = 0; M < NV; ++M) {
                const double t = TB::T(I, L, M);
#pragma unroll
                for (int j = 0; j < D; ++j) w[M][j] += t * U[L][j];
            }
        }
#pragma unroll
        for (int b = 0; b < D; ++b) {
            double s = 0.0;
#pragma unroll
            for (int M = 0; M < NV; ++M)
#pragma unroll
                for (int j = 0; j < D; ++j) s += gM[M][b][j] * w[M][j];
            F[b] += cc * s;
        }
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double pj = 0.0;
#pragma unroll
            for (int M = 0; M < NV; ++M) pj += P[M] * TB::DP(I, j, M);
#pragma unroll
            for (int b = 0; b < D; ++b) F[b] -= cp * pj * g.K[j][b];
        }
#pragma unroll
        for (int b = 0; b < D; ++b) out[I * D + b] = F[b];
    }
}

template <int D, int NL>
__global__ void __launch_bounds__(256)
k_gather_rows(int32_t nrows, const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
              const uint8_t* __restrict__ pair_iloc, const double* __restrict__ scratch, double* __restrict__ b_out) {
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    double F[D];
#pragma unroll
    for (int b = 0; b < D; ++b) F[b] = 0.0;
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const double* src = scratch + ((int64_t)pair_cell[p] * NL + pair_iloc[p]) * D;
#pragma unroll
        for (int b = 0; b < D; ++b) F[b] += src[b];
    }
#pragma unroll
    for (int b = 0; b < D; ++b) b_out[(int64_t)D * row + b] = F[b];
}

}  // namespace

// 0 = launched, 1 = not applicable (no scratch registered), -1 = error
int nsb_launch_residual_cells(nsb_ctx* ctx, const double* ustar, const double* un, const double* unm1,
                              const double* pn, double* b_out, cudaStream_t st) {
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P2P2];
    const int D = ctx->dim, NL = D == 2 ? 6 : 10;
    if (!ctx->scratch || ctx->scratch_doubles < (int64_t)ctx->ncells * NL * D) return 1;
    const int B1 = nsb_blocks(ctx->ncells, 128), B2 = nsb_blocks(P.nrows, 256);
    if (D == 2) {
        k_residual_cells<2><<<B1, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, un, unm1, pn, ctx->coef,
                                               ctx->scratch);
        k_gather_rows<2, 6><<<B2, 256, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->scratch, b_out);
    } else {
        k_residual_cells<3><<<B1, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, un, unm1, pn, ctx->coef,
                                               ctx->scratch);
        k_gather_rows<3, 10><<<B2, 256, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->scratch, b_out);
    }
    NSB_LAUNCH_CHECK(ctx, "k_residual_cells");
    NSB_COUNT(ctx, 1);
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant of the residual assembly (no counterpart in the reference; added
// for callers that cannot allocate: CUDA-graph capture, tight serving loops, harnesses that forbid
// cudaMalloc).  The element-vector scratch of the two-phase assembly lives in a file-scope device array of a fixed,
// modest size instead of the caller-registered workspace of nsb_set_scratch; every GPU that loads the library holds
// these 16 MiB until the process exits.  ALL calls (of all contexts) on one GPU share the array: calls that overlap
// on the GPU (different streams) race.  Nothing in it has to start at zero -- phase 1 writes every entry phase 2
// reads -- so it is never cleared; it keeps the last call's element vectors.
// ---------------------------------------------------------------------------------------------
#define NSB_STATIC_SCRATCH_DOUBLES (2 * 1024 * 1024)      /* 16 MiB: 69 905 tetrahedra or 174 762 triangles */
static __device__ __align__(256) double g_static_scratch[NSB_STATIC_SCRATCH_DOUBLES];

#define NSB_STATIC_MAX_DEVICES 64
static double* g_static_scratch_addr[NSB_STATIC_MAX_DEVICES];   // device address of the array, per device (host code
                                                                // must not pass the symbol itself to a kernel)

extern "C" int64_t nsb_static_scratch_doubles(void) { return NSB_STATIC_SCRATCH_DOUBLES; }

extern "C" int nsb_assemble_ipcs_momentum_residual_noalloc(nsb_ctx* ctx, const double* ustar, const double* un,
                                                           const double* unm1, const double* pn, double* b_out,
                                                           void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, ustar && un && unm1 && pn && b_out, "null vector argument");
    const nsb_pattern_dev& P = ctx->pat[NSB_PAT_P2P2];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr && P.pair_cell && P.pair_iloc, "P2xP2 pattern not set");
    NSB_REQUIRE(ctx, ctx->ncells > 0 && ctx->coords && ctx->cell_nodes && ctx->cell_verts, "mesh not set");
    NSB_REQUIRE(ctx, ctx->dim == 2 || ctx->dim == 3, "dimension must be 2 or 3");
    NSB_REQUIRE(ctx, ctx->form_conv == 0 && ctx->form_visc == 0 && ctx->coriolis[0] == 0.0 && ctx->coriolis[1] == 0.0 &&
                ctx->coriolis[2] == 0.0,
                "the allocation-free residual covers the default weak form only (standard convective term, reduced "
                "viscous term, no Coriolis term): use nsb_assemble_ipcs_momentum");
    const int D = ctx->dim, NL = D == 2 ? 6 : 10;
    const int64_t need = (int64_t)ctx->ncells * NL * D;
    if (need > (int64_t)NSB_STATIC_SCRATCH_DOUBLES) {
        char msg[256];
        snprintf(msg, sizeof(msg), "mesh too large for the allocation-free residual: %d cells need %lld doubles of element "
                 "vectors, the static scratch holds %lld (at most %lld cells in %dD); use nsb_assemble_ipcs_momentum with "
                 "nsb_set_scratch", (int)ctx->ncells, (long long)need, (long long)NSB_STATIC_SCRATCH_DOUBLES,
                 (long long)(NSB_STATIC_SCRATCH_DOUBLES / (NL * D)), D);
        return nsb_fail(ctx, "%s%s", msg, "");
    }
    // set-up, once per device, with that device current: the device address of the static array
    int dev = -1;
    NSB_CUDA(ctx, cudaGetDevice(&dev));
    NSB_REQUIRE(ctx, dev == ctx->device, "the context's device must be current");
    NSB_REQUIRE(ctx, dev >= 0 && dev < NSB_STATIC_MAX_DEVICES, "device index out of range");
    if (g_static_scratch_addr[dev] == nullptr) {
        double* ptr = nullptr;
        NSB_CUDA(ctx, cudaGetSymbolAddress((void**)&ptr, g_static_scratch));
        NSB_REQUIRE(ctx, ptr != nullptr, "cudaGetSymbolAddress returned a null address");
        g_static_scratch_addr[dev] = ptr;
    }
    double* scratch = g_static_scratch_addr[dev];
    cudaStream_t st = (cudaStream_t)stream;
    const int B1 = nsb_blocks(ctx->ncells, 128), B2 = nsb_blocks(P.nrows, 256);
    if (D == 2) {
        k_residual_cells<2><<<B1, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, un, unm1, pn,
                                               ctx->coef, scratch);
        k_gather_rows<2, 6><<<B2, 256, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, scratch, b_out);
    } else {
        k_residual_cells<3><<<B1, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_nodes, ctx->cell_verts, ustar, un, unm1, pn,
                                               ctx->coef, scratch);
        k_gather_rows<3, 10><<<B2, 256, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, scratch, b_out);
    }
    NSB_LAUNCH_CHECK(ctx, "k_residual_cells (static scratch)");
    NSB_COUNT(ctx, 1);
    return 0;
}
