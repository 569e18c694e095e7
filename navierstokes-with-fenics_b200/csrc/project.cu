// project.cu -- SURVEY.md 8f rows 1-3 on the device:
//   * right-hand side of dolfin.project(expr, V|Q) (source/ns_solver_base.py:1151,1168): the Expression is
//     interpolated into P_k on every cell (App. A.7), b_I = detJ sum_J R_IJ f_J with the exact reference
//     tensor R_IJ = int phi_I psi^k_J handed over by the host (navierstokes-with-fenics_b200/lagrange.py);
//     the mass solve is the existing CG.  Atomic-free row gather like every other vector assembly here.
//   * dolfin.errornorm (convergence_test/taylor_green_vortex.py:101-119, App. A.7): both fields
//     interpolated into P_{k+3} per cell, the difference integrated by an exact quadrature rule.
//   * the CFL diagnostic of source/ns_problem.py:554-587: DG-k LocalSolver projection of
//     k |u| dt / h_T (quadrature degree 2k), then the max-norm of the dof vector.  Per cell that projection
//     is ONE fixed matrix (nodes x quadrature points) applied to the integrand's point values.
//   * x -= s (mean-pressure shift, source/ns_solver_base.py:1190-1203).
//   * the DG post-processing fields of source/ns_problem.py:55-102: dolfin.project(curl(u), DG_1) and
//     dolfin.project(grad(p), [DG_0]^d).  The curl of a P2 field is linear and the gradient of a P1 field is
//     constant on every cell, so the cell-wise L2 projection reproduces them: the DG dofs are the vertex values
//     of curl(u_h) / the constant grad(p_h), in closed form from the vertex gradients of the P2 basis.
#include "common.cuh"
#include "tables.h"
#include "fe_geometry.cuh"

#define GRID_STRIDE(i, n) \
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

namespace {

template <int D>
__device__ __forceinline__ double cell_detj(const double* __restrict__ coords, const int32_t* __restrict__ cell_verts, int32_t c) {
    int32_t vtx[D + 1];
#pragma unroll
    for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[(int64_t)c * (D + 1) + M];
    Geo<D> g;
    cell_geometry(coords, vtx, g);
    return g.detJ;
}

// b[row*ncomp + c] = sum_{(cell, I) incident to row} detJ sum_J R[I][J] F[cell][J][c]
template <int D>
__global__ void __launch_bounds__(128)
k_cell_rhs(int32_t nrows, const int32_t* __restrict__ pair_ptr, const int32_t* __restrict__ pair_cell,
           const uint8_t* __restrict__ pair_iloc, const double* __restrict__ coords,
           const int32_t* __restrict__ cell_verts, int nk, int ncomp, const double* __restrict__ R,
           const double* __restrict__ F, double* __restrict__ b_out) {
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nrows) return;
    double acc[3] = {0.0, 0.0, 0.0};
    for (int32_t p = pair_ptr[row]; p < pair_ptr[row + 1]; ++p) {
        const int32_t c = pair_cell[p];
        const double* r = R + (int)pair_iloc[p] * nk;
        const double* f = F + (int64_t)c * nk * ncomp;
        const double dj = cell_detj<D>(coords, cell_verts, c);
        double s[3] = {0.0, 0.0, 0.0};
        for (int J = 0; J < nk; ++J) {
            const double rj = r[J];
            for (int k = 0; k < ncomp; ++k) s[k] += rj * f[J * ncomp + k];
        }
        for (int k = 0; k < ncomp; ++k) acc[k] += dj * s[k];
    }
    for (int k = 0; k < ncomp; ++k) b_out[(int64_t)row * ncomp + k] = acc[k];
}

// sum over cells of detJ sum_q w_q | sum_J Psi[q][J] e_J |^2,  e_J = F[cell][J] - sum_L E[J][L] u_h[cell dof L]
template <int D>
__global__ void __launch_bounds__(128)
k_cell_error(int32_t ncells, const double* __restrict__ coords, const int32_t* __restrict__ cell_verts,
             const int32_t* __restrict__ cell_map, int nloc, int ncomp, const double* __restrict__ uh,
             int nk, const double* __restrict__ E, int nq, const double* __restrict__ Psi,
             const double* __restrict__ wq, const double* __restrict__ F, double* __restrict__ ework,
             double* partials, unsigned int* counter, double* out) {
    double s[1] = {0.0};
    GRID_STRIDE(c, ncells) {
        double* e = ework + (int64_t)c * nk * ncomp;        // caller-owned scratch [ncells][nk][ncomp]
        for (int J = 0; J < nk; ++J)
            for (int k = 0; k < ncomp; ++k) {
                double v = F[((int64_t)c * nk + J) * ncomp + k];
                for (int L = 0; L < nloc; ++L)
                    v -= E[J * nloc + L] * uh[(int64_t)ncomp * cell_map[(int64_t)c * nloc + L] + k];
                e[J * ncomp + k] = v;
            }
        double acc = 0.0;
        for (int q = 0; q < nq; ++q) {
            double v[3] = {0.0, 0.0, 0.0};
            for (int J = 0; J < nk; ++J) {
                const double ps = Psi[q * nk + J];
                for (int k = 0; k < ncomp; ++k) v[k] += ps * e[J * ncomp + k];
            }
            double m = 0.0;
            for (int k = 0; k < ncomp; ++k) m += v[k] * v[k];
            acc += wq[q] * m;
        }
        s[0] += acc * cell_detj<D>(coords, cell_verts, (int32_t)c);
    }
    __shared__ double red[4];
    grid_sum<1, 128>(s, partials, counter, out, red);
}

// max over cells and DG nodes of | sum_q P[node][q] f_q |,  f_q = scale |u_h(x_q)| / h_T,  h_T = longest edge
template <int D, int NL>
__global__ void __launch_bounds__(128)
k_cfl(int32_t ncells, const double* __restrict__ coords, const int32_t* __restrict__ cell_verts,
      const int32_t* __restrict__ cell_nodes, const double* __restrict__ u, double scale, int nq,
      const double* __restrict__ B /* [nq][NL] P2 basis at the quadrature points */,
      const double* __restrict__ P /* [NL][nq] */, double* partial_max) {
    double best = 0.0;
    GRID_STRIDE(c, ncells) {
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[c * (D + 1) + M];
        double h2 = 0.0;
#pragma unroll
        for (int a = 0; a <= D; ++a)
#pragma unroll
            for (int b = a + 1; b <= D; ++b) {
                double s = 0.0;
#pragma unroll
                for (int i = 0; i < D; ++i) {
                    const double t = coords[D * vtx[a] + i] - coords[D * vtx[b] + i];
                    s += t * t;
                }
                h2 = fmax(h2, s);
            }
        const double f0 = scale / sqrt(h2);
        double U[NL][D];
#pragma unroll
        for (int L = 0; L < NL; ++L)
#pragma unroll
            for (int i = 0; i < D; ++i) U[L][i] = u[(int64_t)D * cell_nodes[c * NL + L] + i];
        double coef[NL];
#pragma unroll
        for (int I = 0; I < NL; ++I) coef[I] = 0.0;
        for (int q = 0; q < nq; ++q) {
            double v[D];
#pragma unroll
            for (int i = 0; i < D; ++i) v[i] = 0.0;
#pragma unroll
            for (int L = 0; L < NL; ++L) {
                const double bq = B[q * NL + L];
#pragma unroll
                for (int i = 0; i < D; ++i) v[i] += bq * U[L][i];
            }
            double m = 0.0;
#pragma unroll
            for (int i = 0; i < D; ++i) m += v[i] * v[i];
            const double fq = f0 * sqrt(m);
#pragma unroll
            for (int I = 0; I < NL; ++I) coef[I] += P[I * nq + q] * fq;
        }
#pragma unroll
        for (int I = 0; I < NL; ++I) best = fmax(best, fabs(coef[I]));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
    __shared__ double red[4];
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) partial_max[blockIdx.x] = fmax(fmax(red[0], red[1]), fmax(red[2], red[3]));
}

__global__ void k_max_finish(int n, const double* __restrict__ partial_max, double* out) {
    double best = 0.0;
    for (int i = threadIdx.x; i < n; i += 32) best = fmax(best, partial_max[i]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = fmax(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (threadIdx.x == 0) *out = best;
}

__global__ void k_vec_shift(int64_t n, double s, double* __restrict__ x) {
    GRID_STRIDE(i, n) x[i] -= s;
}

int fetch0(nsb_ctx* ctx, double* out_host, cudaStream_t st) {
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, sizeof(double), cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    *out_host = ctx->h_scalars[0];
    return 0;
}

}  // namespace

extern "C" int nsb_assemble_cell_rhs(nsb_ctx* ctx, int32_t which, int32_t nk, int32_t ncomp, const double* table_dev,
                                     const double* cell_vals_dev, double* b_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which == NSB_PAT_P2P2 || which == NSB_PAT_P1P1, "row space must be P2 nodes or vertices");
    NSB_REQUIRE(ctx, nk >= 1 && ncomp >= 1 && ncomp <= 3 && table_dev && cell_vals_dev && b_out, "bad arguments");
    const nsb_pattern_dev& P = ctx->pat[which];
    NSB_REQUIRE(ctx, P.nrows > 0 && P.pair_ptr, "pattern not set");
    cudaStream_t st = (cudaStream_t)stream;
    const int T = 128, B = nsb_blocks(P.nrows, T);
    NSB_TIMED(ctx, NSB_T_ASM_RHS, st);
    if (ctx->dim == 2)
        k_cell_rhs<2><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords, ctx->cell_verts, nk, ncomp,
                                       table_dev, cell_vals_dev, b_out);
    else
        k_cell_rhs<3><<<B, T, 0, st>>>(P.nrows, P.pair_ptr, P.pair_cell, P.pair_iloc, ctx->coords, ctx->cell_verts, nk, ncomp,
                                       table_dev, cell_vals_dev, b_out);
    NSB_LAUNCH_CHECK(ctx, "k_cell_rhs");
    return 0;
}

extern "C" int nsb_cell_error_norm2(nsb_ctx* ctx, int32_t space, int32_t ncomp, const double* uh, int32_t nk,
                                    const double* E_dev, int32_t nq, const double* Psi_dev, const double* wq_dev,
                                    const double* cell_vals_dev, double* work_dev, double* out_host, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, space == NSB_PAT_P2P2 || space == NSB_PAT_P1P1, "space must be P2 nodes or vertices");
    NSB_REQUIRE(ctx, ncomp >= 1 && ncomp <= 3 && uh && E_dev && Psi_dev && wq_dev && cell_vals_dev && work_dev && out_host,
                "bad arguments");
    NSB_REQUIRE(ctx, ctx->nranks == 1, "errornorm runs on one GPU (replicas only)");
    cudaStream_t st = (cudaStream_t)stream;
    const int D = ctx->dim;
    const int nloc = space == NSB_PAT_P1P1 ? D + 1 : (D == 2 ? 6 : 10);
    const int32_t* map = space == NSB_PAT_P1P1 ? ctx->cell_verts : ctx->cell_nodes;
    int G = nsb_blocks(ctx->ncells, 128);
    if (G > NSB_MAX_PARTIALS) G = NSB_MAX_PARTIALS;
    {
        NSB_TIMED(ctx, NSB_T_VEC, st);
        if (D == 2)
            k_cell_error<2><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, map, nloc, ncomp, uh, nk, E_dev, nq,
                                               Psi_dev, wq_dev, cell_vals_dev, work_dev, ctx->d_partials, ctx->d_counter,
                                               ctx->d_scalars);
        else
            k_cell_error<3><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, map, nloc, ncomp, uh, nk, E_dev, nq,
                                               Psi_dev, wq_dev, cell_vals_dev, work_dev, ctx->d_partials, ctx->d_counter,
                                               ctx->d_scalars);
    }
    NSB_LAUNCH_CHECK(ctx, "k_cell_error");
    return fetch0(ctx, out_host, st);
}

extern "C" int nsb_cfl_max(nsb_ctx* ctx, const double* u, double scale, int32_t nq, const double* B_dev,
                           const double* P_dev, double* out_host, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, u && B_dev && P_dev && out_host && nq >= 1, "bad arguments");
    NSB_REQUIRE(ctx, ctx->ncells > 0 && ctx->cell_nodes, "mesh not set");
    cudaStream_t st = (cudaStream_t)stream;
    int G = nsb_blocks(ctx->ncells, 128);
    if (G > NSB_MAX_PARTIALS) G = NSB_MAX_PARTIALS;
    {
        NSB_TIMED(ctx, NSB_T_VEC, st);
        if (ctx->dim == 2)
            k_cfl<2, 6><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, ctx->cell_nodes, u, scale, nq, B_dev, P_dev,
                                          ctx->d_partials);
        else
            k_cfl<3, 10><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, ctx->cell_nodes, u, scale, nq, B_dev, P_dev,
                                           ctx->d_partials);
        k_max_finish<<<1, 32, 0, st>>>(G, ctx->d_partials, ctx->d_scalars);
    }
    NSB_LAUNCH_CHECK(ctx, "k_cfl");
    NSB_COUNT(ctx, 1);
    if (fetch0(ctx, out_host, st)) return -1;
    if (ctx->nranks > 1) {
        // every rank integrates its own cells + the ghost layer: the max over ranks is the global max.
        // max(a_r) = -min(-a_r); the all-reduce primitive here sums, so exchange through a sum of one-hot slots
        double* ds = ctx->d_scalars;
        NSB_REQUIRE(ctx, ctx->nranks <= 32, "at most 32 ranks");
        for (int r = 0; r < ctx->nranks; ++r) ctx->h_scalars[64 + r] = r == ctx->rank ? *out_host : 0.0;
        NSB_CUDA(ctx, cudaMemcpyAsync(ds + 16, ctx->h_scalars + 64, sizeof(double) * ctx->nranks, cudaMemcpyHostToDevice, st));
        if (nsb_allreduce_sum(ctx, ds + 16, ctx->nranks, st)) return -1;
        NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars + 64, ds + 16, sizeof(double) * ctx->nranks, cudaMemcpyDeviceToHost, st));
        NSB_CUDA(ctx, cudaStreamSynchronize(st));
        double m = 0.0;
        for (int r = 0; r < ctx->nranks; ++r) m = fmax(m, ctx->h_scalars[64 + r]);
        *out_host = m;
    }
    return 0;
}

// ---- DG post-processing fields --------------------------------------------------------------------------
namespace {
template <int D> struct PT;
template <> struct PT<2> {
    static constexpr int NL = 6, NC = 1;
    __device__ static __forceinline__ double DV(int j, int m, int k) { return DV_2D[j][m][k]; }
};
template <> struct PT<3> {
    static constexpr int NL = 10, NC = 3;
    __device__ static __forceinline__ double DV(int j, int m, int k) { return DV_3D[j][m][k]; }
};

// out[cell][M][c]: curl(u_h) at the cell's vertex M (2D: the scalar d_x u_y - d_y u_x)
template <int D>
__global__ void __launch_bounds__(128)
k_cell_vorticity(int32_t ncells, const double* __restrict__ coords, const int32_t* __restrict__ cell_verts,
                 const int32_t* __restrict__ cell_nodes, const double* __restrict__ u, double* __restrict__ out) {
    using T = PT<D>;
    constexpr int NL = T::NL, NC = T::NC;
    GRID_STRIDE(c, ncells) {
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[c * (D + 1) + M];
        Geo<D> g;
        cell_geometry(coords, vtx, g);
        double U[NL][D];
#pragma unroll
        for (int L = 0; L < NL; ++L)
#pragma unroll
            for (int i = 0; i < D; ++i) U[L][i] = u[(int64_t)D * cell_nodes[c * NL + L] + i];
#pragma unroll
        for (int M = 0; M <= D; ++M) {
            double ref[D][D], gr[D][D];      // gr[b][j] = d_j u_b (x_M)
#pragma unroll
            for (int b = 0; b < D; ++b)
#pragma unroll
                for (int k = 0; k < D; ++k) {
                    double s = 0.0;
#pragma unroll
                    for (int L = 0; L < NL; ++L) s += U[L][b] * T::DV(L, M, k);
                    ref[b][k] = s;
                }
#pragma unroll
            for (int b = 0; b < D; ++b)
#pragma unroll
                for (int j = 0; j < D; ++j) {
                    double s = 0.0;
#pragma unroll
                    for (int k = 0; k < D; ++k) s += ref[b][k] * g.K[k][j];
                    gr[b][j] = s;
                }
            double* o = out + ((int64_t)c * (D + 1) + M) * NC;
            if (D == 2) {
                o[0] = gr[1][0] - gr[0][1];
            } else {
                o[0] = gr[2 % D][1] - gr[1][2 % D];
                o[1 % NC] = gr[0][2 % D] - gr[2 % D][0];
                o[2 % NC] = gr[1][0] - gr[0][1];
            }
        }
    }
}

// out[cell][j] = d_j p_h on the cell (P1: constant)
template <int D>
__global__ void __launch_bounds__(128)
k_cell_pressure_gradient(int32_t ncells, const double* __restrict__ coords, const int32_t* __restrict__ cell_verts,
                         const double* __restrict__ p, double* __restrict__ out) {
    GRID_STRIDE(c, ncells) {
        int32_t vtx[D + 1];
#pragma unroll
        for (int M = 0; M <= D; ++M) vtx[M] = cell_verts[c * (D + 1) + M];
        Geo<D> g;
        cell_geometry(coords, vtx, g);
        const double p0 = p[vtx[0]];
        double dp[D];
#pragma unroll
        for (int k = 0; k < D; ++k) dp[k] = p[vtx[k + 1]] - p0;     // reference gradient: lambda_k = xi_k, lambda_0 = 1 - sum
#pragma unroll
        for (int j = 0; j < D; ++j) {
            double s = 0.0;
#pragma unroll
            for (int k = 0; k < D; ++k) s += dp[k] * g.K[k][j];
            out[(int64_t)c * D + j] = s;
        }
    }
}
}  // namespace

extern "C" int nsb_cell_vorticity(nsb_ctx* ctx, const double* u, double* out, void* stream) {
    NSB_REQUIRE(ctx, ctx && u && out, "bad arguments");
    NSB_REQUIRE(ctx, ctx->ncells > 0 && ctx->cell_nodes, "set the mesh first");
    cudaStream_t st = (cudaStream_t)stream;
    const int G = nsb_blocks(ctx->ncells, 128);
    NSB_TIMED(ctx, NSB_T_VEC, st);
    if (ctx->dim == 2) k_cell_vorticity<2><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, ctx->cell_nodes, u, out);
    else k_cell_vorticity<3><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, ctx->cell_nodes, u, out);
    NSB_LAUNCH_CHECK(ctx, "k_cell_vorticity");
    NSB_COUNT(ctx, 1);
    return 0;
}

extern "C" int nsb_cell_pressure_gradient(nsb_ctx* ctx, const double* p, double* out, void* stream) {
    NSB_REQUIRE(ctx, ctx && p && out, "bad arguments");
    NSB_REQUIRE(ctx, ctx->ncells > 0, "set the mesh first");
    cudaStream_t st = (cudaStream_t)stream;
    const int G = nsb_blocks(ctx->ncells, 128);
    NSB_TIMED(ctx, NSB_T_VEC, st);
    if (ctx->dim == 2) k_cell_pressure_gradient<2><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, p, out);
    else k_cell_pressure_gradient<3><<<G, 128, 0, st>>>(ctx->ncells, ctx->coords, ctx->cell_verts, p, out);
    NSB_LAUNCH_CHECK(ctx, "k_cell_pressure_gradient");
    NSB_COUNT(ctx, 1);
    return 0;
}

extern "C" int nsb_vec_shift(nsb_ctx* ctx, int64_t n, double s, double* x, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && n >= 0, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    if (n == 0) return 0;
    int64_t b = (n + 255) / 256;
    if (b > (int64_t)ctx->sm_count * 8) b = (int64_t)ctx->sm_count * 8;
    NSB_TIMED(ctx, NSB_T_VEC, st);
    k_vec_shift<<<(int)b, 256, 0, st>>>(n, s, x);
    NSB_LAUNCH_CHECK(ctx, "k_vec_shift");
    return 0;
}
