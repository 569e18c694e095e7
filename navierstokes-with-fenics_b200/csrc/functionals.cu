// functionals.cu -- exterior-facet functional of the DFG benchmark post-processing (K15, a13):
//   F = - int_Gamma ( -p n + c_visc * sym(grad u) n ) ds,     drag = 2 F_x, lift = 2 F_y
// (demo/dfg_benchmark.py:44-66: traction = -p n + 1/Re dot(0.5 (grad u + grad u^T), n), forces are
// assemble(-traction[i] * ds(cylinder))).  One thread per marked facet; on an affine cell grad u of
// a P2 field and p (P1) are linear along the facet, so the facet-centroid value times the facet
// measure is the exact integral (what FFC's facet quadrature returns up to round-off).  Deterministic
// grid-wide sum, result fetched to the host.
#include "fe_geometry.cuh"

namespace {
template <int D> struct LocalEdges;
template <> struct LocalEdges<2> { static constexpr int N = 3; };
template <> struct LocalEdges<3> { static constexpr int N = 6; };
// UFC local edges (vertices of edge e), SURVEY.md App. B.1
__constant__ int c_edge2[3][2] = {{1, 2}, {0, 2}, {0, 1}};
__constant__ int c_edge3[6][2] = {{2, 3}, {1, 3}, {1, 2}, {0, 3}, {0, 2}, {0, 1}};

template <int D>
__global__ void __launch_bounds__(128)
k_boundary_force(int32_t nfacets, const int32_t* __restrict__ facet_cell, const uint8_t* __restrict__ facet_opp,
                 const double* __restrict__ coords, const int32_t* __restrict__ cell_verts,
                 const int32_t* __restrict__ cell_nodes, const double* __restrict__ u, const double* __restrict__ p,
                 double c_visc, double* partials, unsigned int* counter, double* out) {
    constexpr int NV = D + 1, NE = LocalEdges<D>::N, NL = NV + NE;
    double F[3] = {0.0, 0.0, 0.0};
    const int32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    if (f < nfacets) {
        const int32_t c = facet_cell[f];
        const int opp = facet_opp[f];
        int32_t vtx[NV];
#pragma unroll
        for (int M = 0; M < NV; ++M) vtx[M] = cell_verts[(int64_t)c * NV + M];
        Geo<D> g;
        cell_geometry(coords, vtx, g);
        double gl[NV][D];                  // physical gradients of the barycentric coordinates
#pragma unroll
        for (int M = 0; M < NV; ++M)
#pragma unroll
            for (int i = 0; i < D; ++i) gl[M][i] = grad_lambda<D>(g, M, i);
        // facet centroid: lambda = 1/D on the facet's vertices, 0 at the opposite vertex
        double lam[NV];
#pragma unroll
        for (int M = 0; M < NV; ++M) lam[M] = (M == opp) ? 0.0 : 1.0 / D;
        // n |f| = -D |T| grad(lambda_opp),  |T| = detJ / D!
        const double vol = g.detJ / (D == 2 ? 2.0 : 6.0);
        double nA[D];
#pragma unroll
        for (int i = 0; i < D; ++i) nA[i] = -(double)D * vol * gl[opp][i];
        double pbar = 0.0;
#pragma unroll
        for (int M = 0; M < NV; ++M) pbar += lam[M] * p[vtx[M]];
        // grad u (b, j) = sum_L u_Lb d_j phi_L at the centroid
        double G[D][D];
#pragma unroll
        for (int b = 0; b < D; ++b)
#pragma unroll
            for (int j = 0; j < D; ++j) G[b][j] = 0.0;
        for (int L = 0; L < NL; ++L) {
            double dphi[D];
            if (L < NV) {
#pragma unroll
                for (int j = 0; j < D; ++j) dphi[j] = (4.0 * lam[L] - 1.0) * gl[L][j];
            } else {
                const int e = L - NV;
                const int a = (D == 2) ? c_edge2[e][0] : c_edge3[e][0];
                const int bb = (D == 2) ? c_edge2[e][1] : c_edge3[e][1];
#pragma unroll
                for (int j = 0; j < D; ++j) dphi[j] = 4.0 * (lam[a] * gl[bb][j] + lam[bb] * gl[a][j]);
            }
            const int64_t node = cell_nodes[(int64_t)c * NL + L];
#pragma unroll
            for (int b = 0; b < D; ++b) {
                const double ub = u[node * D + b];
#pragma unroll
                for (int j = 0; j < D; ++j) G[b][j] += ub * dphi[j];
            }
        }
#pragma unroll
        for (int b = 0; b < D; ++b) {
            double s = 0.0;
#pragma unroll
            for (int j = 0; j < D; ++j) s += 0.5 * (G[b][j] + G[j][b]) * nA[j];
            F[b] = pbar * nA[b] - c_visc * s;          // -( -p n + c_visc sym n ) |f|
        }
    }
    __shared__ double red[3 * 4];
    grid_sum<3, 128>(F, partials, counter, out, red);
}
}  // namespace

extern "C" int nsb_boundary_force(nsb_ctx* ctx, int32_t nfacets, const int32_t* facet_cell_dev,
                                  const uint8_t* facet_opp_dev, const double* u, const double* p, double c_visc,
                                  double* force_out_host3, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, nfacets > 0 && facet_cell_dev && facet_opp_dev && u && p && force_out_host3, "bad arguments");
    NSB_REQUIRE(ctx, ctx->coords && ctx->cell_verts && ctx->cell_nodes, "mesh not set");
    cudaStream_t st = (cudaStream_t)stream;
    const int B = nsb_blocks(nfacets, 128);
    NSB_REQUIRE(ctx, B <= NSB_MAX_PARTIALS, "too many boundary facets for one reduction");
    double* out = ctx->d_scalars + 240;
    {
        NSB_TIMED(ctx, NSB_T_VEC, st);
        if (ctx->dim == 2)
            k_boundary_force<2><<<B, 128, 0, st>>>(nfacets, facet_cell_dev, facet_opp_dev, ctx->coords, ctx->cell_verts,
                                                   ctx->cell_nodes, u, p, c_visc, ctx->d_partials, ctx->d_counter, out);
        else
            k_boundary_force<3><<<B, 128, 0, st>>>(nfacets, facet_cell_dev, facet_opp_dev, ctx->coords, ctx->cell_verts,
                                                   ctx->cell_nodes, u, p, c_visc, ctx->d_partials, ctx->d_counter, out);
    }
    NSB_LAUNCH_CHECK(ctx, "k_boundary_force");
    if (nsb_allreduce_sum(ctx, out, 3, st)) return -1;
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars + 240, out, sizeof(double) * 3, cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    for (int i = 0; i < 3; ++i) force_out_host3[i] = ctx->h_scalars[240 + i];
    return 0;
}
