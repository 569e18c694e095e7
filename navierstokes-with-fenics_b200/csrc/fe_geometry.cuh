// fe_geometry.cuh -- affine simplex geometry shared by the assembly kernels.
#pragma once
#include "common.cuh"

// ---- affine geometry: K[j][i] = d xi_j / d x_i, detJ = |det J| -------------------------------
template <int D> struct Geo { double K[D][D]; double detJ; };

__device__ __forceinline__ void cell_geometry(const double* __restrict__ coords, const int32_t* v, Geo<2>& g) {
    const double x0 = coords[2 * v[0]], y0 = coords[2 * v[0] + 1];
    const double j00 = coords[2 * v[1]] - x0, j10 = coords[2 * v[1] + 1] - y0;
    const double j01 = coords[2 * v[2]] - x0, j11 = coords[2 * v[2] + 1] - y0;
    const double det = j00 * j11 - j01 * j10;
    const double id = 1.0 / det;
    g.K[0][0] = j11 * id;  g.K[0][1] = -j01 * id;
    g.K[1][0] = -j10 * id; g.K[1][1] = j00 * id;
    g.detJ = fabs(det);
}

__device__ __forceinline__ void cell_geometry(const double* __restrict__ coords, const int32_t* v, Geo<3>& g) {
    double J[3][3];
    const double x0 = coords[3 * v[0]], y0 = coords[3 * v[0] + 1], z0 = coords[3 * v[0] + 2];
#pragma unroll
    for (int j = 0; j < 3; ++j) {
        J[0][j] = coords[3 * v[j + 1]] - x0;
        J[1][j] = coords[3 * v[j + 1] + 1] - y0;
        J[2][j] = coords[3 * v[j + 1] + 2] - z0;
    }
    const double c00 = J[1][1] * J[2][2] - J[1][2] * J[2][1];
    const double c01 = J[1][2] * J[2][0] - J[1][0] * J[2][2];
    const double c02 = J[1][0] * J[2][1] - J[1][1] * J[2][0];
    const double det = J[0][0] * c00 + J[0][1] * c01 + J[0][2] * c02;
    const double id = 1.0 / det;
    // K = J^-1 = adj(J)/det ; adj[j][i] = cofactor(i,j)
    g.K[0][0] = c00 * id;
    g.K[1][0] = c01 * id;
    g.K[2][0] = c02 * id;
    g.K[0][1] = (J[0][2] * J[2][1] - J[0][1] * J[2][2]) * id;
    g.K[1][1] = (J[0][0] * J[2][2] - J[0][2] * J[2][0]) * id;
    g.K[2][1] = (J[0][1] * J[2][0] - J[0][0] * J[2][1]) * id;
    g.K[0][2] = (J[0][1] * J[1][2] - J[0][2] * J[1][1]) * id;
    g.K[1][2] = (J[0][2] * J[1][0] - J[0][0] * J[1][2]) * id;
    g.K[2][2] = (J[0][0] * J[1][1] - J[0][1] * J[1][0]) * id;
    g.detJ = fabs(det);
}

// physical gradient of barycentric lambda_M: component i
template <int D>
__device__ __forceinline__ double grad_lambda(const Geo<D>& g, int M, int i) {
    if (M > 0) return g.K[M - 1][i];
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < D; ++j) s -= g.K[j][i];
    return s;
}

// G[j][k] = detJ * sum_i K[j][i] K[k][i]
template <int D>
__device__ __forceinline__ void metric(const Geo<D>& g, double (&G)[D][D]) {
#pragma unroll
    for (int j = 0; j < D; ++j)
#pragma unroll
        for (int k = 0; k < D; ++k) {
            double s = 0.0;
#pragma unroll
            for (int i = 0; i < D; ++i) s += g.K[j][i] * g.K[k][i];
            G[j][k] = s * g.detJ;
        }
}

