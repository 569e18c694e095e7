// krylov.cu -- Krylov solvers (K11, K12) replacing the sparse direct LU of the reference
// (SURVEY.md G3): preconditioned CG for the SPD pressure-Poisson / mass systems
// (source/ns_ipcs_solver.py:166-171,189-196) and right-preconditioned restarted GMRES with
// CGS-2 Arnoldi for the velocity / mixed Newton systems (source/ns_ipcs_solver.py:138-147).
// Preconditioners: Jacobi, or a Chebyshev polynomial in D^-1 A on [lambda_min, lambda_max]
// (K12; lambda_max from nsb_lambda_max power iteration).  The polynomial trades Arnoldi steps
// (whose Gram-Schmidt cost grows with the basis) for SpMVs, which run near the HBM roofline.
// Step lengths stay on the device (kernels read alpha/beta from device scalars); inner products
// are warp-shuffle + deterministic last-block reductions; vector updates are fused with the
// reductions that follow them and p.Ap is fused into the SpMV epilogue.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include <vector>
#include "spmv.cuh"

// device scalar slots
enum { S_OUT = 0, S_RZ0 = 1, S_RZ1 = 2, S_PQ = 3, S_RR = 4, S_BB = 5, S_NRM = 8, S_TMP = 11, S_H1 = 32, S_H2 = 96, S_Y = 160 };

static int krylov_grid(nsb_ctx* ctx, int64_t n) {
    int64_t b = (n + 255) / 256;
    const int64_t cap = (int64_t)ctx->sm_count * 8;
    if (b > cap) b = cap;
    return (int)(b < 1 ? 1 : b);
}

#define GRID_STRIDE(i, n) \
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)

// ---------------------------------------------------------------------------------------------
// preconditioner kernels
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_jacobi(int64_t n, const double* __restrict__ dinv, const double* __restrict__ v, double* __restrict__ z) {
    GRID_STRIDE(i, n) z[i] = dinv[i] * v[i];
}

// Chebyshev start (z0 = 0): r = D^-1 v ; d = r / theta ; z = d
__global__ void __launch_bounds__(256)
k_cheb_first(int64_t n, const double* __restrict__ dinv, const double* __restrict__ v, double inv_theta,
             double* __restrict__ r, double* __restrict__ d, double* __restrict__ z) {
    GRID_STRIDE(i, n) {
        const double c = dinv[i] * v[i];
        const double di = c * inv_theta;
        r[i] = c; d[i] = di; z[i] = di;
    }
}

// Chebyshev step: r -= D^-1 q (q = A d) ; d = a d + b r ; z += d
__global__ void __launch_bounds__(256)
k_cheb_step(int64_t n, const double* __restrict__ dinv, const double* __restrict__ q, double a, double b,
            double* __restrict__ r, double* __restrict__ d, double* __restrict__ z) {
    GRID_STRIDE(i, n) {
        const double ri = r[i] - dinv[i] * q[i];
        const double di = a * d[i] + b * ri;
        r[i] = ri; d[i] = di; z[i] += di;
    }
}

// ---- the same recurrences in single precision, one thread per node, vectors padded to NSB_PAD32(BS)
// floats per node (float4 for BS = 3) so that every access is one aligned vector and the SpMV gathers a
// block column with one load.  Input v and the result zout are ordinary FP64 dof vectors.
template <int BS> struct Pad32;
template <> struct Pad32<3> {
    static __device__ __forceinline__ void ld(const float* p, int64_t i, float (&v)[3]) {
        const float4 t = reinterpret_cast<const float4*>(p)[i];
        v[0] = t.x; v[1] = t.y; v[2] = t.z;
    }
    static __device__ __forceinline__ void st(float* p, int64_t i, const float (&v)[3]) {
        reinterpret_cast<float4*>(p)[i] = make_float4(v[0], v[1], v[2], 0.0f);
    }
};
template <> struct Pad32<2> {
    static __device__ __forceinline__ void ld(const float* p, int64_t i, float (&v)[2]) {
        const float2 t = reinterpret_cast<const float2*>(p)[i];
        v[0] = t.x; v[1] = t.y;
    }
    static __device__ __forceinline__ void st(float* p, int64_t i, const float (&v)[2]) {
        reinterpret_cast<float2*>(p)[i] = make_float2(v[0], v[1]);
    }
};

// r = D^-1 v ; d = r / theta ; z = d      (zout != null: degree 1, the result goes straight to zout)
template <int BS>
__global__ void __launch_bounds__(256)
k_cheb32_first(int64_t nnodes, const double* __restrict__ dinv, const double* __restrict__ v, float inv_theta,
               float* __restrict__ r, float* __restrict__ d, float* __restrict__ z, double* __restrict__ zout,
               unsigned long long* __restrict__ d_h = nullptr) {
    GRID_STRIDE(i, nnodes) {
        float c[BS], di[BS];
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            c[b] = (float)(dinv[i * BS + b] * v[i * BS + b]);
            di[b] = c[b] * inv_theta;
        }
        if (zout) {
#pragma unroll
            for (int b = 0; b < BS; ++b) zout[i * BS + b] = (double)di[b];
        } else {
            Pad32<BS>::st(r, i, c); Pad32<BS>::st(d, i, di); Pad32<BS>::st(z, i, di);
            if constexpr (BS == 3) {
                if (d_h) d_h[i] = nsb_pack_bf16x3(di[0], di[1], di[2]);
            }
        }
    }
}

// r -= D^-1 q ; d = a d + b r ; z += d     (zout != null: last step, only zout = z + d is written)
template <int BS>
__global__ void __launch_bounds__(256)
k_cheb32_step(int64_t nnodes, const double* __restrict__ dinv, const float* __restrict__ q, float a, float b,
              float* __restrict__ r, float* __restrict__ d, float* __restrict__ z, double* __restrict__ zout) {
    GRID_STRIDE(i, nnodes) {
        float qi[BS], ri[BS], di[BS], zi[BS];
        Pad32<BS>::ld(q, i, qi); Pad32<BS>::ld(r, i, ri); Pad32<BS>::ld(d, i, di); Pad32<BS>::ld(z, i, zi);
#pragma unroll
        for (int c = 0; c < BS; ++c) {
            ri[c] -= (float)dinv[i * BS + c] * qi[c];
            di[c] = a * di[c] + b * ri[c];
            zi[c] += di[c];
        }
        if (zout) {
#pragma unroll
            for (int c = 0; c < BS; ++c) zout[i * BS + c] = (double)zi[c];
        } else {
            Pad32<BS>::st(r, i, ri); Pad32<BS>::st(d, i, di); Pad32<BS>::st(z, i, zi);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// CG kernels
// ---------------------------------------------------------------------------------------------
// r = b - shift - q ; [JAC: z = dinv r ; p = z] ; sums: r.z (JAC), r.r, (b-shift).(b-shift)
template <bool JAC>
__global__ void __launch_bounds__(256)
k_cg_init(int64_t n, const double* __restrict__ b, const double* __restrict__ q, const double* __restrict__ dinv,
          double shift, double* __restrict__ r, double* __restrict__ z, double* __restrict__ p,
          double* partials, unsigned int* counter, double* scal) {
    double s[3] = {0.0, 0.0, 0.0};
    GRID_STRIDE(i, n) {
        const double bi = b[i] - shift;
        const double ri = bi - q[i];
        r[i] = ri;
        if (JAC) {
            const double zi = dinv[i] * ri;
            z[i] = zi; p[i] = zi;
            s[0] += ri * zi;
        }
        s[1] += ri * ri; s[2] += bi * bi;
    }
    __shared__ double red[3 * 8];
    grid_sum<3, 256>(s, partials, counter, scal + S_TMP, red);
}

__global__ void k_cg_init_finish(double* scal, int jac) {
    if (jac) scal[S_RZ0] = scal[S_TMP];
    scal[S_RR] = scal[S_TMP + 1];
    scal[S_BB] = scal[S_TMP + 2];
}

// alpha = rz/pq ; x += alpha p ; r -= alpha q ; [JAC: z = dinv r ; sum r.z] ; sum r.r
template <bool JAC>
__global__ void __launch_bounds__(256)
k_cg_update1(int64_t n, const double* __restrict__ p, const double* __restrict__ q,
             const double* __restrict__ dinv, double* __restrict__ x, double* __restrict__ r,
             double* __restrict__ z, int slot_rz, double* partials, unsigned int* counter, double* scal) {
    const double pq = scal[S_PQ];
    const double alpha = pq != 0.0 ? scal[slot_rz] / pq : 0.0;   // r = 0 reached (look-ahead iterations): no-op
    double s[2] = {0.0, 0.0};
    GRID_STRIDE(i, n) {
        x[i] += alpha * p[i];
        const double ri = r[i] - alpha * q[i];
        r[i] = ri;
        if (JAC) {
            const double zi = dinv[i] * ri;
            z[i] = zi;
            s[0] += ri * zi;
        }
        s[1] += ri * ri;
    }
    __shared__ double red[2 * 8];
    grid_sum<2, 256>(s, partials, counter, scal + S_TMP, red);
}

__global__ void k_cg_update1_finish(double* scal, int slot_rz_new, int jac) {
    if (jac) scal[slot_rz_new] = scal[S_TMP];
    scal[S_RR] = scal[S_TMP + 1];
}

// beta = rz_new / rz ; p = z + beta p     (first = 1: p = z)
__global__ void __launch_bounds__(256)
k_cg_update2(int64_t n, const double* __restrict__ z, double* __restrict__ p, int slot_rz, int slot_rz_new,
             int first, const double* __restrict__ scal) {
    const double rz_old = scal[slot_rz];
    const double beta = (first || rz_old == 0.0) ? 0.0 : scal[slot_rz_new] / rz_old;
    GRID_STRIDE(i, n) p[i] = z[i] + beta * p[i];
}

__global__ void k_shift(int64_t n, double s, double* __restrict__ x) {
    GRID_STRIDE(i, n) x[i] -= s;
}

template <bool WITH_Y>
__global__ void __launch_bounds__(256)
k_reduce(int64_t n, const double* __restrict__ x, const double* __restrict__ y, double* partials,
         unsigned int* counter, double* out) {
    double s[1] = {0.0};
    GRID_STRIDE(i, n) s[0] += WITH_Y ? x[i] * y[i] : x[i];
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

static int fetch(nsb_ctx* ctx, int slot, int count, cudaStream_t st) {
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars + slot, ctx->d_scalars + slot, sizeof(double) * count,
                                  cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// preconditioner application  z = M^-1 v   (scratch: 3 n doubles for Chebyshev)
// ---------------------------------------------------------------------------------------------
bool nsb_spmv32_available(const nsb_matrix* A);
int nsb_launch_spmv32(nsb_ctx* ctx, const nsb_matrix* A, const float* x, float* y, cudaStream_t st,
                      const nsb_cheb32_epi* epi = nullptr);

// the single-precision Chebyshev polynomial is not exactly linear in v (FP32 rounding): the flexible GMRES may use it,
// and CG on an SPD operator (the polynomial is SPD for every interval end below lambda_max; its rounding perturbs the
// search directions at the 1e-7 level, which costs CG nothing measurable: the residual recurrence stays FP64)
static bool pc_single_precision(const nsb_matrix* A, const nsb_precond* M) {
    return M->kind == NSB_PC_CHEBYSHEV && M->lowp && nsb_spmv32_available(A);
}

template <int BS>
static int apply_cheb32(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* v, double* zout,
                        double* scratch, cudaStream_t st) {
    constexpr int PB = BS == 3 ? 4 : BS;
    const int64_t nn = A->nrows;
    const int G = krylov_grid(ctx, nn);
    float* r = reinterpret_cast<float*>(((uintptr_t)scratch + 15) & ~(uintptr_t)15);   // 4 x PB nn floats <= 3 n doubles
    float *d = r + PB * nn, *q = d + PB * nn, *z = q + PB * nn;
    // bf16 shadows of the two direction buffers (gathered by the block-scalar SpMV); NSB_PC_GATHER=f32 switches them off
    const bool fused = A->diag32 != nullptr;      // block-scalar matrix: the step runs in the SpMV's epilogue
    static const bool shadow_on = !(getenv("NSB_PC_GATHER") && !strcmp(getenv("NSB_PC_GATHER"), "f32"));
    const bool shadow = fused && BS == 3 && shadow_on && M->degree > 1;
    unsigned long long* dh = reinterpret_cast<unsigned long long*>(z + PB * nn);
    unsigned long long* qh = dh + nn;
    const double theta = 0.5 * (M->lambda_max + M->lambda_min), delta = 0.5 * (M->lambda_max - M->lambda_min);
    const double sigma1 = theta / delta;
    double rho_old = 1.0 / sigma1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cheb32_first<BS><<<G, 256, 0, st>>>(nn, M->inv_diag, v, (float)(1.0 / theta), r, d, z, M->degree == 1 ? zout : nullptr,
                                              shadow ? dh : nullptr);
        NSB_LAUNCH_CHECK(ctx, "k_cheb32_first");
    }
    for (int k = 1; k < M->degree; ++k) {
        const double rho = 1.0 / (2.0 * sigma1 - rho_old);
        if (fused) {                              // d and q alternate as old / new direction
            nsb_cheb32_epi E{};
            E.dinv = M->inv_diag; E.r = r; E.dnew = q; E.z = z;
            E.zout = k == M->degree - 1 ? zout : nullptr;
            E.a = (float)(rho * rho_old); E.b = (float)(2.0 * rho / delta);
            E.x_h = shadow ? dh : nullptr; E.dnew_h = shadow ? qh : nullptr;
            if (nsb_launch_spmv32(ctx, A, d, nullptr, st, &E)) return -1;
            float* tmp = d; d = q; q = tmp;
            unsigned long long* th = dh; dh = qh; qh = th;
            rho_old = rho;
            continue;
        }
        if (nsb_launch_spmv32(ctx, A, d, q, st)) return -1;
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cheb32_step<BS><<<G, 256, 0, st>>>(nn, M->inv_diag, q, (float)(rho * rho_old), (float)(2.0 * rho / delta), r, d, z,
                                            k == M->degree - 1 ? zout : nullptr);
        NSB_LAUNCH_CHECK(ctx, "k_cheb32_step");
        rho_old = rho;
    }
    return 0;
}

static int saddle_apply(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* v, double* z,
                        double* scratch, cudaStream_t st);

static int apply_pc(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* v, double* z,
                    double* scratch, int64_t n, cudaStream_t st, bool flexible = false) {
    const int G = krylov_grid(ctx, n);
    if (M->kind == NSB_PC_SADDLE) return saddle_apply(ctx, A, M, v, z, scratch, st);
    if (flexible && pc_single_precision(A, M))
        return A->bs == 3 ? apply_cheb32<3>(ctx, A, M, v, z, scratch, st) : apply_cheb32<2>(ctx, A, M, v, z, scratch, st);
    if (M->kind == NSB_PC_NONE) {
        NSB_CUDA(ctx, cudaMemcpyAsync(z, v, sizeof(double) * n, cudaMemcpyDeviceToDevice, st));
        return 0;
    }
    if (M->kind == NSB_PC_JACOBI) {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_jacobi<<<G, 256, 0, st>>>(n, M->inv_diag, v, z);
        NSB_LAUNCH_CHECK(ctx, "k_jacobi");
        return 0;
    }
    // Chebyshev iteration on D^-1 A z = D^-1 v from z = 0 (Saad, Iterative Methods, Alg. 12.1)
    double *r = scratch, *d = scratch + n, *q = scratch + 2 * n;
    const double theta = 0.5 * (M->lambda_max + M->lambda_min), delta = 0.5 * (M->lambda_max - M->lambda_min);
    const double sigma1 = theta / delta;
    double rho_old = 1.0 / sigma1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cheb_first<<<G, 256, 0, st>>>(n, M->inv_diag, v, 1.0 / theta, r, d, z);
        NSB_LAUNCH_CHECK(ctx, "k_cheb_first");
    }
    const bool fused = A->kind == NSB_CSR;        // CSR kernels carry the step in their epilogue (d, q alternate)
    for (int k = 1; k < M->degree; ++k) {
        const double rho = 1.0 / (2.0 * sigma1 - rho_old);
        if (fused) {
            nsb_cheb64_epi E;
            E.dinv = M->inv_diag; E.r = r; E.dnew = q; E.z = z;
            E.a = rho * rho_old; E.b = 2.0 * rho / delta;
            ctx->spmv_epi = &E;
            if (nsb_launch_spmv(ctx, A, d, q, nullptr, nullptr, st)) return -1;
            double* tmp = d; d = q; q = tmp;
            rho_old = rho;
            continue;
        }
        if (nsb_launch_spmv(ctx, A, d, q, nullptr, nullptr, st, /*lowp=*/true)) return -1;
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cheb_step<<<G, 256, 0, st>>>(n, M->inv_diag, q, rho * rho_old, 2.0 * rho / delta, r, d, z);
        NSB_LAUNCH_CHECK(ctx, "k_cheb_step");
        rho_old = rho;
    }
    return 0;
}

static int check_pc(nsb_ctx* ctx, const nsb_precond* M) {
    NSB_REQUIRE(ctx, M != nullptr, "null preconditioner");
    NSB_REQUIRE(ctx, M->kind == NSB_PC_NONE || M->kind == NSB_PC_JACOBI || M->kind == NSB_PC_CHEBYSHEV ||
                M->kind == NSB_PC_SADDLE, "unknown preconditioner kind");
    if (M->kind == NSB_PC_SADDLE) {
        const nsb_saddle* S = M->saddle;
        NSB_REQUIRE(ctx, S && S->work && S->Mp_inv_diag && S->Lp_inv_diag && S->n_u > 0 && S->n_p > 0,
                    "saddle preconditioner: incomplete description");
        NSB_REQUIRE(ctx, S->vel_degree >= 1 && S->vel_degree <= 64 && S->vel_lambda_min > 0.0 &&
                    S->vel_lambda_max > S->vel_lambda_min, "saddle preconditioner: bad velocity Chebyshev interval");
    }
    NSB_REQUIRE(ctx, M->kind == NSB_PC_NONE || M->inv_diag != nullptr, "preconditioner needs the inverse diagonal");
    if (M->kind == NSB_PC_CHEBYSHEV)
        NSB_REQUIRE(ctx, M->degree >= 1 && M->degree <= 64 && M->lambda_min > 0.0 && M->lambda_max > M->lambda_min,
                    "Chebyshev needs 1 <= degree <= 64 and 0 < lambda_min < lambda_max");
    return 0;
}

// ---- largest eigenvalue of D^-1 A by power iteration (upper end of the Chebyshev interval) ----
__global__ void k_pi_init(int64_t n, double* __restrict__ v) {
    GRID_STRIDE(i, n) {
        uint32_t h = (uint32_t)i * 2654435761u;
        h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        v[i] = 0.5 + (double)(h & 0xffffu) / 65536.0;
    }
}
// v = (D^-1 q) * (1/sqrt(prev)) ; sum v^2
__global__ void __launch_bounds__(256)
k_pi_step(int64_t n, const double* __restrict__ dinv, const double* __restrict__ q, const double* prev2,
          double* __restrict__ v, double* partials, unsigned int* counter, double* out) {
    const double sc = prev2 ? 1.0 / sqrt(*prev2) : 1.0;
    double s[1] = {0.0};
    GRID_STRIDE(i, n) {
        const double vi = dinv[i] * q[i] * sc;
        v[i] = vi;
        s[0] += vi * vi;
    }
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

extern "C" int nsb_lambda_max(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t iters,
                              int32_t warm_start, double* work, double* lambda_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && inv_diag && work && lambda_out && iters >= 1, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = (int64_t)A->nrows * A->bs;
    const int G = krylov_grid(ctx, n);
    double *v = work, *q = work + n;
    double* ds = ctx->d_scalars;
    if (!warm_start) {        // warm start: work[0..n) still holds the iterate a previous call left there
        k_pi_init<<<G, 256, 0, st>>>(n, v);
        NSB_LAUNCH_CHECK(ctx, "k_pi_init");
    }
    // slots S_H1 / S_H1+1 alternate: ||v_k||^2 with v_k = D^-1 A v_{k-1} / ||v_{k-1}||
    for (int k = 0; k <= iters; ++k) {
        if (nsb_launch_spmv(ctx, A, v, q, nullptr, nullptr, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_pi_step<<<G, 256, 0, st>>>(n, inv_diag, q, k == 0 ? nullptr : ds + S_H1 + ((k - 1) & 1), v, ctx->d_partials,
                                         ctx->d_counter, ds + S_H1 + (k & 1));
        }
        NSB_LAUNCH_CHECK(ctx, "k_pi_step");
        if (nsb_allreduce_sum(ctx, ds + S_H1 + (k & 1), 1, st)) return -1;
    }
    // ||v_iters||^2 = || D^-1 A v_{iters-1} ||^2 / ||v_{iters-1}||^2   (the k = 0 step is unnormalised)
    if (fetch(ctx, S_H1, 2, st)) return -1;
    *lambda_out = sqrt(ctx->h_scalars[S_H1 + (iters & 1)]);
    return 0;
}

// ---- the same on the single-precision block-scalar operator the polynomial preconditioner is built from ------
// The Chebyshev iteration of apply_cheb32 runs on S (x) I (nsb_matrix.diag32), so the interval it needs is the
// spectrum of D^-1 (S (x) I), and a power iteration with the 0.27 ms f32 SpMV costs a fifth of the FP64 one.
// work: 2 x PAD32(bs) x nrows floats (<= 2 n doubles); the iterate stays there for warm starts.
template <int BS>
__global__ void k_pi32_init(int64_t nnodes, float* __restrict__ x) {
    GRID_STRIDE(i, nnodes) {
        float c[BS];
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            uint32_t h = (uint32_t)(i * BS + b) * 2654435761u;
            h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
            c[b] = 0.5f + (float)(h & 0xffffu) / 65536.0f;
        }
        Pad32<BS>::st(x, i, c);
    }
}
// x = (D^-1 q) / sqrt(prev) ; out = sum x^2 (accumulated in FP64)
template <int BS>
__global__ void __launch_bounds__(256)
k_pi32_step(int64_t nnodes, const double* __restrict__ dinv, const float* __restrict__ q, const double* prev2,
            float* __restrict__ x, double* partials, unsigned int* counter, double* out) {
    const double sc = prev2 ? 1.0 / sqrt(*prev2) : 1.0;
    double s[1] = {0.0};
    GRID_STRIDE(i, nnodes) {
        float qi[BS], c[BS];
        Pad32<BS>::ld(q, i, qi);
#pragma unroll
        for (int b = 0; b < BS; ++b) {
            const double vi = dinv[i * BS + b] * (double)qi[b] * sc;
            c[b] = (float)vi;
            s[0] += vi * vi;
        }
        Pad32<BS>::st(x, i, c);
    }
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

template <int BS>
static int lambda_max32(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t iters, int32_t warm_start,
                        double* work, double* lambda_out, cudaStream_t st) {
    constexpr int PB = BS == 3 ? 4 : BS;
    const int64_t nn = A->nrows;
    const int G = krylov_grid(ctx, nn);
    float* x = reinterpret_cast<float*>(((uintptr_t)work + 15) & ~(uintptr_t)15);
    float* q = x + PB * nn;
    double* ds = ctx->d_scalars;
    if (!warm_start) {
        k_pi32_init<BS><<<G, 256, 0, st>>>(nn, x);
        NSB_LAUNCH_CHECK(ctx, "k_pi32_init");
    }
    for (int k = 0; k <= iters; ++k) {
        if (nsb_launch_spmv32(ctx, A, x, q, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_pi32_step<BS><<<G, 256, 0, st>>>(nn, inv_diag, q, k == 0 ? nullptr : ds + S_H1 + ((k - 1) & 1), x, ctx->d_partials,
                                               ctx->d_counter, ds + S_H1 + (k & 1));
        }
        NSB_LAUNCH_CHECK(ctx, "k_pi32_step");
        if (nsb_allreduce_sum(ctx, ds + S_H1 + (k & 1), 1, st)) return -1;
    }
    if (fetch(ctx, S_H1, 2, st)) return -1;
    *lambda_out = sqrt(ctx->h_scalars[S_H1 + (iters & 1)]);
    return 0;
}

extern "C" int nsb_lambda_max32(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t iters,
                                int32_t warm_start, double* work, double* lambda_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && inv_diag && work && lambda_out && iters >= 1, "bad arguments");
    NSB_REQUIRE(ctx, A->diag32 && nsb_spmv32_available(A), "nsb_lambda_max32 needs the block-scalar single-precision copy (diag32)");
    cudaStream_t st = (cudaStream_t)stream;
    return A->bs == 3 ? lambda_max32<3>(ctx, A, inv_diag, iters, warm_start, work, lambda_out, st)
                      : lambda_max32<2>(ctx, A, inv_diag, iters, warm_start, work, lambda_out, st);
}

// ---------------------------------------------------------------------------------------------
// CG
// ---------------------------------------------------------------------------------------------
static int ensure_cg_ring(nsb_ctx* ctx) {
    if (ctx->cg_ring) return 0;
    NSB_CUDA(ctx, cudaMallocHost(&ctx->cg_ring, sizeof(double) * NSB_CG_RING));
    for (int i = 0; i < NSB_CG_RING; ++i) NSB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->cg_ev[i], cudaEventDisableTiming));
    return 0;
}

extern "C" int64_t nsb_cg_workspace_doubles(int64_t n) { return 8 * n + 16; }   // r, z, p, q + 4 n of preconditioner scratch

extern "C" int nsb_cg(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* b, double* x,
                      double rtol, double atol, int32_t maxit, int32_t nullspace, double* work,
                      int32_t* iters_out, double* resid_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && b && x && work, "null argument");
    if (check_pc(ctx, M)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    // scalar systems without halo exchanges: the whole solve is one persistent cooperative kernel
    if (!ctx->no_persistent_cg) {
        const int pst = nsb_cg_persistent(ctx, A, M, b, x, rtol, atol, maxit, nullspace, work, iters_out, resid_out, st);
        if (pst <= 0) return pst;
    }
    const int64_t n = (int64_t)A->nrows * A->bs;
    double *r = work, *z = work + n, *p = work + 2 * n, *q = work + 3 * n, *scratch = work + 4 * n;
    double* ds = ctx->d_scalars;
    const int G = krylov_grid(ctx, n);
    const bool jac = M->kind == NSB_PC_JACOBI;
    const double* dinv = M->inv_diag;
    double shift = 0.0;
    // size of the system over all ranks (the zero-mean projection divides by it)
    const double n_all = (ctx->nranks > 1 && A->halo) ? (double)ctx->halo[A->halo].n_global * A->bs : (double)n;
    if (nullspace) {
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_reduce<false><<<G, 256, 0, st>>>(n, b, nullptr, ctx->d_partials, ctx->d_counter, ds + S_OUT);
        }
        NSB_LAUNCH_CHECK(ctx, "k_reduce");
        if (nsb_allreduce_sum(ctx, ds + S_OUT, 1, st)) return -1;
        if (fetch(ctx, S_OUT, 1, st)) return -1;
        shift = ctx->h_scalars[S_OUT] / n_all;
    }
    if (nsb_launch_spmv(ctx, A, x, q, nullptr, nullptr, st)) return -1;
    int cur = S_RZ0, nxt = S_RZ1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        if (jac) k_cg_init<true><<<G, 256, 0, st>>>(n, b, q, dinv, shift, r, z, p, ctx->d_partials, ctx->d_counter, ds);
        else k_cg_init<false><<<G, 256, 0, st>>>(n, b, q, dinv, shift, r, z, p, ctx->d_partials, ctx->d_counter, ds);
    }
    if (nsb_allreduce_sum(ctx, ds + S_TMP, 3, st)) return -1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cg_init_finish<<<1, 1, 0, st>>>(ds, jac ? 1 : 0);
    }
    NSB_LAUNCH_CHECK(ctx, "k_cg_init");
    NSB_COUNT(ctx, 1);
    if (!jac) {
        if (apply_pc(ctx, A, M, r, z, scratch, n, st, pc_single_precision(A, M))) return -1;
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_reduce<true><<<G, 256, 0, st>>>(n, r, z, ctx->d_partials, ctx->d_counter, ds + cur);
        if (nsb_allreduce_sum(ctx, ds + cur, 1, st)) return -1;
        k_cg_update2<<<G, 256, 0, st>>>(n, z, p, cur, cur, 1, ds);
        NSB_LAUNCH_CHECK(ctx, "k_cg_init(pc)");
        NSB_COUNT(ctx, 1);
    }
    if (fetch(ctx, S_RZ0, 5, st)) return -1;
    const double bnorm = sqrt(ctx->h_scalars[S_BB]);
    double rnorm = sqrt(ctx->h_scalars[S_RR]);
    const double tol = fmax(rtol * bnorm, atol);
    int it = 0;
    // Convergence look-ahead: on small systems (the pressure problem: SURVEY.md H10) an iteration is a few
    // 10-us kernels and a blocking read of ||r|| per iteration leaves the GPU idle for as long again.  The
    // norm of iteration k is therefore copied to a pinned ring and examined LAG iterations later, while the
    // next ones are already queued; the loop ends with a blocking read of the newest norm, so the exit
    // criterion ||r|| <= tol is unchanged (at most LAG extra iterations are spent).
    // (pressure-sized systems only: past the tolerance the extra iterations still move the solution in its last digits,
    // and the velocity mass solve -- 1.2 M rows per rank at C5 on 8 GPUs -- must stop at the same iteration for every
    // number of ranks, or the fields of an 8-GPU run differ from the 1-GPU run in the 8th digit)
    const int lag = (n < (int64_t)500000 && maxit > 16) ? NSB_CG_LAG : 0;
    if (lag && ensure_cg_ring(ctx)) return -1;
    while (rnorm > tol && it < maxit) {
        if (nsb_launch_spmv(ctx, A, p, q, p, ds + S_PQ, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            if (jac) k_cg_update1<true><<<G, 256, 0, st>>>(n, p, q, dinv, x, r, z, cur, ctx->d_partials, ctx->d_counter, ds);
            else k_cg_update1<false><<<G, 256, 0, st>>>(n, p, q, dinv, x, r, z, cur, ctx->d_partials, ctx->d_counter, ds);
        }
        if (nsb_allreduce_sum(ctx, ds + S_TMP, 2, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_cg_update1_finish<<<1, 1, 0, st>>>(ds, nxt, jac ? 1 : 0);
        }
        NSB_LAUNCH_CHECK(ctx, "k_cg_update1");
        NSB_COUNT(ctx, 1);
        if (!jac) {
            if (apply_pc(ctx, A, M, r, z, scratch, n, st, pc_single_precision(A, M))) return -1;
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_reduce<true><<<G, 256, 0, st>>>(n, r, z, ctx->d_partials, ctx->d_counter, ds + nxt);
            NSB_LAUNCH_CHECK(ctx, "k_reduce(rz)");
        }
        if (!jac && nsb_allreduce_sum(ctx, ds + nxt, 1, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_cg_update2<<<G, 256, 0, st>>>(n, z, p, cur, nxt, 0, ds);
        }
        NSB_LAUNCH_CHECK(ctx, "k_cg_update2");
        const int t = cur; cur = nxt; nxt = t;
        ++it;
        if (lag == 0) {
            if (fetch(ctx, S_RR, 1, st)) return -1;
            rnorm = sqrt(ctx->h_scalars[S_RR]);
        } else {
            const int slot = (it - 1) % NSB_CG_RING;
            NSB_CUDA(ctx, cudaMemcpyAsync(ctx->cg_ring + slot, ds + S_RR, sizeof(double), cudaMemcpyDeviceToHost, st));
            NSB_CUDA(ctx, cudaEventRecord(ctx->cg_ev[slot], st));
            if (it > lag || it >= maxit) {
                const int old = it >= maxit ? slot : (it - 1 - lag) % NSB_CG_RING;
                NSB_CUDA(ctx, cudaEventSynchronize(ctx->cg_ev[old]));
                double r_old = sqrt(ctx->cg_ring[old]);
                if (r_old <= tol && old != slot) {      // looks converged: confirm on the newest iterate
                    NSB_CUDA(ctx, cudaEventSynchronize(ctx->cg_ev[slot]));
                    r_old = sqrt(ctx->cg_ring[slot]);
                }
                rnorm = r_old;
            }
        }
        if (!(rnorm == rnorm)) return nsb_fail(ctx, "CG breakdown: residual is NaN%s%s");
    }
    if (nullspace) {
        k_reduce<false><<<G, 256, 0, st>>>(n, x, nullptr, ctx->d_partials, ctx->d_counter, ds + S_OUT);
        if (nsb_allreduce_sum(ctx, ds + S_OUT, 1, st)) return -1;
        if (fetch(ctx, S_OUT, 1, st)) return -1;
        k_shift<<<G, 256, 0, st>>>(n, ctx->h_scalars[S_OUT] / n_all, x);
        NSB_LAUNCH_CHECK(ctx, "k_shift");
        NSB_COUNT(ctx, 1);
    }
    if (iters_out) *iters_out = it;
    if (resid_out) *resid_out = rnorm;
    if (rnorm > tol) return nsb_fail(ctx, "CG did not converge within the iteration limit%s%s");
    return 0;
}

// ---------------------------------------------------------------------------------------------
// GMRES kernels
// ---------------------------------------------------------------------------------------------
// r = b - q ; sum r.r
__global__ void __launch_bounds__(256)
k_residual(int64_t n, const double* __restrict__ b, const double* __restrict__ q, double* __restrict__ r,
           double* partials, unsigned int* counter, double* out) {
    double s[1] = {0.0};
    GRID_STRIDE(i, n) {
        const double ri = b[i] - q[i];
        r[i] = ri;
        s[0] += ri * ri;
    }
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

// v = w / sqrt(*nrm2)
__global__ void __launch_bounds__(256)
k_normalize(int64_t n, const double* __restrict__ w, const double* __restrict__ nrm2, double* __restrict__ v) {
    const double inv = 1.0 / sqrt(*nrm2);
    GRID_STRIDE(i, n) v[i] = w[i] * inv;
}

// h[k] = V_k . w for k in [k0, k0+nk), NV accumulators per thread: one pass over w per NV vectors
template <int NV>
__global__ void __launch_bounds__(256)
k_multi_dot(int64_t n, int k0, int nk, const double* __restrict__ V, const double* __restrict__ w,
            double* partials, unsigned int* counter, double* out) {
    double s[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) s[k] = 0.0;
    GRID_STRIDE(i, n) {
        const double wi = w[i];
#pragma unroll
        for (int k = 0; k < NV; ++k)
            if (k < nk) s[k] += V[(int64_t)(k0 + k) * n + i] * wi;
    }
    __shared__ double red[NV * 8];
    grid_sum<NV, 256>(s, partials, counter, out + k0, red);
}

// w -= sum_k h[k] V_k ; sum w.w
__global__ void __launch_bounds__(256)
k_multi_axpy(int64_t n, int nk, const double* __restrict__ V, const double* __restrict__ h,
             double* __restrict__ w, double* partials, unsigned int* counter, double* out) {
    double s[1] = {0.0};
    GRID_STRIDE(i, n) {
        double wi = w[i];
        for (int k = 0; k < nk; ++k) wi -= h[k] * V[(int64_t)k * n + i];
        w[i] = wi;
        s[0] += wi * wi;
    }
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

// CGS-2, middle pass in one sweep over V (nk <= NV):  w -= sum_k h[k] V_k  and, with the V_k entries still
// in registers, h2[k] = V_k . w  of the UPDATED w (the second Gram-Schmidt projection)
template <int NV>
__global__ void __launch_bounds__(256)
k_multi_axpy_dot(int64_t n, int nk, const double* __restrict__ V, const double* __restrict__ h,
                 double* __restrict__ w, double* partials, unsigned int* counter, double* out) {
    double s[NV], hk[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) { s[k] = 0.0; hk[k] = k < nk ? h[k] : 0.0; }
    GRID_STRIDE(i, n) {
        double v[NV];
        double wi = w[i];
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            v[k] = k < nk ? V[(int64_t)k * n + i] : 0.0;
            wi -= hk[k] * v[k];
        }
        w[i] = wi;
#pragma unroll
        for (int k = 0; k < NV; ++k) s[k] += v[k] * wi;
    }
    __shared__ double red[NV * 8];
    grid_sum<NV, 256>(s, partials, counter, out, red);
}

// u = sum_k y[k] V_k
__global__ void __launch_bounds__(256)
k_combine(int64_t n, int nk, const double* __restrict__ V, const double* __restrict__ y, double* __restrict__ u) {
    GRID_STRIDE(i, n) {
        double s = 0.0;
        for (int k = 0; k < nk; ++k) s += y[k] * V[(int64_t)k * n + i];
        u[i] = s;
    }
}

__global__ void __launch_bounds__(256)
k_add(int64_t n, const double* __restrict__ t, double* __restrict__ x) {
    GRID_STRIDE(i, n) x[i] += t[i];
}

// ---------------------------------------------------------------------------------------------
// Successive right-hand sides (Fischer, Comput. Methods Appl. Mech. Engrg. 163, 1998): the pressure Poisson
// problem of source/ns_ipcs_solver.py:149-171 is solved with the SAME operator at every time step.  X holds up to
// lmax A-orthonormal vectors spanning the previous solutions, AX their images.  nsb_rhs_guess gives the best
// approximation of A^-1 b in span(X) in the energy norm, x = sum_k (X_k . b) X_k, as the initial guess of the CG
// (the solve itself still runs to its tolerance: only the iteration count changes); nsb_rhs_update adds the part
// of the new solution that is A-orthogonal to the basis (one SpMV, one Gram-Schmidt pass) and, when the basis is
// full, drops its older half.
// ---------------------------------------------------------------------------------------------
#define NSB_RHS_MAX_BASIS 16

extern "C" int nsb_rhs_guess(nsb_ctx* ctx, int64_t n, int32_t l, const double* X, const double* b, double* x,
                             void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, n > 0 && l >= 0 && l <= NSB_RHS_MAX_BASIS && X && b && x, "bad arguments");
    if (l == 0) return 0;                                   // no history: the caller's x stays the guess
    cudaStream_t st = (cudaStream_t)stream;
    const int G = krylov_grid(ctx, n);
    double* ds = ctx->d_scalars;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_multi_dot<16><<<G, 256, 0, st>>>(n, 0, l, X, b, ctx->d_partials, ctx->d_counter, ds + S_H1);
    }
    NSB_LAUNCH_CHECK(ctx, "nsb_rhs_guess dots");
    if (nsb_allreduce_sum(ctx, ds + S_H1, l, st)) return -1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_combine<<<G, 256, 0, st>>>(n, l, X, ds + S_H1, x);
    }
    NSB_LAUNCH_CHECK(ctx, "nsb_rhs_guess combine");
    NSB_COUNT(ctx, 2);
    return 0;
}

extern "C" int nsb_rhs_update(nsb_ctx* ctx, const nsb_matrix* A, int32_t* l_inout, int32_t lmax, double* X, double* AX,
                              const double* x, double* work, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && l_inout && X && AX && x && work && lmax >= 1 && lmax <= NSB_RHS_MAX_BASIS &&
                *l_inout >= 0 && *l_inout <= lmax, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = (int64_t)A->nrows * A->bs;
    const int G = krylov_grid(ctx, n);
    double* ds = ctx->d_scalars;
    int l = *l_inout;
    if (l == lmax) {
        // full: drop the OLDER half.  A subset of an A-orthonormal set is A-orthonormal, so the newer half moves to
        // the front as it is (lmax == 1: plain restart from the current solution)
        const int keep = lmax / 2;
        for (int k = 0; k < keep; ++k) {
            NSB_CUDA(ctx, cudaMemcpyAsync(X + (int64_t)k * n, X + (int64_t)(lmax - keep + k) * n, sizeof(double) * n,
                                          cudaMemcpyDeviceToDevice, st));
            NSB_CUDA(ctx, cudaMemcpyAsync(AX + (int64_t)k * n, AX + (int64_t)(lmax - keep + k) * n, sizeof(double) * n,
                                          cudaMemcpyDeviceToDevice, st));
        }
        l = keep;
    }
    double *w = work, *aw = work + n;
    NSB_CUDA(ctx, cudaMemcpyAsync(w, x, sizeof(double) * n, cudaMemcpyDeviceToDevice, st));
    if (nsb_launch_spmv(ctx, A, w, aw, nullptr, nullptr, st)) return -1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_reduce<true><<<G, 256, 0, st>>>(n, w, aw, ctx->d_partials, ctx->d_counter, ds + S_TMP);      // x . A x
    }
    if (nsb_allreduce_sum(ctx, ds + S_TMP, 1, st)) return -1;
    if (l > 0) {
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_multi_dot<16><<<G, 256, 0, st>>>(n, 0, l, AX, w, ctx->d_partials, ctx->d_counter, ds + S_H1);
        }
        if (nsb_allreduce_sum(ctx, ds + S_H1, l, st)) return -1;
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_multi_axpy<<<G, 256, 0, st>>>(n, l, X, ds + S_H1, w, ctx->d_partials, ctx->d_counter, ds + S_NRM);
        k_multi_axpy<<<G, 256, 0, st>>>(n, l, AX, ds + S_H1, aw, ctx->d_partials, ctx->d_counter, ds + S_NRM);
    }
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_reduce<true><<<G, 256, 0, st>>>(n, w, aw, ctx->d_partials, ctx->d_counter, ds + S_OUT);      // delta . A delta
    }
    NSB_LAUNCH_CHECK(ctx, "nsb_rhs_update");
    if (nsb_allreduce_sum(ctx, ds + S_OUT, 1, st)) return -1;
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars + S_OUT, ds + S_OUT, sizeof(double), cudaMemcpyDeviceToHost, st));
    if (fetch(ctx, S_TMP, 1, st)) return -1;
    const double s = ctx->h_scalars[S_OUT], xAx = ctx->h_scalars[S_TMP];
    NSB_COUNT(ctx, 6);
    // nothing new (the guess was already the solution), or a direction lost to round-off: keep the basis as it is
    if (!(s > 1e-24 * xAx) || !(xAx > 0.0)) { *l_inout = l; return 0; }
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_normalize<<<G, 256, 0, st>>>(n, w, ds + S_OUT, X + (int64_t)l * n);
        k_normalize<<<G, 256, 0, st>>>(n, aw, ds + S_OUT, AX + (int64_t)l * n);
    }
    NSB_LAUNCH_CHECK(ctx, "nsb_rhs_update normalize");
    *l_inout = l + 1;
    return 0;
}

#define NSB_GMRES_MAX_RESTART 60

// V (restart+1), w, t, Chebyshev scratch (3), and restart more for the preconditioned basis Z of the
// flexible variant (M->lowp)
extern "C" int64_t nsb_gmres_workspace_doubles(int64_t n, int32_t restart, int32_t flexible) {
    return (int64_t)(restart + 7 + (flexible ? restart : 0)) * n;
}

extern "C" int nsb_gmres(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* b, double* x,
                         double rtol, double atol, int32_t maxit, int32_t restart, int32_t x0_is_zero, double* work,
                         int32_t* iters_out, double* resid_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && b && x && work, "null argument");
    NSB_REQUIRE(ctx, restart >= 1 && restart <= NSB_GMRES_MAX_RESTART, "restart must be in [1, 60]");
    if (check_pc(ctx, M)) return -1;
    NSB_REQUIRE(ctx, M->kind != NSB_PC_SADDLE || ((int64_t)M->saddle->n_u + M->saddle->n_p == (int64_t)A->nrows * A->bs &&
                                                   A->kind == NSB_CSR && A->bs == 1 && ctx->nranks == 1),
                "saddle preconditioner needs the scalar CSR mixed matrix on one GPU");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = (int64_t)A->nrows * A->bs;
    const int m = restart;
    double* V = work;                       // (m+1) x n
    double* w = work + (int64_t)(m + 1) * n;
    double* t = w + n;
    double* scratch = t + n;                // 4 n (Chebyshev: 3 n of recurrence vectors + the bf16 shadows of the f32 variant)
    // Flexible GMRES (Saad 1993) when the preconditioner runs in single precision: z_j = M^-1 v_j is kept
    // (Z, restart x n behind the scratch) and x += Z y, so the preconditioner need not be a fixed linear
    // operator; Arnoldi, the Hessenberg system and the residual test stay FP64 on the FP64 matrix.
    const bool flex = pc_single_precision(A, M) || M->kind == NSB_PC_SADDLE;
    double* Z = scratch + 4 * n;
    double* ds = ctx->d_scalars;
    double* hs = ctx->h_scalars;
    const int G = krylov_grid(ctx, n);
    std::vector<double> H((size_t)(m + 1) * m, 0.0), cs(m), sn(m), g(m + 1), y(m);

    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_reduce<true><<<G, 256, 0, st>>>(n, b, b, ctx->d_partials, ctx->d_counter, ds + S_OUT);
    }
    NSB_LAUNCH_CHECK(ctx, "k_reduce");
    if (nsb_allreduce_sum(ctx, ds + S_OUT, 1, st)) return -1;
    if (fetch(ctx, S_OUT, 1, st)) return -1;
    const double bnorm = sqrt(hs[S_OUT]);
    const double tol = fmax(rtol * bnorm, atol);
    int it = 0;
    double rnorm = 0.0;
    bool first_cycle = true;
    while (true) {
        const bool r_is_b = first_cycle && x0_is_zero;       // x = 0: r = b and ||r|| = ||b||, no SpMV needed
        first_cycle = false;
        if (r_is_b) {
            rnorm = bnorm;
        } else {
            // true residual r = b - A x -> w
            if (nsb_launch_spmv(ctx, A, x, t, nullptr, nullptr, st)) return -1;
            {
                NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                k_residual<<<G, 256, 0, st>>>(n, b, t, w, ctx->d_partials, ctx->d_counter, ds + S_NRM);
            }
            NSB_LAUNCH_CHECK(ctx, "k_residual");
            if (nsb_allreduce_sum(ctx, ds + S_NRM, 1, st)) return -1;
            if (fetch(ctx, S_NRM, 1, st)) return -1;
            rnorm = sqrt(hs[S_NRM]);
        }
        if (!(rnorm == rnorm)) return nsb_fail(ctx, "GMRES breakdown: residual is NaN%s%s");
        if (rnorm <= tol || it >= maxit) break;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_normalize<<<G, 256, 0, st>>>(n, r_is_b ? b : w, ds + (r_is_b ? S_OUT : S_NRM), V);
        }
        NSB_LAUNCH_CHECK(ctx, "k_normalize");
        g.assign(m + 1, 0.0);
        g[0] = rnorm;
        int j = 0;
        for (; j < m && it < maxit; ++j) {
            // w = A M^-1 v_j
            double* zj = flex ? Z + (int64_t)j * n : t;
            if (apply_pc(ctx, A, M, V + (int64_t)j * n, zj, scratch, n, st, flex)) return -1;
            if (nsb_launch_spmv(ctx, A, zj, w, nullptr, nullptr, st)) return -1;
            const int nk = j + 1;
            // CGS-2.  Up to 16 basis vectors: three sweeps over V (dot | axpy + second dot | axpy + norm)
            // instead of four, and one all-reduce less.
            const bool fused = nk <= 16;
            if (fused) {
                {
                    NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                    if (nk <= 4) k_multi_dot<4><<<G, 256, 0, st>>>(n, 0, nk, V, w, ctx->d_partials, ctx->d_counter, ds + S_H1);
                    else if (nk <= 8) k_multi_dot<8><<<G, 256, 0, st>>>(n, 0, nk, V, w, ctx->d_partials, ctx->d_counter, ds + S_H1);
                    else k_multi_dot<16><<<G, 256, 0, st>>>(n, 0, nk, V, w, ctx->d_partials, ctx->d_counter, ds + S_H1);
                }
                if (nsb_allreduce_sum(ctx, ds + S_H1, nk, st)) return -1;
                {
                    NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                    if (nk <= 4) k_multi_axpy_dot<4><<<G, 256, 0, st>>>(n, nk, V, ds + S_H1, w, ctx->d_partials, ctx->d_counter, ds + S_H2);
                    else if (nk <= 8) k_multi_axpy_dot<8><<<G, 256, 0, st>>>(n, nk, V, ds + S_H1, w, ctx->d_partials, ctx->d_counter, ds + S_H2);
                    else k_multi_axpy_dot<16><<<G, 256, 0, st>>>(n, nk, V, ds + S_H1, w, ctx->d_partials, ctx->d_counter, ds + S_H2);
                }
                if (nsb_allreduce_sum(ctx, ds + S_H2, nk, st)) return -1;
                {
                    NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                    k_multi_axpy<<<G, 256, 0, st>>>(n, nk, V, ds + S_H2, w, ctx->d_partials, ctx->d_counter, ds + S_NRM);
                }
                if (nsb_allreduce_sum(ctx, ds + S_NRM, 1, st)) return -1;
                NSB_COUNT(ctx, 1);
            }
            for (int pass = 0; pass < 2 && !fused; ++pass) {   // CGS-2, more than 16 vectors: chunked dots
                double* hdev = ds + (pass == 0 ? S_H1 : S_H2);
                {
                    NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                    for (int k0 = 0; k0 < nk; k0 += 16) {
                        const int c = nk - k0 < 16 ? nk - k0 : 16;
                        if (c <= 4) k_multi_dot<4><<<G, 256, 0, st>>>(n, k0, c, V, w, ctx->d_partials, ctx->d_counter, hdev);
                        else if (c <= 8) k_multi_dot<8><<<G, 256, 0, st>>>(n, k0, c, V, w, ctx->d_partials, ctx->d_counter, hdev);
                        else k_multi_dot<16><<<G, 256, 0, st>>>(n, k0, c, V, w, ctx->d_partials, ctx->d_counter, hdev);
                        NSB_COUNT(ctx, 1);
                    }
                }
                if (nsb_allreduce_sum(ctx, hdev, nk, st)) return -1;
                {
                    NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                    k_multi_axpy<<<G, 256, 0, st>>>(n, nk, V, hdev, w, ctx->d_partials, ctx->d_counter, ds + S_NRM);
                }
                if (nsb_allreduce_sum(ctx, ds + S_NRM, 1, st)) return -1;
            }
            {
                NSB_TIMED(ctx, NSB_T_KRYLOV, st);
                k_normalize<<<G, 256, 0, st>>>(n, w, ds + S_NRM, V + (int64_t)(j + 1) * n);
            }
            NSB_LAUNCH_CHECK(ctx, "gmres arnoldi");
            NSB_COUNT(ctx, 2);
            // one host round trip per iteration: h1, h2, ||w||^2
            NSB_CUDA(ctx, cudaMemcpyAsync(hs + S_NRM, ds + S_NRM, sizeof(double) * (S_H2 + nk + 8 - S_NRM),
                                          cudaMemcpyDeviceToHost, st));
            NSB_CUDA(ctx, cudaStreamSynchronize(st));
            double* Hj = &H[(size_t)j * (m + 1)];
            for (int k = 0; k < nk; ++k) Hj[k] = hs[S_H1 + k] + hs[S_H2 + k];
            Hj[nk] = sqrt(hs[S_NRM]);
            for (int k = 0; k < j; ++k) {   // previous rotations
                const double a = cs[k] * Hj[k] + sn[k] * Hj[k + 1];
                Hj[k + 1] = -sn[k] * Hj[k] + cs[k] * Hj[k + 1];
                Hj[k] = a;
            }
            const double d = hypot(Hj[j], Hj[j + 1]);
            if (d == 0.0) return nsb_fail(ctx, "GMRES breakdown: zero Hessenberg column%s%s");
            cs[j] = Hj[j] / d; sn[j] = Hj[j + 1] / d;
            Hj[j] = d; Hj[j + 1] = 0.0;
            g[j + 1] = -sn[j] * g[j];
            g[j] = cs[j] * g[j];
            ++it;
            rnorm = fabs(g[j + 1]);
            if (!(rnorm == rnorm)) return nsb_fail(ctx, "GMRES breakdown: NaN in the Arnoldi process%s%s");
            if (rnorm <= tol) { ++j; break; }
        }
        // back substitution, x += M^-1 (V y)   [flexible: x += Z y]
        const int nk = j;
        for (int i = nk - 1; i >= 0; --i) {
            double s = g[i];
            for (int k = i + 1; k < nk; ++k) s -= H[(size_t)k * (m + 1) + i] * y[k];
            y[i] = s / H[(size_t)i * (m + 1) + i];
        }
        for (int k = 0; k < nk; ++k) hs[S_Y + k] = y[k];
        NSB_CUDA(ctx, cudaMemcpyAsync(ds + S_Y, hs + S_Y, sizeof(double) * nk, cudaMemcpyHostToDevice, st));
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_combine<<<G, 256, 0, st>>>(n, nk, flex ? Z : V, ds + S_Y, flex ? t : w);
        }
        NSB_LAUNCH_CHECK(ctx, "k_combine");
        if (!flex && apply_pc(ctx, A, M, w, t, scratch, n, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_add<<<G, 256, 0, st>>>(n, t, x);
        }
        NSB_LAUNCH_CHECK(ctx, "k_add");
        NSB_CUDA(ctx, cudaStreamSynchronize(st));   // hs[S_Y..] is reused next cycle
        // Converged by the Arnoldi residual norm |g_{j+1}|: done, as library GMRES implementations (PETSc KSPGMRES)
        // do; the reference itself solves with sparse LU -- with CGS-2 the recurrence tracks ||b - A x|| to round-off, and in the flexible variant
        // w = A z_j makes it exact by construction.  Only a restart recomputes the true residual.
        if (rnorm <= tol) break;
    }
    if (iters_out) *iters_out = it;
    if (resid_out) *resid_out = rnorm;
    if (rnorm > tol) return nsb_fail(ctx, "GMRES did not converge within the iteration limit%s%s");
    return 0;
}

// ---------------------------------------------------------------------------------------------
// K16: block-triangular preconditioner of the mixed Newton system (include/nsb200.h, nsb_saddle)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
k_mask_copy(int64_t n, const double* __restrict__ src, const uint8_t* __restrict__ m1, const uint8_t* __restrict__ m2,
            double* __restrict__ dst) {
    GRID_STRIDE(i, n) dst[i] = ((m1 && m1[i]) || (m2 && m2[i])) ? 0.0 : src[i];
}
// z_p = -(cm t1 + cl t2); identity rows of A: z_p = v_p
__global__ void __launch_bounds__(256)
k_schur_combine(int64_t n, const double* __restrict__ t1, const double* __restrict__ t2, double cm, double cl,
                const double* __restrict__ vp, const uint8_t* __restrict__ ident, double* __restrict__ zp) {
    GRID_STRIDE(i, n) zp[i] = (ident && ident[i]) ? vp[i] : -(cm * t1[i] + (t2 ? cl * t2[i] : 0.0));
}
// t = v - t
__global__ void __launch_bounds__(256)
k_rsub(int64_t n, const double* __restrict__ v, double* __restrict__ t) {
    GRID_STRIDE(i, n) t[i] = v[i] - t[i];
}

static int saddle_apply(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* v, double* z,
                        double* scratch, cudaStream_t st) {
    nsb_saddle* S = const_cast<nsb_saddle*>(M->saddle);
    const int64_t nu = S->n_u, np_ = S->n_p, n = nu + np_;
    double* t = S->work;                       // [n]   B z_p, then the velocity right-hand side
    double* vm = t + n;                        // [n_p] masked copy of v_p
    double* t1 = vm + np_;                     // [n_p] Mp^-1 v_p
    double* t2 = t1 + np_;                     // [n_p] Lp^-1 v_p
    double* cgw = t2 + np_;                    // [7 n_p]
    const double* vp = v + nu;
    const int Gp = krylov_grid(ctx, np_), Gu = krylov_grid(ctx, nu);
    int32_t its = 0;
    // ---- pressure: z_p = -(c_mass Mp^-1 + c_lap Lp^-1) v_p -------------------------------------------
    nsb_precond jac;
    memset(&jac, 0, sizeof jac);
    jac.kind = NSB_PC_JACOBI; jac.degree = 1; jac.inv_diag = S->Mp_inv_diag;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_mask_copy<<<Gp, 256, 0, st>>>(np_, vp, S->Mp.bc_mask, nullptr, vm);
        NSB_LAUNCH_CHECK(ctx, "k_mask_copy");
    }
    NSB_CUDA(ctx, cudaMemsetAsync(t1, 0, sizeof(double) * np_, st));
    if (nsb_cg(ctx, &S->Mp, &jac, vm, t1, S->inner_rtol, 0.0, S->inner_maxit, 0, cgw, &its, nullptr, st)) return -1;
    S->inner_iterations += its;
    const bool lap = S->c_lap != 0.0;
    if (lap) {
        nsb_precond pc;
        memset(&pc, 0, sizeof pc);
        pc.inv_diag = S->Lp_inv_diag;
        if (S->lp_degree >= 2) {
            pc.kind = NSB_PC_CHEBYSHEV; pc.degree = S->lp_degree; pc.lambda_min = S->lp_lambda_min; pc.lambda_max = S->lp_lambda_max;
        } else {
            pc.kind = NSB_PC_JACOBI; pc.degree = 1;
        }
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_mask_copy<<<Gp, 256, 0, st>>>(np_, vp, S->Lp.bc_mask, nullptr, vm);
            NSB_LAUNCH_CHECK(ctx, "k_mask_copy");
        }
        NSB_CUDA(ctx, cudaMemsetAsync(t2, 0, sizeof(double) * np_, st));
        if (nsb_cg(ctx, &S->Lp, &pc, vm, t2, S->inner_rtol, 0.0, S->inner_maxit, S->lp_nullspace, cgw, &its, nullptr, st)) return -1;
        S->inner_iterations += its;
    }
    NSB_CUDA(ctx, cudaMemsetAsync(z, 0, sizeof(double) * nu, st));
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_schur_combine<<<Gp, 256, 0, st>>>(np_, t1, lap ? t2 : nullptr, S->c_mass, S->c_lap, vp, S->p_identity, z + nu);
        NSB_LAUNCH_CHECK(ctx, "k_schur_combine");
    }
    // ---- velocity: z_u = Cheb(D^-1 F) D^-1 (v_u - B z_p);  (A [0; z_p])_u = B z_p ---------------------
    if (nsb_launch_spmv(ctx, A, z, t, nullptr, nullptr, st)) return -1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_rsub<<<Gu, 256, 0, st>>>(nu, v, t);
        NSB_LAUNCH_CHECK(ctx, "k_rsub");
    }
    double *r = scratch, *d = scratch + n, *q = scratch + 2 * n;
    NSB_CUDA(ctx, cudaMemsetAsync(d + nu, 0, sizeof(double) * np_, st));      // the direction's pressure part stays 0
    const double theta = 0.5 * (S->vel_lambda_max + S->vel_lambda_min), delta = 0.5 * (S->vel_lambda_max - S->vel_lambda_min);
    const double sigma1 = theta / delta;
    double rho_old = 1.0 / sigma1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cheb_first<<<Gu, 256, 0, st>>>(nu, M->inv_diag, t, 1.0 / theta, r, d, z);
        NSB_LAUNCH_CHECK(ctx, "k_cheb_first");
    }
    for (int k = 1; k < S->vel_degree; ++k) {
        const double rho = 1.0 / (2.0 * sigma1 - rho_old);
        if (nsb_launch_spmv(ctx, A, d, q, nullptr, nullptr, st)) return -1;
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cheb_step<<<Gu, 256, 0, st>>>(nu, M->inv_diag, q, rho * rho_old, 2.0 * rho / delta, r, d, z);
        NSB_LAUNCH_CHECK(ctx, "k_cheb_step");
        rho_old = rho;
    }
    return 0;
}

extern "C" int nsb_saddle_lambda_max(nsb_ctx* ctx, const nsb_matrix* A, const double* inv_diag, int32_t n_u, int32_t iters,
                                     double* work, double* lambda_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && inv_diag && work && lambda_out && iters >= 1 && A->kind == NSB_CSR && A->bs == 1 && n_u > 0 &&
                n_u <= A->nrows, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = A->nrows;
    const int G = krylov_grid(ctx, n_u);
    double *v = work, *q = work + n;
    double* ds = ctx->d_scalars;
    NSB_CUDA(ctx, cudaMemsetAsync(v, 0, sizeof(double) * n, st));
    k_pi_init<<<G, 256, 0, st>>>(n_u, v);
    NSB_LAUNCH_CHECK(ctx, "k_pi_init");
    for (int k = 0; k <= iters; ++k) {
        if (nsb_launch_spmv(ctx, A, v, q, nullptr, nullptr, st)) return -1;
        k_pi_step<<<G, 256, 0, st>>>(n_u, inv_diag, q, k == 0 ? nullptr : ds + S_H1 + ((k - 1) & 1), v, ctx->d_partials,
                                     ctx->d_counter, ds + S_H1 + (k & 1));
        NSB_LAUNCH_CHECK(ctx, "k_pi_step");
    }
    if (fetch(ctx, S_H1, 2, st)) return -1;
    *lambda_out = sqrt(ctx->h_scalars[S_H1 + (iters & 1)]);
    return 0;
}
