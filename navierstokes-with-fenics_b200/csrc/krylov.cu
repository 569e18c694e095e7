// krylov.cu -- Krylov solvers (K11) replacing the sparse direct LU of the reference
// (SURVEY.md G3): Jacobi-preconditioned CG for the SPD pressure-Poisson / mass systems
// (source/ns_ipcs_solver.py:166-171,189-196) and right-Jacobi-preconditioned restarted GMRES
// with CGS-2 Arnoldi for the velocity / mixed Newton systems (source/ns_ipcs_solver.py:138-147).
// Step lengths stay on the device (kernels read alpha/beta from device scalars); inner products
// are warp-shuffle + deterministic last-block reductions; vector updates are fused with the
// reductions that follow them and p.Ap is fused into the SpMV epilogue.
#include <math.h>
#include <vector>
#include "spmv.cuh"

// device scalar slots
enum { S_OUT = 0, S_RZ0 = 1, S_RZ1 = 2, S_PQ = 3, S_RR = 4, S_BB = 5, S_NRM = 8, S_H1 = 32, S_H2 = 96, S_Y = 160 };

static int krylov_grid(nsb_ctx* ctx, int64_t n) {
    int64_t b = (n + 255) / 256;
    const int64_t cap = (int64_t)ctx->sm_count * 8;
    if (b > cap) b = cap;
    return (int)(b < 1 ? 1 : b);
}

// ---------------------------------------------------------------------------------------------
// CG kernels
// ---------------------------------------------------------------------------------------------
// r = b - shift - q ; z = dinv r ; p = z ; sums: r.z, r.r, (b-shift).(b-shift)
__global__ void __launch_bounds__(256)
k_cg_init(int64_t n, const double* __restrict__ b, const double* __restrict__ q, const double* __restrict__ dinv,
          double shift, double* __restrict__ r, double* __restrict__ z, double* __restrict__ p,
          double* partials, unsigned int* counter, double* scal) {
    double s[3] = {0.0, 0.0, 0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double bi = b[i] - shift;
        const double ri = bi - q[i];
        const double zi = dinv[i] * ri;
        r[i] = ri; z[i] = zi; p[i] = zi;
        s[0] += ri * zi; s[1] += ri * ri; s[2] += bi * bi;
    }
    __shared__ double red[3 * 8];
    double out3[3];
    (void)out3;
    grid_sum<3, 256>(s, partials, counter, scal + S_PQ + 8, red);   // staged below
}

// the 3 sums of k_cg_init land in scal[S_PQ+8 .. +10]; move them to their slots
__global__ void k_cg_init_finish(double* scal) {
    scal[S_RZ0] = scal[S_PQ + 8];
    scal[S_RR] = scal[S_PQ + 9];
    scal[S_BB] = scal[S_PQ + 10];
}

// alpha = rz/pq ; x += alpha p ; r -= alpha q ; z = dinv r ; sums r.z -> rz_new, r.r
__global__ void __launch_bounds__(256)
k_cg_update1(int64_t n, const double* __restrict__ p, const double* __restrict__ q,
             const double* __restrict__ dinv, double* __restrict__ x, double* __restrict__ r,
             double* __restrict__ z, int slot_rz, double* partials, unsigned int* counter, double* scal) {
    const double alpha = scal[slot_rz] / scal[S_PQ];
    double s[2] = {0.0, 0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        x[i] += alpha * p[i];
        const double ri = r[i] - alpha * q[i];
        const double zi = dinv[i] * ri;
        r[i] = ri; z[i] = zi;
        s[0] += ri * zi; s[1] += ri * ri;
    }
    __shared__ double red[2 * 8];
    grid_sum<2, 256>(s, partials, counter, scal + S_PQ + 8, red);
}

__global__ void k_cg_update1_finish(double* scal, int slot_rz_new) {
    scal[slot_rz_new] = scal[S_PQ + 8];
    scal[S_RR] = scal[S_PQ + 9];
}

// beta = rz_new / rz ; p = z + beta p
__global__ void __launch_bounds__(256)
k_cg_update2(int64_t n, const double* __restrict__ z, double* __restrict__ p, int slot_rz, int slot_rz_new,
             const double* __restrict__ scal) {
    const double beta = scal[slot_rz_new] / scal[slot_rz];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        p[i] = z[i] + beta * p[i];
}

__global__ void k_shift(int64_t n, double s, double* __restrict__ x) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        x[i] -= s;
}

template <bool WITH_Y>
__global__ void __launch_bounds__(256)
k_reduce(int64_t n, const double* __restrict__ x, const double* __restrict__ y, double* partials,
         unsigned int* counter, double* out) {
    double s[1] = {0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        s[0] += WITH_Y ? x[i] * y[i] : x[i];
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

static int fetch(nsb_ctx* ctx, int slot, int count, cudaStream_t st) {
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars + slot, ctx->d_scalars + slot, sizeof(double) * count,
                                  cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    return 0;
}

extern "C" int64_t nsb_cg_workspace_doubles(int64_t n) { return 4 * n; }

extern "C" int nsb_cg(nsb_ctx* ctx, const nsb_matrix* A, const double* dinv, const double* b, double* x,
                      double rtol, double atol, int32_t maxit, int32_t nullspace, double* work,
                      int32_t* iters_out, double* resid_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && dinv && b && x && work, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = (int64_t)A->nrows * A->bs;
    double *r = work, *z = work + n, *p = work + 2 * n, *q = work + 3 * n;
    double* ds = ctx->d_scalars;
    const int G = krylov_grid(ctx, n);
    double shift = 0.0;
    if (nullspace) {
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_reduce<false><<<G, 256, 0, st>>>(n, b, nullptr, ctx->d_partials, ctx->d_counter, ds + S_OUT);
        }
        NSB_LAUNCH_CHECK(ctx, "k_reduce");
        if (fetch(ctx, S_OUT, 1, st)) return -1;
        shift = ctx->h_scalars[S_OUT] / (double)n;
    }
    if (nsb_launch_spmv(ctx, A, x, q, nullptr, nullptr, st)) return -1;
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_cg_init<<<G, 256, 0, st>>>(n, b, q, dinv, shift, r, z, p, ctx->d_partials, ctx->d_counter, ds);
        k_cg_init_finish<<<1, 1, 0, st>>>(ds);
    }
    NSB_LAUNCH_CHECK(ctx, "k_cg_init");
    NSB_COUNT(ctx, 1);
    if (fetch(ctx, S_RZ0, 5, st)) return -1;
    const double bnorm = sqrt(ctx->h_scalars[S_BB]);
    double rnorm = sqrt(ctx->h_scalars[S_RR]);
    const double tol = fmax(rtol * bnorm, atol);
    int it = 0;
    int cur = S_RZ0, nxt = S_RZ1;
    while (rnorm > tol && it < maxit) {
        if (nsb_launch_spmv(ctx, A, p, q, p, ds + S_PQ, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_cg_update1<<<G, 256, 0, st>>>(n, p, q, dinv, x, r, z, cur, ctx->d_partials, ctx->d_counter, ds);
            k_cg_update1_finish<<<1, 1, 0, st>>>(ds, nxt);
            k_cg_update2<<<G, 256, 0, st>>>(n, z, p, cur, nxt, ds);
        }
        NSB_LAUNCH_CHECK(ctx, "k_cg_update");
        NSB_COUNT(ctx, 2);
        if (fetch(ctx, S_RR, 1, st)) return -1;
        rnorm = sqrt(ctx->h_scalars[S_RR]);
        if (!(rnorm == rnorm)) return nsb_fail(ctx, "CG breakdown: residual is NaN%s%s");
        const int t = cur; cur = nxt; nxt = t;
        ++it;
    }
    if (nullspace) {
        k_reduce<false><<<G, 256, 0, st>>>(n, x, nullptr, ctx->d_partials, ctx->d_counter, ds + S_OUT);
        if (fetch(ctx, S_OUT, 1, st)) return -1;
        k_shift<<<G, 256, 0, st>>>(n, ctx->h_scalars[S_OUT] / (double)n, x);
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
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double ri = b[i] - q[i];
        r[i] = ri;
        s[0] += ri * ri;
    }
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

// v = w / sqrt(scal[slot]) ; t = dinv * v   (next Arnoldi input, right preconditioning)
__global__ void __launch_bounds__(256)
k_normalize(int64_t n, const double* __restrict__ w, const double* __restrict__ nrm2, const double* __restrict__ dinv,
            double* __restrict__ v, double* __restrict__ t) {
    const double inv = 1.0 / sqrt(*nrm2);
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const double vi = w[i] * inv;
        v[i] = vi;
        t[i] = dinv[i] * vi;
    }
}

// h[k] = V_k . w for k in [k0, k0+NV)
template <int NV>
__global__ void __launch_bounds__(256)
k_multi_dot(int64_t n, int k0, int nk, const double* __restrict__ V, const double* __restrict__ w,
            double* partials, unsigned int* counter, double* out) {
    double s[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) s[k] = 0.0;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
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
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double wi = w[i];
        for (int k = 0; k < nk; ++k) wi -= h[k] * V[(int64_t)k * n + i];
        w[i] = wi;
        s[0] += wi * wi;
    }
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

// x += dinv * sum_k y[k] V_k
__global__ void __launch_bounds__(256)
k_gmres_update(int64_t n, int nk, const double* __restrict__ V, const double* __restrict__ y,
               const double* __restrict__ dinv, double* __restrict__ x) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        double s = 0.0;
        for (int k = 0; k < nk; ++k) s += y[k] * V[(int64_t)k * n + i];
        x[i] += dinv[i] * s;
    }
}

#define NSB_GMRES_MAX_RESTART 60

extern "C" int64_t nsb_gmres_workspace_doubles(int64_t n, int32_t restart) { return (int64_t)(restart + 3) * n; }

extern "C" int nsb_gmres(nsb_ctx* ctx, const nsb_matrix* A, const double* dinv, const double* b, double* x,
                         double rtol, double atol, int32_t maxit, int32_t restart, double* work,
                         int32_t* iters_out, double* resid_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && dinv && b && x && work, "null argument");
    NSB_REQUIRE(ctx, restart >= 1 && restart <= NSB_GMRES_MAX_RESTART, "restart must be in [1, 60]");
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = (int64_t)A->nrows * A->bs;
    const int m = restart;
    double* V = work;                       // (m+1) x n
    double* w = work + (int64_t)(m + 1) * n;
    double* t = w + n;
    double* ds = ctx->d_scalars;
    double* hs = ctx->h_scalars;
    const int G = krylov_grid(ctx, n);
    std::vector<double> H((size_t)(m + 1) * m, 0.0), cs(m), sn(m), g(m + 1), y(m);

    // ||b||
    {
        NSB_TIMED(ctx, NSB_T_KRYLOV, st);
        k_reduce<true><<<G, 256, 0, st>>>(n, b, b, ctx->d_partials, ctx->d_counter, ds + S_OUT);
    }
    NSB_LAUNCH_CHECK(ctx, "k_reduce");
    if (fetch(ctx, S_OUT, 1, st)) return -1;
    const double bnorm = sqrt(hs[S_OUT]);
    const double tol = fmax(rtol * bnorm, atol);
    int it = 0;
    double rnorm = 0.0;
    while (true) {
        // true residual r = b - A x -> w
        if (nsb_launch_spmv(ctx, A, x, t, nullptr, nullptr, st)) return -1;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_residual<<<G, 256, 0, st>>>(n, b, t, w, ctx->d_partials, ctx->d_counter, ds + S_NRM);
        }
        NSB_LAUNCH_CHECK(ctx, "k_residual");
        if (fetch(ctx, S_NRM, 1, st)) return -1;
        rnorm = sqrt(hs[S_NRM]);
        if (!(rnorm == rnorm)) return nsb_fail(ctx, "GMRES breakdown: residual is NaN%s%s");
        if (rnorm <= tol || it >= maxit) break;
        {
            NSB_TIMED(ctx, NSB_T_KRYLOV, st);
            k_normalize<<<G, 256, 0, st>>>(n, w, ds + S_NRM, dinv, V, t);
        }
        NSB_COUNT(ctx, 1);
        g.assign(m + 1, 0.0);
        g[0] = rnorm;
        int j = 0;
        for (; j < m && it < maxit; ++j) {
            if (nsb_launch_spmv(ctx, A, t, w, nullptr, nullptr, st)) return -1;
            const int nk = j + 1;
            nsb_timed_region* arnoldi = new nsb_timed_region(ctx, NSB_T_KRYLOV, st);
            for (int pass = 0; pass < 2; ++pass) {   // CGS-2
                double* hdev = ds + (pass == 0 ? S_H1 : S_H2);
                for (int k0 = 0; k0 < nk; k0 += 8) {
                    const int c = nk - k0 < 8 ? nk - k0 : 8;
                    k_multi_dot<8><<<G, 256, 0, st>>>(n, k0, c, V, w, ctx->d_partials, ctx->d_counter, hdev);
                    NSB_COUNT(ctx, 1);
                }
                k_multi_axpy<<<G, 256, 0, st>>>(n, nk, V, hdev, w, ctx->d_partials, ctx->d_counter, ds + S_NRM);
            }
            k_normalize<<<G, 256, 0, st>>>(n, w, ds + S_NRM, dinv, V + (int64_t)(j + 1) * n, t);
            delete arnoldi;
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
        // back substitution, x += Dinv V y
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
            k_gmres_update<<<G, 256, 0, st>>>(n, nk, V, ds + S_Y, dinv, x);
        }
        NSB_LAUNCH_CHECK(ctx, "k_gmres_update");
        NSB_CUDA(ctx, cudaStreamSynchronize(st));   // hs[S_Y..] is reused next cycle
    }
    if (iters_out) *iters_out = it;
    if (resid_out) *resid_out = rnorm;
    if (rnorm > tol) return nsb_fail(ctx, "GMRES did not converge within the iteration limit%s%s");
    return 0;
}
