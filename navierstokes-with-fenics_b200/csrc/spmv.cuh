// spmv.cuh -- FP64 SpMV kernels (K10): Kronecker-CSR (scalar matrix applied to BS interleaved
// components) and BSR-BS.  LANES cooperating lanes own one (block) row; value and column
// streams are read coalesced along the row; x is gathered through L2 (x fits the 126 MB L2 for
// every config of SURVEY.md 8).  Rows flagged in bc_mask act as identity rows (dolfin `ident`).
// On a partitioned mesh the rows are the owned prefix and a column >= nown is a ghost, read from
// xg (= ghost buffer - BS*nown, so that xg[BS*col + c] addresses it); one GPU: nown = INT32_MAX.
// Optional fused epilogue: deterministic grid-wide dot(x-like vector, y)  (CG: p.Ap).
#pragma once
#include "common.cuh"

#define NSB_SPMV_THREADS 256

template <int BS, int LANES, bool PAIRS = false>
__device__ __forceinline__ void csr_row(int32_t row, int lane, const int32_t* __restrict__ rowptr,
                                        const int32_t* __restrict__ col, const double* __restrict__ vals,
                                        const double* __restrict__ x, int32_t nown,
                                        const double* __restrict__ xg, double (&acc)[BS], int32_t nlast = 0) {
#pragma unroll
    for (int b = 0; b < BS; ++b) acc[b] = 0.0;
    const int32_t end = rowptr[row + 1];
    // rows are short (15 .. 30 entries): offsets -> columns -> gather is a chain of dependent loads, so a
    // lane keeps U independent entries in flight (the kernels are latency bound, not bandwidth bound)
    constexpr int U = 4;
    for (int32_t s = rowptr[row] + lane; s < end; s += LANES * U) {
        double v[U];
        int32_t cn[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int32_t ss = s + u * LANES;
            const bool ok = ss < end;
            v[u] = ok ? vals[ss] : 0.0;
            cn[u] = ok ? col[ss] : row;
        }
        double xv[U][BS];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (BS == 3 && PAIRS && cn[u] < nlast) {
                // three interleaved components = 24 unaligned bytes: two aligned 16-byte loads cover them (the node
                // behind the last one is never touched: cn < nlast), one L1 request less per entry
                const int64_t e0 = (int64_t)cn[u] * 3, a0 = e0 & ~(int64_t)1;
                const double2 lo = *reinterpret_cast<const double2*>(x + a0);
                const double2 hi = *reinterpret_cast<const double2*>(x + a0 + 2);
                const bool odd = (e0 & 1) != 0;
                xv[u][0] = odd ? lo.y : lo.x;
                xv[u][1] = odd ? hi.x : lo.y;
                xv[u][2 % BS] = odd ? hi.y : hi.x;
            } else {
                const double* xs = (cn[u] < nown ? x : xg) + (int64_t)cn[u] * BS;
#pragma unroll
                for (int b = 0; b < BS; ++b) xv[u][b] = xs[b];
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int b = 0; b < BS; ++b) acc[b] += v[u] * xv[u][b];
    }
}

template <int BS, int LANES>
__device__ __forceinline__ void bsr_row(int32_t row, int lane, const int32_t* __restrict__ rowptr,
                                        const int32_t* __restrict__ col, const double* __restrict__ vals,
                                        const double* __restrict__ x, int32_t nown,
                                        const double* __restrict__ xg, double (&acc)[BS]) {
    constexpr int BB = BS * BS;
#pragma unroll
    for (int b = 0; b < BS; ++b) acc[b] = 0.0;
    const int32_t r0 = rowptr[row];
    const int n = (rowptr[row + 1] - r0) * BB;
    const double* v = vals + (int64_t)r0 * BB;
    for (int e = lane; e < n; e += LANES) {
        const int blk = e / BB, rc = e - blk * BB;
        const int r = rc / BS, c = rc - r * BS;
        const int32_t cn = col[r0 + blk];
        const double t = v[e] * (cn < nown ? x : xg)[(int64_t)cn * BS + c];
#pragma unroll
        for (int b = 0; b < BS; ++b) acc[b] += (b == r) ? t : 0.0;
    }
}

// y = A x ; if DOT: scalars_out[0] = sum_i d[i] * y[i]  (d = dot_with)
// CHEB: instead of storing y = A x the row applies the Chebyshev step of krylov.cu to it (x = d_old):
// r -= D^-1 y ; d_new = a d_old + b r ; z += d_new
template <int KIND, int BS, int LANES, bool DOT, bool CHEB = false>
__global__ void __launch_bounds__(NSB_SPMV_THREADS)
k_spmv(int32_t nrows, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
       const double* __restrict__ vals, const uint8_t* __restrict__ mask, const double* __restrict__ x,
       int32_t nown, const double* __restrict__ xg,
       double* __restrict__ y, const double* __restrict__ dot_with, double* partials,
       unsigned int* counter, double* scalar_out, nsb_cheb64_epi E = nsb_cheb64_epi{}) {
    constexpr int ROWS_PER_BLOCK = NSB_SPMV_THREADS / LANES;
    const int lane = threadIdx.x % LANES;
    double dsum[1] = {0.0};
    for (int64_t row0 = (int64_t)blockIdx.x * ROWS_PER_BLOCK; row0 < nrows;
         row0 += (int64_t)gridDim.x * ROWS_PER_BLOCK) {
        const int64_t row = row0 + threadIdx.x / LANES;
        const bool valid = row < nrows;
        double acc[BS];
        if (valid) {
            if (KIND == NSB_CSR) {
                // paired 16-byte gathers need an aligned x; columns are owned nodes < nrows (square operators)
                const int32_t nlast = (reinterpret_cast<uintptr_t>(x) & 15) == 0 ? nrows - 1 : 0;
                // (measured on B200, C5 mass matrix: 0.70 ms with the paired loads against 0.57 ms with three 8-byte
                // loads -- the two 16-byte requests touch more L1 wavefronts than they save instructions: left off)
                csr_row<BS, LANES, false>((int32_t)row, lane, rowptr, col, vals, x, nown, xg, acc, nlast);
            }
            else bsr_row<BS, LANES>((int32_t)row, lane, rowptr, col, vals, x, nown, xg, acc);
        } else {
#pragma unroll
            for (int b = 0; b < BS; ++b) acc[b] = 0.0;
        }
#pragma unroll
        for (int b = 0; b < BS; ++b) acc[b] = group_sum<LANES>(acc[b]);
        if (valid && lane == 0) {
#pragma unroll
            for (int b = 0; b < BS; ++b) {
                const int64_t d = row * BS + b;
                double v = acc[b];
                if (mask != nullptr && mask[d]) v = x[d];
                if (CHEB) {
                    const double ri = E.r[d] - E.dinv[d] * v;
                    const double dn = E.a * x[d] + E.b * ri;
                    E.r[d] = ri; E.dnew[d] = dn; E.z[d] += dn;
                } else {
                    y[d] = v;
                }
                if (DOT) dsum[0] += dot_with[d] * v;
            }
        }
    }
    if (DOT) {
        __shared__ double red[NSB_SPMV_THREADS / 32];
        grid_sum<1, NSB_SPMV_THREADS>(dsum, partials, counter, scalar_out, red);
    }
}

// host-side launcher (defined in linalg.cu)
// lowp: the caller accepts the single-precision value copy A->vals32 (preconditioner applications only)
int nsb_launch_spmv(nsb_ctx* ctx, const nsb_matrix* A, const double* x, double* y,
                    const double* dot_with, double* scalar_out, cudaStream_t st, bool lowp = false);
