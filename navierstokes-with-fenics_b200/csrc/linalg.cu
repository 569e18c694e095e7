// linalg.cu -- SpMV launcher, Jacobi diagonal (K12), vector kernels and the BDF history shift
// (K13), reductions (K15).
#include "spmv.cuh"
#include "spmv_tma.cuh"

// ---- SpMV dispatch ---------------------------------------------------------------------------
template <int KIND, int BS, int LANES>
static int launch_spmv_t(nsb_ctx* ctx, const nsb_matrix* A, const double* x, int32_t nown, const double* xg,
                         double* y, const double* dot_with, double* scalar_out, cudaStream_t st) {
    constexpr int RPB = NSB_SPMV_THREADS / LANES;
    int64_t blocks = ((int64_t)A->nrows + RPB - 1) / RPB;
    nsb_timed_region region(ctx, KIND == NSB_BSR ? NSB_T_SPMV_BSR : NSB_T_SPMV_CSR, st);
    if (dot_with) {
        const int64_t cap = (int64_t)ctx->sm_count * 8;
        if (blocks > cap) blocks = cap;
        k_spmv<KIND, BS, LANES, true><<<(int)blocks, NSB_SPMV_THREADS, 0, st>>>(
            A->nrows, A->rowptr, A->col, A->vals, A->bc_mask, x, nown, xg, y, dot_with, ctx->d_partials,
            ctx->d_counter, scalar_out);
    } else if (KIND == NSB_CSR && ctx->spmv_epi) {
        const nsb_cheb64_epi E = *ctx->spmv_epi;
        ctx->spmv_epi = nullptr;
        k_spmv<KIND, BS, LANES, false, true><<<(int)blocks, NSB_SPMV_THREADS, 0, st>>>(
            A->nrows, A->rowptr, A->col, A->vals, A->bc_mask, x, nown, xg, y, nullptr, nullptr, nullptr, nullptr, E);
    } else {
        k_spmv<KIND, BS, LANES, false><<<(int)blocks, NSB_SPMV_THREADS, 0, st>>>(
            A->nrows, A->rowptr, A->col, A->vals, A->bc_mask, x, nown, xg, y, nullptr, nullptr, nullptr, nullptr);
    }
    NSB_LAUNCH_CHECK(ctx, "k_spmv");
    return 0;
}

template <int KIND, int BS>
static int launch_spmv_bs(nsb_ctx* ctx, const nsb_matrix* A, const double* x, int32_t nown, const double* xg,
                          double* y, const double* dot_with, double* scalar_out, cudaStream_t st) {
    double per_row = A->nrows > 0 ? (double)A->nnz / (double)A->nrows : 0.0;
    if (KIND == NSB_BSR) per_row *= BS * BS;
    if (KIND == NSB_CSR) {      // csr_row keeps 4 entries per lane in flight: fewer lanes per row
        if (per_row <= 40.0) return launch_spmv_t<KIND, BS, 4>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
        if (per_row <= 80.0) return launch_spmv_t<KIND, BS, 8>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
        return launch_spmv_t<KIND, BS, 16>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
    }
    if (per_row <= 10.0) return launch_spmv_t<KIND, BS, 4>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
    if (per_row <= 24.0) return launch_spmv_t<KIND, BS, 8>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
    if (per_row <= 64.0) return launch_spmv_t<KIND, BS, 16>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
    return launch_spmv_t<KIND, BS, 32>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
}

static int spmv_dispatch(nsb_ctx* ctx, const nsb_matrix* A, const double* x, int32_t nown, const double* xg,
                         double* y, const double* dot_with, double* scalar_out, cudaStream_t st, bool lowp);

int nsb_launch_spmv(nsb_ctx* ctx, const nsb_matrix* A, const double* x, double* y,
                    const double* dot_with, double* scalar_out, cudaStream_t st, bool lowp) {
    NSB_REQUIRE(ctx, A && A->rowptr && A->col && A->vals && A->nrows > 0, "bad matrix descriptor");
    NSB_REQUIRE(ctx, x && y && x != y, "bad vectors (x must not alias y)");
    int32_t nown = INT32_MAX;
    const double* xg = x;
    if (ctx->nranks > 1) {
        // partitioned mesh: refresh the ghost columns of x from their owners, then multiply the owned rows
        NSB_REQUIRE(ctx, A->halo == NSB_HALO_NODES || A->halo == NSB_HALO_VERTS, "matrix of a partitioned run needs a halo index space");
        const nsb_halo& H = ctx->halo[A->halo];
        NSB_REQUIRE(ctx, H.n_owned == A->nrows, "matrix rows are not the owned prefix of the halo plan");
        if (nsb_halo_exchange(ctx, A->halo, A->bs, x, nullptr, st)) return -1;
        nown = H.n_owned;
        xg = H.ghostbuf - (int64_t)nown * A->bs;
    }
    if (spmv_dispatch(ctx, A, x, nown, xg, y, dot_with, scalar_out, st, lowp)) return -1;
    if (ctx->spmv_epi) {
        ctx->spmv_epi = nullptr;
        return nsb_fail(ctx, "internal: fused Chebyshev epilogue requested for a SpMV kernel that has none%s%s");
    }
    if (dot_with && nsb_allreduce_sum(ctx, scalar_out, 1, st)) return -1;
    return 0;
}

// y = A32 x in single precision throughout (FP32 Chebyshev preconditioner, krylov.cu): x, y hold
// NSB_PAD32(bs) floats per node.  On a partitioned mesh the padded node is moved as 8-byte words by the
// ordinary halo exchange (a float4 = 2 doubles, a float2 = 1).
bool nsb_spmv32_available(const nsb_matrix* A) {
    if (A->bs != 3 && A->bs != 2) return false;
    // a Kronecker CSR matrix (scalar values applied to bs interleaved components) IS block-scalar: its float copy in
    // diag32 serves the same kernel (the velocity mass matrix of the correction step)
    if (A->kind == NSB_CSR) return A->diag32 != nullptr;
    if (A->kind != NSB_BSR) return false;
    if (A->diag32) return true;
    return A->vals32 && A->tiles && A->ntiles > 0 &&
           ((((uintptr_t)A->vals32) | ((uintptr_t)A->col) | ((uintptr_t)A->tiles)) & 15) == 0;
}

// Block-scalar approximation S (x) I of a BSR matrix, S_ij = mean of the diagonal of block ij, in single
// precision (nsb_matrix.diag32): y_i = sum_j S_ij x_j on NSB_PAD32(BS)-padded node vectors.  8 bytes per
// block entry instead of 40: the momentum Jacobian is (mass/dt + diffusion + convection) (x) I plus the
// lower-order Newton coupling, so as a PRECONDITIONER the scalar part does as well as the full blocks
// (same GMRES iteration counts, tools/probe_scalar_pc.py).  LANES lanes per row, coalesced value / column
// streams, one 16-byte gather per entry.
template <int BS, int LANES, int U, bool CHEB, bool H16 = false>
__global__ void __launch_bounds__(256)
k_spmv_kron32(int32_t nrows, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
              const float* __restrict__ s, const float* __restrict__ x, int32_t nown, const float* __restrict__ xg,
              float* __restrict__ y, nsb_cheb32_epi E, const unsigned long long* __restrict__ xgh = nullptr) {
    constexpr int PB = NSB_PAD32(BS);
    const int lane = threadIdx.x % LANES;
    const int64_t row = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) / LANES;
    if (row >= nrows) return;                      // whole lane groups leave together
    float acc[BS];
#pragma unroll
    for (int b = 0; b < BS; ++b) acc[b] = 0.0f;
    const int32_t end = rowptr[row + 1];
    // The kernel is latency bound (a row is ~28 entries: row offsets -> columns -> gather is a chain of three
    // dependent loads and then the warp is done), so every lane keeps U independent gathers in flight.
    for (int32_t k = rowptr[row] + lane; k < end; k += LANES * U) {
        int32_t cn[U];
        float v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int32_t kk = k + u * LANES;
            const bool ok = kk < end;
            cn[u] = ok ? col[kk] : (int32_t)row;
            v[u] = ok ? s[kk] : 0.0f;
        }
        if constexpr (BS == 3 && H16) {
            // bf16 shadow: one 8-byte gather per entry (ghost columns: the shadow's ghost block, E.x_h is [owned | ghost])
            unsigned long long t[U];
#pragma unroll
            for (int u = 0; u < U; ++u) t[u] = __ldg((cn[u] < nown ? E.x_h : xgh) + cn[u]);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const unsigned int lo = (unsigned int)t[u], hi = (unsigned int)(t[u] >> 32);
                acc[0] += v[u] * __uint_as_float(lo << 16);
                acc[1] += v[u] * __uint_as_float(lo & 0xffff0000u);
                acc[2] += v[u] * __uint_as_float(hi << 16);
            }
        } else if constexpr (BS == 3) {
            float4 t[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                t[u] = __ldg(reinterpret_cast<const float4*>((cn[u] < nown ? x : xg) + (int64_t)cn[u] * PB));
#pragma unroll
            for (int u = 0; u < U; ++u) { acc[0] += v[u] * t[u].x; acc[1] += v[u] * t[u].y; acc[2] += v[u] * t[u].z; }
        } else {
            float2 t[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                t[u] = __ldg(reinterpret_cast<const float2*>((cn[u] < nown ? x : xg) + (int64_t)cn[u] * PB));
#pragma unroll
            for (int u = 0; u < U; ++u) { acc[0] += v[u] * t[u].x; acc[1] += v[u] * t[u].y; }
        }
    }
#pragma unroll
    for (int b = 0; b < BS; ++b)
#pragma unroll
        for (int o = LANES / 2; o > 0; o >>= 1) acc[b] += __shfl_xor_sync(0xffffffffu, acc[b], o, LANES);
    if (lane == 0) {
        if constexpr (CHEB) {
            float ri[BS], dold[BS], zi[BS];
            if constexpr (BS == 3) {
                const float4 a4 = reinterpret_cast<const float4*>(E.r)[row], b4 = reinterpret_cast<const float4*>(x)[row],
                             c4 = reinterpret_cast<const float4*>(E.z)[row];
                ri[0] = a4.x; ri[1] = a4.y; ri[2] = a4.z; dold[0] = b4.x; dold[1] = b4.y; dold[2] = b4.z;
                zi[0] = c4.x; zi[1] = c4.y; zi[2] = c4.z;
            } else {
                const float2 a2 = reinterpret_cast<const float2*>(E.r)[row], b2 = reinterpret_cast<const float2*>(x)[row],
                             c2 = reinterpret_cast<const float2*>(E.z)[row];
                ri[0] = a2.x; ri[1] = a2.y; dold[0] = b2.x; dold[1] = b2.y; zi[0] = c2.x; zi[1] = c2.y;
            }
            if (E.mask != nullptr) {
#pragma unroll
                for (int c = 0; c < BS; ++c)
                    if (E.mask[row * BS + c]) acc[c] = dold[c];
            }
#pragma unroll
            for (int c = 0; c < BS; ++c) {
                ri[c] -= (float)E.dinv[row * BS + c] * acc[c];
                acc[c] = E.a * dold[c] + E.b * ri[c];          // d_new
                zi[c] += acc[c];
            }
            if (E.zout) {
#pragma unroll
                for (int c = 0; c < BS; ++c) E.zout[row * BS + c] = (double)zi[c];
            } else if constexpr (BS == 3) {
                reinterpret_cast<float4*>(E.r)[row] = make_float4(ri[0], ri[1], ri[2], 0.0f);
                reinterpret_cast<float4*>(E.dnew)[row] = make_float4(acc[0], acc[1], acc[2], 0.0f);
                if constexpr (H16) E.dnew_h[row] = nsb_pack_bf16x3(acc[0], acc[1], acc[2]);
                reinterpret_cast<float4*>(E.z)[row] = make_float4(zi[0], zi[1], zi[2], 0.0f);
            } else {
                reinterpret_cast<float2*>(E.r)[row] = make_float2(ri[0], ri[1]);
                reinterpret_cast<float2*>(E.dnew)[row] = make_float2(acc[0], acc[1]);
                reinterpret_cast<float2*>(E.z)[row] = make_float2(zi[0], zi[1]);
            }
        } else {
            if (E.mask != nullptr) {
#pragma unroll
                for (int c = 0; c < BS; ++c)
                    if (E.mask[row * BS + c]) acc[c] = x[row * PB + c];
            }
            if constexpr (BS == 3) reinterpret_cast<float4*>(y)[row] = make_float4(acc[0], acc[1], acc[2], 0.0f);
            else reinterpret_cast<float2*>(y)[row] = make_float2(acc[0], acc[1]);
        }
    }
}

template <int BS>
static int launch_spmv_kron32(nsb_ctx* ctx, const nsb_matrix* A, const float* x, int32_t nown, const float* xg, float* y,
                              const nsb_cheb32_epi* epi, cudaStream_t st, const unsigned long long* xgh = nullptr) {
    const double per_row = (double)A->nnz / (double)A->nrows;
    // measured on B200 at C5 (28 entries per row), preconditioner SpMV ms per step: 4 lanes x 4 gathers 34.0,
    // 8 x 4 38.7, 2 x 4 39.1, 4 x 8 39.5, 8 x 8 56.5, 16 x 2 66.8, 32 x 2 121
    const int lanes = per_row <= 40.0 ? 4 : 8;
    const int blocks = nsb_blocks((int64_t)A->nrows * lanes, 256);
    nsb_cheb32_epi E = epi ? *epi : nsb_cheb32_epi{};
    E.mask = A->bc_mask;
#define NSB_KRON(LANES, U)                                                                                                  \
    do {                                                                                                                    \
        if (epi) k_spmv_kron32<BS, LANES, U, true><<<blocks, 256, 0, st>>>(A->nrows, A->rowptr, A->col, A->diag32, x, nown, xg, y, E);  \
        else k_spmv_kron32<BS, LANES, U, false><<<blocks, 256, 0, st>>>(A->nrows, A->rowptr, A->col, A->diag32, x, nown, xg, y, E); \
    } while (0)
    if constexpr (BS == 3) {
        if (epi && epi->x_h) {        // bf16 shadow gathers (fused Chebyshev step only)
            if (lanes == 4) k_spmv_kron32<3, 4, 4, true, true><<<blocks, 256, 0, st>>>(A->nrows, A->rowptr, A->col, A->diag32, x, nown, xg, y, E, xgh);
            else k_spmv_kron32<3, 8, 4, true, true><<<blocks, 256, 0, st>>>(A->nrows, A->rowptr, A->col, A->diag32, x, nown, xg, y, E, xgh);
            NSB_LAUNCH_CHECK(ctx, "k_spmv_kron32(bf16)");
            return 0;
        }
    }
    if (lanes == 4) NSB_KRON(4, 4);
    else NSB_KRON(8, 4);
#undef NSB_KRON
    NSB_LAUNCH_CHECK(ctx, "k_spmv_kron32");
    return 0;
}

// out[k] = mean of the diagonal of block k, for the blocks of the listed block rows only (after nsb_bc_apply_matrix
// changed those rows)
__global__ void __launch_bounds__(256)
k_block_scalar_rows(int32_t nlisted, const int32_t* __restrict__ rows, const int32_t* __restrict__ rowptr, int bs,
                    const double* __restrict__ vals, float* __restrict__ out) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = t >> 3;                              // 8 lanes per listed row
    if (r >= nlisted) return;
    const int32_t row = rows[r];
    for (int32_t k = rowptr[row] + (int)(t & 7); k < rowptr[row + 1]; k += 8) {
        const double* b = vals + (int64_t)k * bs * bs;
        double sdiag = 0.0;
        for (int i = 0; i < bs; ++i) sdiag += b[i * (bs + 1)];
        out[k] = (float)(sdiag / bs);
    }
}
extern "C" int nsb_bsr_block_scalar_rows_f32(nsb_ctx* ctx, int32_t nlisted, const int32_t* block_rows, const int32_t* rowptr,
                                             int32_t bs, const double* vals, float* out, void* stream) {
    NSB_REQUIRE(ctx, ctx && nlisted >= 0 && bs >= 1 && bs <= 3 && (nlisted == 0 || (block_rows && rowptr && vals && out)),
                "bad arguments");
    if (nlisted == 0) return 0;
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_block_scalar_rows<<<nsb_blocks((int64_t)nlisted * 8, 256), 256, 0, (cudaStream_t)stream>>>(nlisted, block_rows, rowptr, bs, vals, out);
    NSB_LAUNCH_CHECK(ctx, "k_block_scalar_rows");
    NSB_COUNT(ctx, 1);
    return 0;
}

extern "C" int nsb_set_block_scalar_out(nsb_ctx* ctx, float* out, int64_t nnz_blocks) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, out == nullptr || nnz_blocks > 0, "the block count of the P2 x P2 pattern is needed");
    ctx->jac_block_scalar_out = out;
    ctx->jac_block_scalar_nnz = out ? nnz_blocks : 0;
    return 0;
}

// out[k] = mean of the diagonal of block k
__global__ void __launch_bounds__(256)
k_block_scalar(int64_t nnz, int bs, const double* __restrict__ vals, float* __restrict__ out) {
    const int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= nnz) return;
    const double* b = vals + k * bs * bs;
    double sdiag = 0.0;
    for (int i = 0; i < bs; ++i) sdiag += b[i * (bs + 1)];
    out[k] = (float)(sdiag / bs);
}
extern "C" int nsb_bsr_block_scalar_f32(nsb_ctx* ctx, int64_t nnz, int32_t bs, const double* vals, float* out, void* stream) {
    NSB_REQUIRE(ctx, ctx && vals && out && nnz >= 0 && bs >= 1 && bs <= 3, "bad arguments");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_block_scalar<<<nsb_blocks(nnz, 256), 256, 0, (cudaStream_t)stream>>>(nnz, bs, vals, out);
    NSB_LAUNCH_CHECK(ctx, "k_block_scalar");
    return 0;
}

int nsb_launch_spmv32(nsb_ctx* ctx, const nsb_matrix* A, const float* x, float* y, cudaStream_t st,
                      const nsb_cheb32_epi* epi) {
    NSB_REQUIRE(ctx, A && nsb_spmv32_available(A), "matrix has no single-precision copy (diag32, or vals32 + row tiles)");
    NSB_REQUIRE(ctx, x && (epi || y) && x != y, "bad vectors (x must not alias y)");
    NSB_REQUIRE(ctx, !epi || (A->diag32 && epi->dinv && epi->r && epi->z && (epi->zout || epi->dnew) && epi->dnew != x),
                "fused Chebyshev step needs the block-scalar matrix and a second direction buffer");
    const int pb = NSB_PAD32(A->bs);
    int32_t nown = INT32_MAX;
    const float* xg = x;
    const bool shadow = epi && epi->x_h && A->bs == 3 && A->diag32;
    const unsigned long long* xgh = shadow ? epi->x_h : nullptr;
    if (ctx->nranks > 1) {
        NSB_REQUIRE(ctx, A->halo == NSB_HALO_NODES || A->halo == NSB_HALO_VERTS, "matrix of a partitioned run needs a halo index space");
        const nsb_halo& H = ctx->halo[A->halo];
        NSB_REQUIRE(ctx, H.n_owned == A->nrows, "matrix rows are not the owned prefix of the halo plan");
        nown = H.n_owned;
        if (shadow) {     // only the gathered copy travels: 8 bytes per node instead of 16
            if (nsb_halo_exchange(ctx, A->halo, 1, reinterpret_cast<const double*>(epi->x_h), nullptr, st)) return -1;
            xgh = reinterpret_cast<const unsigned long long*>(H.ghostbuf) - (int64_t)nown;
        } else {
            if (nsb_halo_exchange(ctx, A->halo, pb / 2, reinterpret_cast<const double*>(x), nullptr, st)) return -1;
            xg = reinterpret_cast<const float*>(H.ghostbuf) - (int64_t)nown * pb;
        }
    }
    NSB_TIMED(ctx, NSB_T_SPMV_BSR32, st);
    if (A->diag32) return A->bs == 3 ? launch_spmv_kron32<3>(ctx, A, x, nown, xg, y, epi, st, xgh) : launch_spmv_kron32<2>(ctx, A, x, nown, xg, y, epi, st);
    NSB_COUNT(ctx, 1);
    if (A->bs == 3) return launch_spmv_bsr_tma<3, float, float>(ctx, A, A->vals32, x, nown, xg, y, st);
    return launch_spmv_bsr_tma<2, float, float>(ctx, A, A->vals32, x, nown, xg, y, st);
}

static int spmv_dispatch(nsb_ctx* ctx, const nsb_matrix* A, const double* x, int32_t nown, const double* xg,
                         double* y, const double* dot_with, double* scalar_out, cudaStream_t st, bool lowp) {
    if (A->kind == NSB_CSR) {
        switch (A->bs) {
            case 1: return launch_spmv_bs<NSB_CSR, 1>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
            case 2: return launch_spmv_bs<NSB_CSR, 2>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
            case 3: return launch_spmv_bs<NSB_CSR, 3>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
        }
    } else if (A->kind == NSB_BSR) {
        const bool tma_ok = A->tiles && A->ntiles > 0 && !dot_with &&
                            ((((uintptr_t)A->vals) | ((uintptr_t)A->col) | ((uintptr_t)A->tiles)) & 15) == 0;
        if (tma_ok) {
            const bool f32 = lowp && A->vals32 && (((uintptr_t)A->vals32) & 15) == 0;
            nsb_timed_region region(ctx, f32 ? NSB_T_SPMV_BSR32 : NSB_T_SPMV_BSR, st);
            // preconditioner-only single-precision copy of the values (see spmv_tma.cuh)
            if (f32) {
                if (A->bs == 3) return launch_spmv_bsr_tma<3, float>(ctx, A, A->vals32, x, nown, xg, y, st);
                if (A->bs == 2) return launch_spmv_bsr_tma<2, float>(ctx, A, A->vals32, x, nown, xg, y, st);
            }
            if (A->bs == 3) return launch_spmv_bsr_tma<3, double>(ctx, A, A->vals, x, nown, xg, y, st);
            if (A->bs == 2) return launch_spmv_bsr_tma<2, double>(ctx, A, A->vals, x, nown, xg, y, st);
        }
        switch (A->bs) {
            case 2: return launch_spmv_bs<NSB_BSR, 2>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
            case 3: return launch_spmv_bs<NSB_BSR, 3>(ctx, A, x, nown, xg, y, dot_with, scalar_out, st);
        }
    }
    return nsb_fail(ctx, "unsupported matrix kind / block size%s%s");
}

extern "C" int nsb_spmv(nsb_ctx* ctx, const nsb_matrix* A, const double* x, double* y, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    return nsb_launch_spmv(ctx, A, x, y, nullptr, nullptr, (cudaStream_t)stream, false);
}

// ---- Jacobi: inverse diagonal ------------------------------------------------------------------
__global__ void k_inv_diag(int kind, int bs, int32_t nrows, const int32_t* __restrict__ rowptr,
                           const int32_t* __restrict__ col, const double* __restrict__ vals,
                           const uint8_t* __restrict__ mask, double* __restrict__ out) {
    const int64_t d = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= (int64_t)nrows * bs) return;
    const int32_t row = (int32_t)(d / bs);
    const int comp = (int)(d % bs);
    double diag = 0.0;
    int lo = rowptr[row], hi = rowptr[row + 1] - 1;
    while (lo <= hi) {   // columns ascending: binary search for the diagonal
        const int mid = (lo + hi) >> 1;
        const int32_t c = col[mid];
        if (c == row) {
            diag = (kind == NSB_CSR) ? vals[mid] : vals[(int64_t)mid * bs * bs + comp * bs + comp];
            break;
        }
        if (c < row) lo = mid + 1; else hi = mid - 1;
    }
    if (mask != nullptr && mask[d]) diag = 1.0;
    out[d] = (diag != 0.0) ? 1.0 / diag : 1.0;
}

extern "C" int nsb_inv_diagonal(nsb_ctx* ctx, const nsb_matrix* A, double* inv_diag, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, A && A->rowptr && A->col && A->vals && inv_diag, "bad arguments");
    const int64_t n = (int64_t)A->nrows * A->bs;
    nsb_timed_region region(ctx, NSB_T_KRYLOV, (cudaStream_t)stream);
    k_inv_diag<<<nsb_blocks(n, 256), 256, 0, (cudaStream_t)stream>>>(A->kind, A->bs, A->nrows, A->rowptr,
                                                                       A->col, A->vals, A->bc_mask, inv_diag);
    NSB_LAUNCH_CHECK(ctx, "k_inv_diag");
    return 0;
}

// ---- vector kernels ------------------------------------------------------------------------------
__global__ void k_axpy(int64_t n, double a, const double* __restrict__ x, double* __restrict__ y) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] += a * x[i];
}
__global__ void k_scale(int64_t n, double a, double* __restrict__ x) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        x[i] *= a;
}
__global__ void k_scatter(int64_t n, const int32_t* __restrict__ idx, const double* __restrict__ x,
                          double* __restrict__ y) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[idx[i]] = x[i];
}
__global__ void k_gather(int64_t n, const int32_t* __restrict__ idx, const double* __restrict__ x,
                         double* __restrict__ y) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        y[i] = x[idx[i]];
}
// u2 <- u1 <- u0 and p_old <- p in one pass (source/ns_ipcs_solver.py:35-43)
__global__ void k_ipcs_advance(int64_t nv, const double* __restrict__ u0, double* __restrict__ u1,
                               double* __restrict__ u2, int64_t nq, const double* __restrict__ p,
                               double* __restrict__ pold) {
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += stride) {
        const double a = u1[i];
        u1[i] = u0[i];
        u2[i] = a;
    }
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nq; i += stride) pold[i] = p[i];
}

template <bool WITH_Y>
__global__ void __launch_bounds__(256)
k_dot(int64_t n, const double* __restrict__ x, const double* __restrict__ y, double* partials,
      unsigned int* counter, double* out) {
    double s[1] = {0.0};
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        s[0] += WITH_Y ? x[i] * y[i] : x[i];
    __shared__ double red[8];
    grid_sum<1, 256>(s, partials, counter, out, red);
}

static int grid_for(nsb_ctx* ctx, int64_t n, int threads) {
    int64_t b = (n + threads - 1) / threads;
    const int64_t cap = (int64_t)ctx->sm_count * 8;
    if (b > cap) b = cap;
    if (b < 1) b = 1;
    return (int)b;
}

extern "C" int nsb_vec_copy(nsb_ctx* ctx, int64_t n, const double* x, double* y, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && y && n >= 0, "bad arguments");
    NSB_CUDA(ctx, cudaMemcpyAsync(y, x, sizeof(double) * n, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
    return 0;
}
extern "C" int nsb_vec_axpy(nsb_ctx* ctx, int64_t n, double a, const double* x, double* y, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && y && n >= 0, "bad arguments");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_axpy<<<grid_for(ctx, n, 256), 256, 0, (cudaStream_t)stream>>>(n, a, x, y);
    NSB_LAUNCH_CHECK(ctx, "k_axpy");
    return 0;
}
extern "C" int nsb_vec_scale(nsb_ctx* ctx, int64_t n, double a, double* x, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && n >= 0, "bad arguments");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_scale<<<grid_for(ctx, n, 256), 256, 0, (cudaStream_t)stream>>>(n, a, x);
    NSB_LAUNCH_CHECK(ctx, "k_scale");
    return 0;
}
// y = (float) x, 4 entries per thread (the FP32 value copy read by the Chebyshev SpMVs)
__global__ void __launch_bounds__(256)
k_to_f32(int64_t n, const double* __restrict__ x, float* __restrict__ y) {
    const int64_t n4 = n >> 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const double2 a = reinterpret_cast<const double2*>(x)[2 * i], b = reinterpret_cast<const double2*>(x)[2 * i + 1];
        reinterpret_cast<float4*>(y)[i] = make_float4((float)a.x, (float)a.y, (float)b.x, (float)b.y);
    }
    if (blockIdx.x == 0 && threadIdx.x < (n & 3)) y[(n4 << 2) + threadIdx.x] = (float)x[(n4 << 2) + threadIdx.x];
}
extern "C" int nsb_vec_to_f32(nsb_ctx* ctx, int64_t n, const double* x, float* y, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && y && n >= 0, "bad arguments");
    NSB_REQUIRE(ctx, ((((uintptr_t)x) | ((uintptr_t)y)) & 15) == 0, "buffers must be 16-byte aligned");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_to_f32<<<grid_for(ctx, (n >> 2) + 1, 256), 256, 0, (cudaStream_t)stream>>>(n, x, y);
    NSB_LAUNCH_CHECK(ctx, "k_to_f32");
    return 0;
}
extern "C" int nsb_vec_scatter(nsb_ctx* ctx, int64_t n, const int32_t* idx, const double* x, double* y, void* stream) {
    NSB_REQUIRE(ctx, ctx && idx && x && y && n >= 0, "bad arguments");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_scatter<<<grid_for(ctx, n, 256), 256, 0, (cudaStream_t)stream>>>(n, idx, x, y);
    NSB_LAUNCH_CHECK(ctx, "k_scatter");
    return 0;
}
extern "C" int nsb_vec_gather(nsb_ctx* ctx, int64_t n, const int32_t* idx, const double* x, double* y, void* stream) {
    NSB_REQUIRE(ctx, ctx && idx && x && y && n >= 0, "bad arguments");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_gather<<<grid_for(ctx, n, 256), 256, 0, (cudaStream_t)stream>>>(n, idx, x, y);
    NSB_LAUNCH_CHECK(ctx, "k_gather");
    return 0;
}
extern "C" int nsb_ipcs_advance(nsb_ctx* ctx, int64_t nv, const double* u0, double* u1, double* u2,
                                int64_t nq, const double* p, double* p_old, void* stream) {
    NSB_REQUIRE(ctx, ctx && u0 && u1 && u2 && p && p_old, "bad arguments");
    nsb_timed_region region(ctx, NSB_T_VEC, (cudaStream_t)stream);
    k_ipcs_advance<<<grid_for(ctx, nv, 256), 256, 0, (cudaStream_t)stream>>>(nv, u0, u1, u2, nq, p, p_old);
    NSB_LAUNCH_CHECK(ctx, "k_ipcs_advance");
    return 0;
}

static int fetch_scalar(nsb_ctx* ctx, double* out_host, cudaStream_t st) {
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, sizeof(double), cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    *out_host = ctx->h_scalars[0];
    return 0;
}
extern "C" int nsb_vec_dot(nsb_ctx* ctx, int64_t n, const double* x, const double* y, double* out_host, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && y && out_host && n >= 0, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    {
        NSB_TIMED(ctx, NSB_T_VEC, st);
        k_dot<true><<<grid_for(ctx, n, 256), 256, 0, st>>>(n, x, y, ctx->d_partials, ctx->d_counter, ctx->d_scalars);
    }
    NSB_LAUNCH_CHECK(ctx, "k_dot");
    if (nsb_allreduce_sum(ctx, ctx->d_scalars, 1, st)) return -1;
    return fetch_scalar(ctx, out_host, st);
}
extern "C" int nsb_vec_sum(nsb_ctx* ctx, int64_t n, const double* x, double* out_host, void* stream) {
    NSB_REQUIRE(ctx, ctx && x && out_host && n >= 0, "bad arguments");
    cudaStream_t st = (cudaStream_t)stream;
    {
        NSB_TIMED(ctx, NSB_T_VEC, st);
        k_dot<false><<<grid_for(ctx, n, 256), 256, 0, st>>>(n, x, nullptr, ctx->d_partials, ctx->d_counter, ctx->d_scalars);
    }
    NSB_LAUNCH_CHECK(ctx, "k_dot(sum)");
    if (nsb_allreduce_sum(ctx, ctx->d_scalars, 1, st)) return -1;
    return fetch_scalar(ctx, out_host, st);
}
