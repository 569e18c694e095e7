// context.cu -- life cycle of nsb_ctx, registration of device-resident mesh / pattern arrays,
// Dirichlet-row kernels (K9) and the FP64 peak probe.
#include <stdlib.h>
#include "common.cuh"

extern "C" int nsb_version(void) { return NSB_VERSION; }

extern "C" const char* nsb_last_error(const nsb_ctx* ctx) {
    return ctx ? ctx->err.c_str() : g_nsb_last_error.c_str();
}

extern "C" int nsb_create(nsb_ctx** out, int device) {
    if (!out) return nsb_fail(nullptr, "nsb_create: null output%s%s");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return nsb_fail(nullptr, "nsb_create: no CUDA device (%s); libnsb200 has no CPU fallback%s",
                        cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return nsb_fail(nullptr, "nsb_create: bad device index%s%s");
    nsb_ctx* ctx = new nsb_ctx();
    ctx->device = device;
    NSB_CUDA(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    NSB_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    {
        const char* e = getenv("NSB_CG");
        ctx->no_persistent_cg = e && !strcmp(e, "multi");
    }
    NSB_CUDA(ctx, cudaMalloc(&ctx->d_partials, sizeof(double) * NSB_MAX_PARTIALS * 8));
    NSB_CUDA(ctx, cudaMalloc(&ctx->d_scalars, sizeof(double) * 256));
    NSB_CUDA(ctx, cudaMalloc(&ctx->d_counter, sizeof(unsigned int) * 4));
    NSB_CUDA(ctx, cudaMemset(ctx->d_counter, 0, sizeof(unsigned int) * 4));
    NSB_CUDA(ctx, cudaMemset(ctx->d_scalars, 0, sizeof(double) * 256));
    NSB_CUDA(ctx, cudaMallocHost(&ctx->h_scalars, sizeof(double) * 256));
    *out = ctx;
    return 0;
}

extern "C" int nsb_destroy(nsb_ctx* ctx) {
    if (!ctx) return 0;
    cudaSetDevice(ctx->device);
    nsb_comm_destroy(ctx);
    cudaFree(ctx->d_partials);
    cudaFree(ctx->d_scalars);
    cudaFree(ctx->d_counter);
    cudaFreeHost(ctx->h_scalars);
    if (ctx->cg_ring) {
        cudaFreeHost(ctx->cg_ring);
        for (int i = 0; i < NSB_CG_RING; ++i) cudaEventDestroy(ctx->cg_ev[i]);
    }
    delete ctx;
    return 0;
}

extern "C" int nsb_set_mesh(nsb_ctx* ctx, int32_t dim, int32_t nverts, const double* coords,
                            int32_t ncells, const int32_t* cell_verts, int32_t nnodes,
                            const int32_t* cell_nodes) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, dim == 2 || dim == 3, "dim must be 2 or 3");
    NSB_REQUIRE(ctx, coords && cell_verts && cell_nodes && nverts > 0 && ncells > 0 && nnodes >= nverts,
                "bad mesh arrays");
    ctx->dim = dim; ctx->nverts = nverts; ctx->ncells = ncells; ctx->nnodes = nnodes;
    ctx->coords = coords; ctx->cell_verts = cell_verts; ctx->cell_nodes = cell_nodes;
    return 0;
}

extern "C" int nsb_set_pattern(nsb_ctx* ctx, int32_t which, int32_t nrows, const int32_t* rowptr,
                               const int32_t* col, const int32_t* pair_ptr, const int32_t* pair_cell,
                               const uint8_t* pair_iloc, const uint8_t* pair_slots) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which >= 0 && which < 3, "unknown pattern id");
    NSB_REQUIRE(ctx, nrows > 0 && rowptr && col && pair_ptr && pair_cell && pair_iloc && pair_slots,
                "bad pattern arrays");
    nsb_pattern_dev& P = ctx->pat[which];
    P.nrows = nrows; P.rowptr = rowptr; P.col = col; P.pair_ptr = pair_ptr; P.pair_cell = pair_cell;
    P.pair_iloc = pair_iloc; P.pair_slots = pair_slots;
    return 0;
}

extern "C" int nsb_set_schedule(nsb_ctx* ctx, int32_t which, int32_t ngroups, int32_t group_max_entries,
                                const int32_t* group_row, const int32_t* chunk_first, const int32_t* round_ptr,
                                const uint16_t* sched, const int32_t* fold_ptr, const uint32_t* fold) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which >= 0 && which < 3, "unknown pattern id");
    NSB_REQUIRE(ctx, ngroups > 0 && group_max_entries > 0 && group_row && chunk_first && round_ptr && sched && fold_ptr,
                "bad schedule arrays");
    nsb_pattern_dev& P = ctx->pat[which];
    NSB_REQUIRE(ctx, P.nrows > 0, "set the pattern before its schedule");
    P.ngroups = ngroups; P.group_max_entries = group_max_entries; P.group_row = group_row;
    P.chunk_first = chunk_first; P.round_ptr = round_ptr; P.sched = sched;
    P.fold_ptr = fold_ptr; P.fold = fold;
    return 0;
}

extern "C" int nsb_set_scratch(nsb_ctx* ctx, double* scratch, int64_t ndoubles) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, (scratch == nullptr) == (ndoubles == 0) && ndoubles >= 0, "bad scratch");
    ctx->scratch = scratch; ctx->scratch_doubles = ndoubles;
    return 0;
}

extern "C" int nsb_set_mixed_dofmap(nsb_ctx* ctx, int32_t ndofs, const int32_t* cell_dofs) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, ndofs > 0 && cell_dofs, "bad mixed dofmap");
    ctx->ndofs_w = ndofs; ctx->cell_dofs_w = cell_dofs;
    return 0;
}

extern "C" int nsb_set_mixed_bc(nsb_ctx* ctx, const uint8_t* mask, const double* g) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    ctx->w_bc_mask = mask; ctx->w_bc_g = g;
    return 0;
}

extern "C" int nsb_set_coefficients(nsb_ctx* ctx, const double* c) {
    NSB_REQUIRE(ctx, ctx != nullptr && c != nullptr, "null argument");
    NSB_REQUIRE(ctx, c[0] > 0.0, "step size k must be positive");
    NSB_REQUIRE(ctx, c[1] != 0.0, "alpha_0 must be non-zero");
    ctx->coef = {c[0], c[1], c[2], c[3], c[4], c[5], c[6], c[7]};
    return 0;
}

// ---- K9 Dirichlet rows ---------------------------------------------------------------------
// DirichletBC.apply(A): dolfin zeroes the row (pattern kept) and puts 1.0 on the diagonal;
// columns are untouched.  One 8-lane group per BC dof.
__global__ void k_bc_rows_csr(int32_t nbc, const int32_t* __restrict__ dofs, const int32_t* __restrict__ rowptr,
                              const int32_t* __restrict__ col, double* __restrict__ vals) {
    const int gid = (blockIdx.x * blockDim.x + threadIdx.x) >> 3, lane = threadIdx.x & 7;
    if (gid >= nbc) return;
    const int32_t r = dofs[gid];
    for (int32_t s = rowptr[r] + lane; s < rowptr[r + 1]; s += 8) vals[s] = (col[s] == r) ? 1.0 : 0.0;
}

__global__ void k_bc_rows_bsr(int32_t nbc, int bs, const int32_t* __restrict__ dofs,
                              const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                              double* __restrict__ vals) {
    const int gid = (blockIdx.x * blockDim.x + threadIdx.x) >> 3, lane = threadIdx.x & 7;
    if (gid >= nbc) return;
    const int32_t dof = dofs[gid];
    const int32_t node = dof / bs, comp = dof % bs;
    const int nb = rowptr[node + 1] - rowptr[node];
    const int64_t base = (int64_t)rowptr[node] * bs * bs;
    for (int e = lane; e < nb * bs; e += 8) {
        const int blk = e / bs, a = e % bs;
        vals[base + (int64_t)blk * bs * bs + comp * bs + a] =
            (col[rowptr[node] + blk] == node && a == comp) ? 1.0 : 0.0;
    }
}

extern "C" int nsb_bc_apply_matrix(nsb_ctx* ctx, int32_t kind, int32_t bs, const int32_t* rowptr,
                                   const int32_t* col, double* vals, int32_t nbc, const int32_t* bc_dofs,
                                   void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    if (nbc == 0) return 0;
    NSB_REQUIRE(ctx, rowptr && col && vals && bc_dofs && nbc > 0 && bs >= 1, "bad arguments");
    const int T = 256, B = nsb_blocks((int64_t)nbc * 8, T);
    NSB_TIMED(ctx, NSB_T_BC, (cudaStream_t)stream);
    if (kind == NSB_CSR) {
        NSB_REQUIRE(ctx, bs == 1, "CSR Dirichlet rows need an explicit (bs = 1) matrix; use bc_mask for Kronecker CSR");
        k_bc_rows_csr<<<B, T, 0, (cudaStream_t)stream>>>(nbc, bc_dofs, rowptr, col, vals);
    } else {
        k_bc_rows_bsr<<<B, T, 0, (cudaStream_t)stream>>>(nbc, bs, bc_dofs, rowptr, col, vals);
    }
    NSB_LAUNCH_CHECK(ctx, "k_bc_rows");
    return 0;
}

__global__ void k_bc_vector(int32_t nbc, int mode, const int32_t* __restrict__ dofs,
                            const double* __restrict__ g, const double* __restrict__ x, double* __restrict__ b) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nbc) return;
    const int32_t d = dofs[i];
    b[d] = (mode == NSB_BC_NEWTON) ? x[d] - g[i] : g[i];
}

extern "C" int nsb_bc_apply_vector(nsb_ctx* ctx, int32_t mode, double* b, const double* x, int32_t nbc,
                                   const int32_t* bc_dofs, const double* bc_vals, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    if (nbc == 0) return 0;
    NSB_REQUIRE(ctx, b && bc_dofs && bc_vals && nbc > 0, "bad arguments");
    NSB_REQUIRE(ctx, mode == NSB_BC_LINEAR || (mode == NSB_BC_NEWTON && x), "Newton mode needs x");
    NSB_TIMED(ctx, NSB_T_BC, (cudaStream_t)stream);
    k_bc_vector<<<nsb_blocks(nbc, 256), 256, 0, (cudaStream_t)stream>>>(nbc, mode, bc_dofs, bc_vals, x, b);
    NSB_LAUNCH_CHECK(ctx, "k_bc_vector");
    return 0;
}

// ---- FP64 FMA peak probe: roofline denominator for the element kernels -----------------------
__global__ void __launch_bounds__(256) k_fp64_peak(double* out, int iters) {
    double a0 = threadIdx.x * 1e-9, a1 = a0 + 1., a2 = a0 + 2., a3 = a0 + 3., a4 = a0 + 4., a5 = a0 + 5.,
           a6 = a0 + 6., a7 = a0 + 7.;
    const double m = 1.0000001, c = 1e-9;
    for (int i = 0; i < iters; ++i) {
        a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
        a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    }
    const double s = a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7;
    if (s == 12345.678) out[0] = s;
}

extern "C" int nsb_fp64_peak(nsb_ctx* ctx, double* tflops_out) {
    NSB_REQUIRE(ctx, ctx != nullptr && tflops_out != nullptr, "null argument");
    const int iters = 1 << 14, T = 256, B = ctx->sm_count * 8;
    cudaEvent_t e0, e1;
    NSB_CUDA(ctx, cudaEventCreate(&e0));
    NSB_CUDA(ctx, cudaEventCreate(&e1));
    k_fp64_peak<<<B, T>>>(ctx->d_partials, iters);
    NSB_CUDA(ctx, cudaDeviceSynchronize());
    float best = 1e30f;
    for (int r = 0; r < 5; ++r) {
        NSB_CUDA(ctx, cudaEventRecord(e0));
        k_fp64_peak<<<B, T>>>(ctx->d_partials, iters);
        NSB_CUDA(ctx, cudaEventRecord(e1));
        NSB_CUDA(ctx, cudaEventSynchronize(e1));
        float ms;
        NSB_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
        if (ms < best) best = ms;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    const double flops = 2.0 * 8.0 * (double)iters * (double)T * (double)B;
    *tflops_out = flops / (best * 1e-3) / 1e12;
    return 0;
}

// ---- instrumentation: per-class CUDA-event timers and the launch counter -----------------------
int nsb_timer_flush(nsb_ctx* ctx) {
    for (int c = 0; c < NSB_T_NCLASS; ++c) {
        nsb_timer_class& t = ctx->timers[c];
        for (size_t i = 0; i + 1 < t.used; i += 2) {
            if (cudaEventSynchronize(t.ev[i + 1]) != cudaSuccess) return -1;
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, t.ev[i], t.ev[i + 1]) != cudaSuccess) return -1;
            t.ms += ms;
        }
        t.used = 0;
    }
    return 0;
}

extern "C" int nsb_timer_enable(nsb_ctx* ctx, int32_t enable) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    if (nsb_timer_flush(ctx)) return nsb_fail(ctx, "timer flush failed%s%s");
    ctx->timing = enable != 0;
    ctx->timing_mask = enable == 1 ? 0xffffffffu : ((uint32_t)enable >> 1);   // 1: every class; else bit (c+1) selects class c
    return 0;
}

extern "C" int nsb_timer_report(nsb_ctx* ctx, double* ms_out, int64_t* regions_out) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    if (nsb_timer_flush(ctx)) return nsb_fail(ctx, "timer flush failed%s%s");
    for (int c = 0; c < NSB_T_NCLASS; ++c) {
        if (ms_out) ms_out[c] = ctx->timers[c].ms;
        if (regions_out) regions_out[c] = ctx->timers[c].regions;
        ctx->timers[c].ms = 0.0;
        ctx->timers[c].regions = 0;
    }
    return 0;
}

extern "C" int nsb_launch_count(nsb_ctx* ctx, int64_t* count_out) {
    NSB_REQUIRE(ctx, ctx != nullptr && count_out != nullptr, "null argument");
    *count_out = ctx->launch_count;
    return 0;
}

// doubles per cell of the caller-owned scratch of the two-phase residual assembly: NL2 * d
extern "C" int nsb_scratch_doubles_per_cell(int32_t dim) { return dim == 2 ? 12 : 30; }
