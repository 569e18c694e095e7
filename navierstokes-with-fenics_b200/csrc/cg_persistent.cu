// cg_persistent.cu -- the whole preconditioned CG solve of a scalar CSR system as ONE persistent cooperative
// kernel (K10 + K11 + K12 fused): the pressure Poisson problem of source/ns_ipcs_solver.py:166-171,205.
//
// Why: at C5 that system has 405 224 rows and a 79.5 MB operator -- it fits the 126 MB L2 of a B200 whole --
// and a solve is ~80 CG iterations x (1 + 7 polynomial) SpMVs x a handful of vector kernels.  As separate
// launches every SpMV costs 21-25 us against an L2-resident floor of a few us, and the ~1 200 launches per
// solve bound the step (19 % of the step in `spmv_csr`, 12 % in `krylov_vec`).  Here one CTA per
// SM owns a contiguous row range for the whole solve; the phases of an iteration are separated by
// grid-wide barriers (cooperative launch), every inner product is a per-CTA partial + one barrier after which
// each CTA sums the partials in CTA order (deterministic, bit-identical on every CTA, so the convergence
// decision is uniform without a broadcast), the Chebyshev recurrence and all CG vector updates live in the
// SpMV epilogues.  degree + 2 barriers per iteration, no host round trip before the solve ends.
//
// Memory-model notes: vectors written inside the kernel are read with ordinary (L1-cacheable) loads: the grid
// barrier is a gpu-scope release / acquire (thread 0 of every CTA fences after the spin, which invalidates the
// SM's L1, and the CTA barrier that follows orders the other threads behind it), exactly the guarantee
// grid.sync() documents.  That matters: a gathered x_j is used by ~15 rows, nearly all of them in the same
// CTA, so the gathers hit L1 -- with L2-only loads (ld.global.cg, the first version) every 8-byte gather
// cost a 32-byte sector request and the SM <-> L2 port bounded the SpMV phase at ~15 us (measured: 206 us per
// iteration).  Only truly read-only data (matrix, D^-1, b) goes through the non-coherent path (__ldg).
#include <cooperative_groups.h>
#include <math.h>
#include "common.cuh"

namespace cg = cooperative_groups;

#define PCG_THREADS 1024
#define PCG_LANES 4
#define PCG_MAXDEG 32

struct pcg_params {
    int32_t nrows;
    const int32_t* rowptr;
    const int32_t* col;
    const double* vals;
    // SELL-32 copy (nsb_sell): one thread per row, coalesced streams.  s_vals32 / s_col16: the 6-bytes-per-entry copy
    // read ONLY by the SpMVs inside the Chebyshev polynomial
    const int32_t* s_ptr;
    const int32_t* s_col;
    const double* s_vals;
    const float* s_vals32;
    const int16_t* s_col16;
    const uint8_t* mask;
    const double* dinv;
    const double* b;
    double *x, *r, *z, *p, *q, *rc, *d0, *d1;
    double* partials;          // [6][gridDim.x]
    double* out;               // [4]: iterations, ||r||^2, ||b||^2
    unsigned int* barrier;     // grid barrier counter, zero on entry
    double rtol, atol, n_all, inv_theta;
    int32_t maxit, nullspace, degree;     // degree 0: Jacobi; >= 2: Chebyshev polynomial with degree terms
    double ca[PCG_MAXDEG], cb[PCG_MAXDEG];
};

namespace {

__device__ __forceinline__ double ldv(const double* p) { return *p; }

// Grid-wide barrier on one monotonically increasing counter (zeroed before the launch; all CTAs are co-resident:
// cooperative launch).  cooperative_groups' grid.sync() measured ~8-10 us per barrier here (1024-thread CTAs on 148
// SMs) -- with degree + 2 barriers per CG iteration that was most of the solve.  Release: every thread's writes are
// ordered before thread 0's fence by the CTA barrier; acquire: thread 0 spins with ld.acquire.gpu, fences (which
// invalidates the SM's L1) and the second CTA barrier orders the other threads' ordinary loads behind it.
struct GridBarrier {
    unsigned int* counter;
    unsigned int target;
    __device__ __forceinline__ void sync() {
        __syncthreads();
        if (threadIdx.x == 0) {
            target += gridDim.x;
            __threadfence();
            atomicAdd(counter, 1u);
            unsigned int v;
            do {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(counter) : "memory");
            } while ((int)(v - target) < 0);
            __threadfence();
        }
        __syncthreads();
    }
};

// per-CTA partials -> barrier -> every CTA sums all partials in CTA order
template <int NV>
__device__ __forceinline__ void grid_reduce(GridBarrier& grid, double (&v)[NV], double* slots, double* smem) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    constexpr int NW = PCG_THREADS / 32;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    __syncthreads();                       // smem reuse
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) smem[i * NW + w] = v[i];
    }
    __syncthreads();
    if (w == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            double s = lane < NW ? smem[i * NW + lane] : 0.0;
            s = warp_sum(s);
            if (lane == 0) slots[(size_t)i * gridDim.x + blockIdx.x] = s;
        }
    }
    grid.sync();
    if (w == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            double s = 0.0;
            for (int j = lane; j < (int)gridDim.x; j += 32) s += ldv(slots + (size_t)i * gridDim.x + j);
            s = warp_sum(s);
            if (lane == 0) smem[NV * NW + i] = s;
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = smem[NV * NW + i];
}

// One SpMV phase over the CTA's rows [lo, hi): PCG_LANES lanes per row, PCG_THREADS / PCG_LANES rows per pass.
// The pass is software-pipelined: while the x values of pass k are gathered, the value / column entries of pass
// k+1 and the row offsets of pass k+2 are already in flight, and so are the vector entries the epilogue of pass k
// needs (pre()), so that a pass costs ONE memory latency instead of the chain offsets -> entries -> gather ->
// epilogue operands.  VT = double | float (values), IT = int32_t (absolute columns) | int16_t (col - row).
// post(row, (A x)_row, pre-loaded operands) runs in lane 0 of the row's group.
template <typename VT, typename IT>
struct RowEntries {
    VT v[4];
    int32_t c[4];
};

template <typename VT, typename IT>
__device__ __forceinline__ void load_entries(const VT* __restrict__ vals, const IT* __restrict__ cols, int32_t row,
                                             int32_t safe, int32_t beg, int32_t end, int lane, RowEntries<VT, IT>& e) {
    // `safe`: a valid column for the lanes without an entry (their gather is issued anyway, with value 0)
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int32_t s = beg + lane + u * PCG_LANES;
        const bool ok = s < end;
        e.v[u] = ok ? __ldg(vals + s) : (VT)0;
        if (sizeof(IT) == 2) e.c[u] = ok ? row + (int32_t)__ldg(cols + s) : safe;
        else e.c[u] = ok ? (int32_t)__ldg(cols + s) : safe;
    }
}

template <typename VT, typename IT, typename Pre, typename Post>
__device__ __forceinline__ void spmv_phase(const int32_t* __restrict__ rowptr, const VT* __restrict__ vals,
                                           const IT* __restrict__ cols, const double* x, int32_t lo, int32_t hi, int grp,
                                           int lane, Pre&& pre, Post&& post) {
    constexpr int GROUPS = PCG_THREADS / PCG_LANES;
    const int npass = (hi - lo + GROUPS - 1) / GROUPS;
    auto offsets = [&](int k, int32_t& b, int32_t& e) {
        const int32_t row = lo + grp + k * GROUPS;
        if (k < npass && row < hi) { b = __ldg(rowptr + row); e = __ldg(rowptr + row + 1); }
        else { b = 0; e = 0; }
    };
    int32_t beg_c, end_c, beg_n, end_n;
    offsets(0, beg_c, end_c);
    offsets(1, beg_n, end_n);
    RowEntries<VT, IT> cur, nxt;
    load_entries<VT, IT>(vals, cols, lo + grp, lo, beg_c, end_c, lane, cur);
    for (int k = 0; k < npass; ++k) {
        const int32_t row = lo + grp + k * GROUPS;
        const bool mine = row < hi && lane == 0;
        // everything issued here is independent: one round trip per pass
        double xv[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) xv[u] = ldv(x + cur.c[u]);
        auto operands = pre(row, mine);
        load_entries<VT, IT>(vals, cols, row + GROUPS, lo, beg_n, end_n, lane, nxt);
        int32_t beg_nn, end_nn;
        offsets(k + 2, beg_nn, end_nn);
        double acc = 0.0;
#pragma unroll
        for (int u = 0; u < 4; ++u) acc += (double)cur.v[u] * xv[u];
        // rows longer than 4 * PCG_LANES entries (never on P1 meshes of bounded valence): plain tail loop
        for (int32_t s = beg_c + lane + 4 * PCG_LANES; s < end_c; s += PCG_LANES) {
            const int32_t cn = sizeof(IT) == 2 ? row + (int32_t)__ldg(cols + s) : (int32_t)__ldg(cols + s);
            acc += (double)__ldg(vals + s) * ldv(x + cn);
        }
        __syncwarp();
        acc = group_sum<PCG_LANES>(acc);
        if (mine) post(row, acc, operands);
        cur = nxt;
        beg_c = beg_n; end_c = end_n;
        beg_n = beg_nn; end_n = end_nn;
    }
}

// One SpMV phase in the SELL-32 layout: thread = row, warp = slice, every load of the value / column streams is a
// fully coalesced 32-lane access; ~6 instructions per entry instead of the ~35 of the 4-lanes-per-row CSR pass
// (which was ISSUE bound: measured 1 us per 256-row pass).  rows [lo, hi), lo a multiple of 32.
template <typename VT, typename IT, typename Pre, typename Post>
__device__ __forceinline__ void sell_phase(const int32_t* __restrict__ sptr, const VT* __restrict__ vals,
                                           const IT* __restrict__ cols, const double* x, int32_t lo, int32_t hi,
                                           Pre&& pre, Post&& post) {
    for (int32_t row = lo + (int32_t)threadIdx.x; row < hi; row += PCG_THREADS) {
        const int32_t slice = row >> 5, l = row & 31;
        const int32_t base = __ldg(sptr + slice), width = (__ldg(sptr + slice + 1) - base) >> 5;
        auto operands = pre(row, true);
        const VT* v = vals + base + l;
        const IT* c = cols + base + l;
        double acc0 = 0.0, acc1 = 0.0;
        int k = 0;
        // 8 entries in flight per thread: a row of 15 is two dependent round trips (columns -> gathers), not eight
        for (; k + 8 <= width; k += 8) {
            VT vv[8];
            int32_t cc[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                vv[u] = __ldg(v + (k + u) * 32);
                cc[u] = sizeof(IT) == 2 ? row + (int32_t)__ldg(c + (k + u) * 32) : (int32_t)__ldg(c + (k + u) * 32);
            }
            double xv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) xv[u] = ldv(x + cc[u]);
#pragma unroll
            for (int u = 0; u < 8; u += 2) { acc0 += (double)vv[u] * xv[u]; acc1 += (double)vv[u + 1] * xv[u + 1]; }
        }
        {   // remaining (< 8) entries, still issued together
            VT vv[8];
            int32_t cc[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const bool ok = k + u < width;
                vv[u] = ok ? __ldg(v + (k + u) * 32) : (VT)0;
                cc[u] = ok ? (sizeof(IT) == 2 ? row + (int32_t)__ldg(c + (k + u) * 32) : (int32_t)__ldg(c + (k + u) * 32)) : row;
            }
            double xv[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) xv[u] = ldv(x + cc[u]);
#pragma unroll
            for (int u = 0; u < 8; u += 2) { acc0 += (double)vv[u] * xv[u]; acc1 += (double)vv[u + 1] * xv[u + 1]; }
        }
        post(row, acc0 + acc1, operands);
    }
}

}  // namespace

__global__ void __launch_bounds__(PCG_THREADS, 1) k_pcg_persistent(const pcg_params P) {
    GridBarrier grid = {P.barrier, 0u};
    __shared__ double smem[3 * (PCG_THREADS / 32) + 4];
    const int32_t n = P.nrows;
    const bool sell = P.s_ptr != nullptr;
    int32_t per = (n + (int32_t)gridDim.x - 1) / (int32_t)gridDim.x;
    if (sell) per = (per + 31) & ~31;                  // whole slices per CTA
    const int32_t lo = min(n, (int32_t)blockIdx.x * per), hi = min(n, lo + per);
    const int lane = threadIdx.x % PCG_LANES, grp = threadIdx.x / PCG_LANES;
    double* slot_a = P.partials;                       // 3 values
    double* slot_b = P.partials + 3 * (size_t)gridDim.x;   // 3 values (alternates with slot_a)
    const bool cheb = P.degree >= 2;

    // ---- zero-mean projection of the right-hand side (constant null space) --------------------------
    double shift = 0.0;
    if (P.nullspace) {
        double s[1] = {0.0};
        for (int32_t i = lo + threadIdx.x; i < hi; i += PCG_THREADS) s[0] += __ldg(P.b + i);
        grid_reduce<1>(grid, s, slot_a, smem);
        shift = s[0] / P.n_all;
    }
    // ---- r = (b - shift) - A x ; preconditioner start for the first iteration --------------------------
    double red[3] = {0.0, 0.0, 0.0};                   // rr, bb, rz (Jacobi)
    struct Op0 { double xi, bi, dinv; };
    {
        auto pre0 = [&](int32_t row, bool mine) {
            Op0 o = {0.0, 0.0, 0.0};
            if (mine) { o.xi = ldv(P.x + row); o.bi = __ldg(P.b + row); o.dinv = __ldg(P.dinv + row); }
            return o;
        };
        auto post0 = [&](int32_t row, double v, const Op0& o) {
            if (P.mask != nullptr && P.mask[row]) v = o.xi;
            const double bi = o.bi - shift;
            const double ri = bi - v;
            P.r[row] = ri;
            red[0] += ri * ri;
            red[1] += bi * bi;
            const double di = o.dinv * ri;
            if (cheb) {
                P.rc[row] = di;
                const double d = di * P.inv_theta;
                P.d0[row] = d;
                P.z[row] = d;
            } else {
                P.z[row] = di;
                red[2] += ri * di;
            }
        };
        if (sell) sell_phase<double, int32_t>(P.s_ptr, P.s_vals, P.s_col, P.x, lo, hi, pre0, post0);
        else spmv_phase<double, int32_t>(P.rowptr, P.vals, P.col, P.x, lo, hi, grp, lane, pre0, post0);
    }
    grid_reduce<3>(grid, red, slot_b, smem);
    double rr = red[0];
    const double bb = red[1];
    double rz_next = red[2];
    const double tol = fmax(P.rtol * sqrt(bb), P.atol);
    const double tol2 = tol * tol;
    double rz_old = 0.0;
    int it = 0;
    while (rr > tol2 && it < P.maxit) {
        double rz;
        if (cheb) {
            // z = p_degree(D^-1 A) D^-1 r : degree-1 SpMVs, each with the recurrence in its epilogue
            double* dold = P.d0;
            double* dnew = P.d1;
            double s[1] = {0.0};
            for (int k = 1; k < P.degree; ++k) {
                const double a = P.ca[k], bcoef = P.cb[k];
                const bool last = k == P.degree - 1;
                struct OpC { double dk, rc, dinv, z, r; };
                auto pre = [&](int32_t row, bool mine) {
                    OpC o = {0.0, 0.0, 0.0, 0.0, 0.0};
                    if (mine) {
                        o.dk = ldv(dold + row); o.rc = ldv(P.rc + row); o.dinv = __ldg(P.dinv + row); o.z = ldv(P.z + row);
                        if (last) o.r = ldv(P.r + row);
                    }
                    return o;
                };
                auto post = [&](int32_t row, double v, const OpC& o) {
                    if (P.mask != nullptr && P.mask[row]) v = o.dk;
                    const double rci = o.rc - o.dinv * v;
                    const double dn = a * o.dk + bcoef * rci;
                    const double zi = o.z + dn;
                    P.rc[row] = rci;
                    dnew[row] = dn;
                    P.z[row] = zi;
                    if (last) s[0] += o.r * zi;
                };
                if (sell && P.s_vals32 != nullptr && P.s_col16 != nullptr)
                    sell_phase<float, int16_t>(P.s_ptr, P.s_vals32, P.s_col16, dold, lo, hi, pre, post);
                else if (sell && P.s_vals32 != nullptr)
                    sell_phase<float, int32_t>(P.s_ptr, P.s_vals32, P.s_col, dold, lo, hi, pre, post);
                else if (sell)
                    sell_phase<double, int32_t>(P.s_ptr, P.s_vals, P.s_col, dold, lo, hi, pre, post);
                else
                    spmv_phase<double, int32_t>(P.rowptr, P.vals, P.col, dold, lo, hi, grp, lane, pre, post);
                if (!last) grid.sync();
                double* t = dold; dold = dnew; dnew = t;
            }
            grid_reduce<1>(grid, s, slot_a, smem);
            rz = s[0];
        } else {
            rz = rz_next;
        }
        // p = z + beta p
        const double beta = (it == 0 || rz_old == 0.0) ? 0.0 : rz / rz_old;
        for (int32_t i = lo + threadIdx.x; i < hi; i += PCG_THREADS) {
            const double zi = ldv(P.z + i);
            P.p[i] = it == 0 ? zi : zi + beta * ldv(P.p + i);
        }
        grid.sync();
        // q = A p ; p.q
        double s1[1] = {0.0};
        struct OpQ { double pi; };
        {
            auto preq = [&](int32_t row, bool mine) {
                OpQ o = {0.0};
                if (mine) o.pi = ldv(P.p + row);
                return o;
            };
            auto postq = [&](int32_t row, double v, const OpQ& o) {
                if (P.mask != nullptr && P.mask[row]) v = o.pi;
                P.q[row] = v;
                s1[0] += o.pi * v;
            };
            if (sell) sell_phase<double, int32_t>(P.s_ptr, P.s_vals, P.s_col, P.p, lo, hi, preq, postq);
            else spmv_phase<double, int32_t>(P.rowptr, P.vals, P.col, P.p, lo, hi, grp, lane, preq, postq);
        }
        grid_reduce<1>(grid, s1, slot_b, smem);
        const double pq = s1[0];
        const double alpha = pq != 0.0 ? rz / pq : 0.0;
        // x += alpha p ; r -= alpha q ; r.r ; preconditioner start of the next iteration
        double s2[2] = {0.0, 0.0};
        for (int32_t i = lo + threadIdx.x; i < hi; i += PCG_THREADS) {
            P.x[i] = ldv(P.x + i) + alpha * ldv(P.p + i);
            const double ri = ldv(P.r + i) - alpha * ldv(P.q + i);
            P.r[i] = ri;
            s2[0] += ri * ri;
            const double di = __ldg(P.dinv + i) * ri;
            if (cheb) {
                P.rc[i] = di;
                const double d = di * P.inv_theta;
                P.d0[i] = d;
                P.z[i] = d;
            } else {
                P.z[i] = di;
                s2[1] += ri * di;
            }
        }
        grid_reduce<2>(grid, s2, slot_a, smem);
        rr = s2[0];
        rz_next = s2[1];
        rz_old = rz;
        ++it;
        if (!(rr == rr)) break;            // NaN: uniform on every CTA
    }
    if (P.nullspace) {
        double s[1] = {0.0};
        for (int32_t i = lo + threadIdx.x; i < hi; i += PCG_THREADS) s[0] += ldv(P.x + i);
        grid_reduce<1>(grid, s, slot_b, smem);
        const double mean = s[0] / P.n_all;
        for (int32_t i = lo + threadIdx.x; i < hi; i += PCG_THREADS) P.x[i] = ldv(P.x + i) - mean;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        P.out[0] = (double)it;
        P.out[1] = rr;
        P.out[2] = bb;
    }
}

// 0 = solved (or failed to converge: status via the outputs), 1 = not applicable (caller keeps the multi-kernel
// path), -1 = error.  work: 7 n doubles (nsb_cg_workspace_doubles).
int nsb_cg_persistent(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* b, double* x, double rtol,
                      double atol, int32_t maxit, int32_t nullspace, double* work, int32_t* iters_out, double* resid_out,
                      cudaStream_t st) {
    if (A->kind != NSB_CSR || A->bs != 1) return 1;
    if (ctx->nranks > 1 && A->halo != NSB_HALO_NONE) return 1;           // partitioned operator: needs halo exchanges
    if (!(M->kind == NSB_PC_JACOBI || (M->kind == NSB_PC_CHEBYSHEV && M->degree >= 2 && M->degree <= PCG_MAXDEG))) return 1;
    static int coop = -1, per_sm = 0;
    if (coop < 0) {
        int v = 0;
        cudaDeviceGetAttribute(&v, cudaDevAttrCooperativeLaunch, ctx->device);
        coop = v;
        if (coop) {
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_pcg_persistent, PCG_THREADS, 0) != cudaSuccess) per_sm = 0;
        }
    }
    if (!coop || per_sm < 1) return 1;
    const int64_t n = A->nrows;
    int grid = ctx->sm_count;                       // one CTA per SM: the barrier cost grows with the CTA count
    const int64_t need = (n + 255) / 256;           // at least 256 rows per CTA
    if (grid > need) grid = (int)(need < 1 ? 1 : need);
    if (6 * (int64_t)grid > NSB_MAX_PARTIALS * 8) return 1;
    pcg_params P;
    memset(&P, 0, sizeof P);
    P.nrows = A->nrows; P.rowptr = A->rowptr; P.col = A->col; P.vals = A->vals; P.mask = A->bc_mask;
    if (A->sell && A->sell->slice_ptr && A->sell->col && A->sell->vals && A->sell->nslices == (A->nrows + 31) / 32) {
        P.s_ptr = A->sell->slice_ptr; P.s_col = A->sell->col; P.s_vals = A->sell->vals;
        P.s_vals32 = A->sell->vals32; P.s_col16 = A->sell->col16;
    }
    P.dinv = M->inv_diag; P.b = b; P.x = x;
    P.r = work; P.z = work + n; P.p = work + 2 * n; P.q = work + 3 * n; P.rc = work + 4 * n; P.d0 = work + 5 * n; P.d1 = work + 6 * n;
    P.partials = ctx->d_partials;
    P.out = ctx->d_scalars;
    P.barrier = ctx->d_counter + 2;
    NSB_CUDA(ctx, cudaMemsetAsync(P.barrier, 0, sizeof(unsigned int), st));
    P.rtol = rtol; P.atol = atol; P.maxit = maxit; P.nullspace = nullspace;
    P.n_all = (double)n;
    P.degree = M->kind == NSB_PC_CHEBYSHEV ? M->degree : 0;
    if (P.degree) {
        const double theta = 0.5 * (M->lambda_max + M->lambda_min), delta = 0.5 * (M->lambda_max - M->lambda_min);
        const double sigma1 = theta / delta;
        double rho_old = 1.0 / sigma1;
        P.inv_theta = 1.0 / theta;
        for (int k = 1; k < P.degree; ++k) {
            const double rho = 1.0 / (2.0 * sigma1 - rho_old);
            P.ca[k] = rho * rho_old;
            P.cb[k] = 2.0 * rho / delta;
            rho_old = rho;
        }
    }
    void* args[] = {(void*)&P};
    {
        NSB_TIMED(ctx, NSB_T_CG_FUSED, st);
        NSB_CUDA(ctx, cudaLaunchCooperativeKernel((const void*)k_pcg_persistent, dim3(grid), dim3(PCG_THREADS), args, 0, st));
    }
    NSB_LAUNCH_CHECK(ctx, "k_pcg_persistent");
    NSB_CUDA(ctx, cudaMemcpyAsync(ctx->h_scalars, ctx->d_scalars, sizeof(double) * 3, cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    const int it = (int)ctx->h_scalars[0];
    const double rnorm = sqrt(ctx->h_scalars[1]), bnorm = sqrt(ctx->h_scalars[2]);
    if (iters_out) *iters_out = it;
    if (resid_out) *resid_out = rnorm;
    if (!(rnorm == rnorm)) return nsb_fail(ctx, "CG breakdown: residual is NaN%s%s");
    if (rnorm > fmax(rtol * bnorm, atol)) return nsb_fail(ctx, "CG did not converge within the iteration limit%s%s");
    return 0;
}

// SELL-32 fill: thread = row; entry k of the row goes to slice_ptr[row / 32] + 32 k + row % 32, padding gets value 0
// and the row's own index
__global__ void k_sell_fill(int32_t nrows, int32_t nslices, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                            const double* __restrict__ vals, const int32_t* __restrict__ sptr, int32_t* __restrict__ scol,
                            double* __restrict__ svals, float* __restrict__ svals32, int16_t* __restrict__ scol16, int* flag) {
    const int32_t row = blockIdx.x * blockDim.x + threadIdx.x;
    if (row >= nslices * 32) return;
    const int32_t slice = row >> 5, l = row & 31;
    const int32_t base = sptr[slice], width = (sptr[slice + 1] - base) >> 5;
    const int32_t beg = row < nrows ? rowptr[row] : 0, len = row < nrows ? rowptr[row + 1] - beg : 0;
    for (int k = 0; k < width; ++k) {
        const int32_t at = base + k * 32 + l;
        const bool real = k < len;
        const int32_t c = real ? col[beg + k] : min(row, nrows - 1);
        const double v = real ? vals[beg + k] : 0.0;
        scol[at] = c;
        svals[at] = v;
        if (svals32) svals32[at] = (float)v;
        if (scol16) {
            const int32_t d = c - row;
            if (d > 32767 || d < -32768) *flag = 1;
            scol16[at] = (int16_t)d;
        }
    }
}

extern "C" int nsb_sell_fill(nsb_ctx* ctx, const nsb_matrix* A, int32_t nslices, const int32_t* slice_ptr_dev, int32_t* col_out,
                             double* vals_out, float* vals32_out, int16_t* col16_out, int32_t* overflow_out, void* stream) {
    NSB_REQUIRE(ctx, ctx && A && A->kind == NSB_CSR && A->bs == 1 && slice_ptr_dev && col_out && vals_out && overflow_out,
                "bad arguments");
    NSB_REQUIRE(ctx, nslices == (A->nrows + 31) / 32, "nslices must be ceil(nrows / 32)");
    cudaStream_t st = (cudaStream_t)stream;
    int* flag = reinterpret_cast<int*>(ctx->d_counter + 1);
    NSB_CUDA(ctx, cudaMemsetAsync(flag, 0, sizeof(int), st));
    k_sell_fill<<<nsb_blocks((int64_t)nslices * 32, 256), 256, 0, st>>>(A->nrows, nslices, A->rowptr, A->col, A->vals, slice_ptr_dev,
                                                                       col_out, vals_out, vals32_out, col16_out, flag);
    NSB_LAUNCH_CHECK(ctx, "k_sell_fill");
    int h = 0;
    NSB_CUDA(ctx, cudaMemcpyAsync(&h, flag, sizeof(int), cudaMemcpyDeviceToHost, st));
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    *overflow_out = h;
    return 0;
}
