// common.cuh -- context, error plumbing and device helpers shared by the nsb200 kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <string>
#include <vector>
#include "../../include/nsb200.h"

#define NSB_MAX_PARTIALS 4096
#define NSB_CG_RING 8
#define NSB_CG_LAG 4

struct nsb_pattern_dev {
    int32_t nrows = 0;
    const int32_t* rowptr = nullptr;
    const int32_t* col = nullptr;
    const int32_t* pair_ptr = nullptr;
    const int32_t* pair_cell = nullptr;
    const uint8_t* pair_iloc = nullptr;
    const uint8_t* pair_slots = nullptr;
    // scatter schedule of the warp-cooperative matrix assembly (nsb_sched_build / nsb_set_schedule)
    int32_t ngroups = 0, group_max_entries = 0;
    const int32_t* group_row = nullptr;     // [ngroups+1]
    const int32_t* chunk_first = nullptr;   // [ngroups+1]
    const int32_t* round_ptr = nullptr;     // [nchunks+1]
    const uint16_t* sched = nullptr;        // [32 * nrounds]
    const int32_t* fold_ptr = nullptr;      // [ngroups+1]
    const uint32_t* fold = nullptr;         // [nfold] real entry << 16 | shadow slot
};

struct nsb_coef {
    double k, a0, a1, a2, c_conv, c_p, c_visc, c_body;
};

// kernel classes of the built-in CUDA-event timers (nsb_timer_*), see include/nsb200.h
struct nsb_timer_class {
    std::vector<cudaEvent_t> ev;   // start/stop pairs, recycled after every report
    size_t used = 0;               // events in use (2 per timed region)
    double ms = 0.0;               // accumulated device time since the last reset
    int64_t regions = 0;           // timed regions since the last reset
};

// owner -> ghost exchange plan of one index space (comm.cu, nsb_set_halo)
struct nsb_halo {
    int32_t n_owned = 0, n_ghost = 0;
    int64_t n_global = 0;
    std::vector<int32_t> peers, send_ptr, recv_ptr;   // host copies
    const int32_t* send_idx = nullptr;                // dev [send_ptr.back()] local owned indices
    double* sendbuf = nullptr;                        // dev [send_ptr.back() * 3], caller-owned
    double* ghostbuf = nullptr;                       // dev [n_ghost * 3], caller-owned (SpMV ghosts)
    // peer-memory path (comm.cu, nsb_p2p_*): a library-owned, IPC-shared block per rank = header of
    // uint64 acknowledgement words (ack[j][b] = newest of my epochs block b of neighbour j has consumed)
    // followed by a ring of NSB_P2P_RING receive slots of p2p_cap 16-byte lines {lo, epoch, hi, epoch},
    // each slot shaped like the ghost block (x3 components)
    void* p2p_block = nullptr;
    int64_t p2p_cap = 0;
    uint64_t p2p_epoch = 0;
    bool p2p_ready = false;
    std::vector<void*> peer_block;                    // mapped blocks of the neighbours
    std::vector<long long> peer_cap;                  // doubles per ring slot of neighbour i
    std::vector<int32_t> peer_off;                    // offset (entries) of my part inside neighbour i's ghost block
    std::vector<int32_t> my_slot;                     // my index in neighbour i's peer list
};
#define NSB_P2P_RING 4
#define NSB_P2P_HDR_BYTES 8192
#define NSB_AR_RING 4
#define NSB_AR_MAX 64

// ---- peer-memory lines (comm.cu): one double travels as a 16-byte line {lo, epoch, hi, epoch}; a reader
// polls until both flags carry the epoch it waits for (8-byte store atomicity is all this needs)
#ifdef __CUDACC__
__device__ __forceinline__ void nsb_st_line(uint4* p, double v, unsigned int flag) {
    const unsigned long long u = (unsigned long long)__double_as_longlong(v);
    asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"((unsigned int)u), "r"(flag),
                 "r"((unsigned int)(u >> 32)), "r"(flag) : "memory");
}
__device__ __forceinline__ uint4 nsb_ld_line(const uint4* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
#define NSB_P2P_SPIN_CYCLES 60000000000ll        // bounded waits: ~30 s, then the error word is raised
// wait for line p of epoch `flag`; on a timeout (or once the error word is set) return what is there
__device__ __forceinline__ double nsb_wait_line(const uint4* p, unsigned int flag, int* err, int code) {
    uint4 v = nsb_ld_line(p);
    if (v.y != flag || v.w != flag) {
        if (*(volatile int*)err == 0) {
            const long long t0 = clock64();
            do {
                if (clock64() - t0 > NSB_P2P_SPIN_CYCLES) { *err = code; break; }
                v = nsb_ld_line(p);
            } while (v.y != flag || v.w != flag);
        }
    }
    return __longlong_as_double((long long)(((unsigned long long)v.z << 32) | v.x));
}
#endif

// Chebyshev step fused into the epilogue of the single-precision block-scalar SpMV (linalg.cu
// k_spmv_kron32, krylov.cu apply_cheb32): with q = S d_old the row computes r -= D^-1 q,
// d_new = a d_old + b r, z += d_new (zout != null: last step, only zout = z + d_new is written, in FP64)
struct nsb_cheb32_epi {
    const double* dinv;
    float *r, *dnew, *z;
    double* zout;
    float a, b;
    // bf16 shadow of the direction vectors (BS = 3: 4 x bf16 = 8 bytes per node): the SpMV GATHERS from x_h instead of
    // the float4 vector (half the bytes per gathered node -- the kernel is bound by L1TEX sectors of exactly those
    // gathers) and writes the shadow of d_new next to d_new; the recurrences themselves stay FP32 on the exact vectors.
    // nullptr: gather from the float4 vector.
    const unsigned long long* x_h;
    unsigned long long* dnew_h;
    // dofs flagged here are identity rows of the operator, (A x)_i = x_i (nsb_matrix.bc_mask of the cached operators)
    const uint8_t* mask;
};
__device__ __forceinline__ unsigned long long nsb_pack_bf16x3(float a, float b, float c) {
    // round-to-nearest-even bf16 = upper 16 bits of the float
    auto rn = [](float f) -> unsigned int {
        unsigned int u = __float_as_uint(f);
        u += 0x7fffu + ((u >> 16) & 1u);
        return u >> 16;
    };
    const unsigned int lo = rn(a) | (rn(b) << 16), hi = rn(c);
    return ((unsigned long long)hi << 32) | lo;
}

// the same step in FP64 in the epilogue of the CSR SpMV kernels (pressure Chebyshev-CG): z += d_new in place
struct nsb_cheb64_epi {
    const double* dinv;
    double *r, *dnew, *z;
    double a, b;
};

struct nsb_ctx {
    int device = 0;
    const nsb_cheb64_epi* spmv_epi = nullptr;   // set by apply_pc for the very next CSR SpMV, consumed by its launcher
    // multi-GPU (comm.cu): one rank per GPU, NCCL communicator created by nsb_comm_init
    int rank = 0, nranks = 1;
    void* nccl_comm = nullptr;
    int* d_p2p_err = nullptr;         // device error word of the peer-memory kernels (0 = fine)
    // peer-memory all-reduce of the Krylov scalars (comm.cu): IPC block = ring of NSB_AR_RING slots x
    // nranks senders x NSB_AR_MAX 16-byte lines; ar_peer[r] = rank r's block (own block at [rank])
    void* ar_block = nullptr;
    std::vector<void*> ar_peer;
    uint64_t ar_epoch = 0;
    bool ar_ready = false;
    nsb_halo halo[3];                 // [NSB_HALO_NODES], [NSB_HALO_VERTS]; [0] unused
    int sm_count = 148;
    bool no_persistent_cg = false;    // NSB_CG=multi: keep the multi-kernel CG (A/B measurements)
    std::string err;
    // mesh
    int32_t dim = 0, nverts = 0, ncells = 0, nnodes = 0;
    const double* coords = nullptr;
    const int32_t* cell_verts = nullptr;
    const int32_t* cell_nodes = nullptr;
    nsb_pattern_dev pat[3];
    // mixed space
    int32_t ndofs_w = 0;
    const int32_t* cell_dofs_w = nullptr;
    const uint8_t* w_bc_mask = nullptr;
    const double* w_bc_g = nullptr;
    nsb_coef coef = {1.0, 1.0, -1.0, 0.0, 1.0, 1.0, 1.0, 1.0};
    // weak-form variants (nsb_set_forms): convective form, viscous form, Coriolis vector 2 c_Omega Omega
    int32_t form_conv = 0, form_visc = 0;
    double coriolis[3] = {0.0, 0.0, 0.0};
    bool force_generic_assembly = false;
    float* jac_block_scalar_out = nullptr;   // nsb_set_block_scalar_out: diag32 written alongside the momentum Jacobian
    int64_t jac_block_scalar_nnz = 0;
    // reduction scratch (device) + pinned host mirror
    double* d_partials = nullptr;     // [NSB_MAX_PARTIALS * 8]
    double* d_scalars = nullptr;      // [64] device-resident Krylov scalars
    unsigned int* d_counter = nullptr;
    double* h_scalars = nullptr;      // pinned [256]
    // look-ahead convergence check of CG (krylov.cu): pinned ring of residual norms + their events
    double* cg_ring = nullptr;
    cudaEvent_t cg_ev[8] = {};
    // caller-owned workspace of the two-phase (cell -> row) assembly kernels (nsb_set_scratch)
    double* scratch = nullptr;
    int64_t scratch_doubles = 0;
    // K16 library leg (direct.cu): accumulated wall time of the sparse direct solves, kept apart
    // from the kernel-class timers
    double direct_seconds = 0.0;
    int64_t direct_solves = 0;
    // instrumentation
    int64_t launch_count = 0;         // kernels launched by this context (all classes)
    bool timing = false;
    uint32_t timing_mask = 0xffffffffu;   // bit c: class c is timed (nsb_timer_enable)
    nsb_timer_class timers[NSB_T_NCLASS];
};

int nsb_timer_flush(nsb_ctx* ctx);
// cg_persistent.cu: 0 = solved, 1 = not applicable (caller keeps the multi-kernel path), -1 = error
int nsb_cg_persistent(nsb_ctx* ctx, const nsb_matrix* A, const nsb_precond* M, const double* b, double* x, double rtol,
                      double atol, int32_t maxit, int32_t nullspace, double* work, int32_t* iters_out, double* resid_out,
                      cudaStream_t st);
// comm.cu: no-ops (return 0) on a single rank
int nsb_allreduce_sum(nsb_ctx* ctx, double* dev_buf, int count, cudaStream_t st);
int nsb_comm_destroy(nsb_ctx* ctx);
// assemble_jac.cu: 0 = launched, 1 = not applicable (caller falls back), -1 = error
int nsb_launch_residual_cells(nsb_ctx* ctx, const double* ustar, const double* un, const double* unm1,
                              const double* pn, double* b_out, cudaStream_t st);   // assemble_res.cu
int nsb_launch_jacobian_rowgroups(nsb_ctx* ctx, const double* ustar, double mass_coef, int newton, double* A_out, float* S32_out,
                                  cudaStream_t st);   // context.cu: fold finished event pairs into timers[].ms

// RAII region: records a start/stop event pair on `st` around the launches of one kernel class
struct nsb_timed_region {
    nsb_ctx* ctx; int cls; cudaStream_t st; bool on;
    nsb_timed_region(nsb_ctx* c, int k, cudaStream_t s) : ctx(c), cls(k), st(s), on(c && c->timing && ((c->timing_mask >> k) & 1u)) {
        if (!on) return;
        nsb_timer_class& t = ctx->timers[cls];
        if (t.used + 2 > t.ev.size()) {
            if (t.ev.size() >= (1u << 16)) nsb_timer_flush(ctx);
            if (t.used + 2 > t.ev.size()) {
                cudaEvent_t a, b;
                cudaEventCreate(&a); cudaEventCreate(&b);
                t.ev.push_back(a); t.ev.push_back(b);
            }
        }
        cudaEventRecord(t.ev[t.used], st);
    }
    ~nsb_timed_region() {
        if (!on) return;
        nsb_timer_class& t = ctx->timers[cls];
        cudaEventRecord(t.ev[t.used + 1], st);
        t.used += 2;
        t.regions += 1;
    }
};
#define NSB_TIMED(ctx, cls, st) nsb_timed_region nsb_region_##cls(ctx, cls, st)

extern thread_local std::string g_nsb_last_error;

static inline int nsb_fail(nsb_ctx* ctx, const char* fmt, const char* a = "", const char* b = "") {
    char buf[512];
    snprintf(buf, sizeof buf, fmt, a, b);
    if (ctx) ctx->err = buf;
    g_nsb_last_error = buf;
    return -1;
}

#define NSB_CUDA(ctx, call)                                                        \
    do {                                                                           \
        cudaError_t e_ = (call);                                                   \
        if (e_ != cudaSuccess) return nsb_fail(ctx, "CUDA error %s at %s", cudaGetErrorString(e_), #call); \
    } while (0)

#define NSB_LAUNCH_CHECK(ctx, name)                                                \
    do {                                                                           \
        cudaError_t e_ = cudaGetLastError();                                       \
        if (e_ != cudaSuccess) return nsb_fail(ctx, "launch of %s failed: %s", name, cudaGetErrorString(e_)); \
        (ctx)->launch_count += 1;                                                  \
    } while (0)

// extra launches covered by one NSB_LAUNCH_CHECK
#define NSB_COUNT(ctx, n) ((ctx)->launch_count += (n))

#define NSB_REQUIRE(ctx, cond, msg)                                                \
    do {                                                                           \
        if (!(cond)) return nsb_fail(ctx, "%s (%s)", msg, #cond);                  \
    } while (0)

static inline int nsb_blocks(int64_t n, int threads) { return (int)((n + threads - 1) / threads); }

// ---- warp helpers ------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int LANES>
__device__ __forceinline__ double group_sum(double v) {
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Deterministic block reduction of NV values; result valid in thread 0.
template <int NV, int THREADS>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* smem /* [NV * THREADS/32] */) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
#pragma unroll
    for (int i = 0; i < NV; ++i) v[i] = warp_sum(v[i]);
    if (lane == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) smem[i * (THREADS / 32) + w] = v[i];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            double s = 0.0;
            for (int j = 0; j < THREADS / 32; ++j) s += smem[i * (THREADS / 32) + j];
            v[i] = s;
        }
    }
}

// Grid-wide deterministic reduction: every block writes its partials, the last block to
// finish sums them in block order and stores into out[0..NV).  `counter` must be 0 on entry
// and is reset to 0 on exit.
template <int NV, int THREADS>
__device__ __forceinline__ void grid_sum(double (&v)[NV], double* partials, unsigned int* counter,
                                         double* out, double* smem) {
    block_sum<NV, THREADS>(v, smem);
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < NV; ++i) partials[(size_t)i * gridDim.x + blockIdx.x] = v[i];
        __threadfence();
        unsigned int t = atomicAdd(counter, 1u);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (is_last) {
        __threadfence();
        double s[NV];
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            s[i] = 0.0;
            for (int j = threadIdx.x; j < (int)gridDim.x; j += THREADS)
                s[i] += ((volatile double*)partials)[(size_t)i * gridDim.x + j];
        }
        __syncthreads();
        block_sum<NV, THREADS>(s, smem);
        if (threadIdx.x == 0) {
#pragma unroll
            for (int i = 0; i < NV; ++i) out[i] = s[i];
            *counter = 0u;
        }
    }
}
