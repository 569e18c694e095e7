// spmv_tma.cuh -- BSR-BS SpMV with the matrix streamed through shared memory by the sm_100a
// bulk-copy engine (K10, TMA 1-D form).
//
// One CTA per row tile (setup_host.cu::nsb_tiles_build).  Thread 0 arms an mbarrier with the
// tile's byte count and issues three cp.async.bulk copies: the 256-byte tile record (row
// offsets), the contiguous value chunk (entries * BS*BS * 8 B) and the contiguous column chunk
// (entries * 4 B).  No registers stage the matrix, so the bytes in flight per SM are
// resident CTAs x tile bytes (7 x 29 KB for BS = 3) -- far above the ~35 KB/SM Little's-law
// requirement of HBM3e -- independent of how short the rows are.  After the barrier flips,
// each warp owns rows w, w+NW, ... of the tile: a lane owns whole BS x BS blocks (LDS.64 at a
// 72-byte stride is bank-conflict free for BS = 3), gathers x through L1/L2, and a shuffle tree
// finishes the row.  Rows flagged in bc_mask act as identity rows (dolfin `ident`).
#pragma once
#include "common.cuh"
#include "tma.cuh"

#define NSB_TILE_WORDS 64

// VT = double: the operator itself.  VT = float: a single-precision COPY of the values, used only
// inside the Chebyshev preconditioner (krylov.cu): the polynomial is a fixed linear operator either
// way, the outer GMRES keeps the FP64 matrix, so the solution accuracy is unchanged while three of
// the four SpMVs per Arnoldi step stream 40 instead of 76 bytes per block.
template <int BS, typename VT = double>
struct SpmvTmaCfg {
    static constexpr int BB = BS * BS;
    static constexpr int TILE = (BS == 3) ? 384 : 768;     // entries per tile (== _lib.TILE_ENTRIES)
    static constexpr int THREADS = 256;
    static constexpr int CAP = TILE + 8;                   // + the 4-entry alignment hull on both sides
    static constexpr size_t VAL_BYTES = (size_t)CAP * BB * sizeof(VT);
    static constexpr size_t COL_BYTES = (size_t)CAP * 4;
    static constexpr size_t SMEM = VAL_BYTES + COL_BYTES + NSB_TILE_WORDS * 4 + 16;
};

// XT = double: x, y are ordinary dof vectors (BS doubles per node).  XT = float (with VT = float): the
// all-single-precision SpMV of the FP32 Chebyshev preconditioner -- x and y hold NSB_PAD32(BS) floats per
// node (BS = 3 padded to a float4), so a lane gathers a block column with ONE 16-byte load instead of
// three scattered 8-byte ones (the FP64-vector kernel is L1TEX-wavefront bound on exactly those),
// multiplies in FP32 and lane 0 stores the row as one vector.
#define NSB_PAD32(BS) ((BS) == 3 ? 4 : (BS))

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int BS, typename VT, typename XT>
__global__ void __launch_bounds__(SpmvTmaCfg<BS, VT>::THREADS)
k_spmv_bsr_tma(const uint32_t* __restrict__ tiles, uint32_t nnz, const int32_t* __restrict__ col,
               const VT* __restrict__ vals, const uint8_t* __restrict__ mask, const XT* __restrict__ x,
               int32_t nown, const XT* __restrict__ xg, XT* __restrict__ y) {
    using Cfg = SpmvTmaCfg<BS, VT>;
    constexpr int BB = Cfg::BB, NW = Cfg::THREADS / 32;
    constexpr bool X32 = sizeof(XT) == 4;
    constexpr int PB = X32 ? NSB_PAD32(BS) : BS;              // components stored per node in x / y
    extern __shared__ __align__(128) unsigned char smem[];
    VT* sv = reinterpret_cast<VT*>(smem);
    int32_t* sc = reinterpret_cast<int32_t*>(smem + Cfg::VAL_BYTES);
    uint32_t* rec = reinterpret_cast<uint32_t*>(smem + Cfg::VAL_BYTES + Cfg::COL_BYTES);
    uint64_t* bar = reinterpret_cast<uint64_t*>(rec + NSB_TILE_WORDS);
    const uint32_t* trec = tiles + (size_t)blockIdx.x * NSB_TILE_WORDS;
    const uint32_t row0 = __ldg(trec), nrows = __ldg(trec + 1), ent0 = __ldg(trec + 2), nent = __ldg(trec + 3);
    // bulk-copied hull [a0, a1): 4-entry (16-byte) aligned, clipped to the array; entry k of the tile
    // lands at shared index shift + k; a tail [a1, ent0 + nent) exists only at the very end of the matrix
    const uint32_t a0 = ent0 & ~3u, shift = ent0 - a0;
    uint32_t a1 = (ent0 + nent + 3u) & ~3u;
    if (a1 > nnz) a1 = nnz & ~3u;
    const uint32_t nal = a1 > a0 ? a1 - a0 : 0u;
    if (threadIdx.x == 0) {
        mbar_init(bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(bar, NSB_TILE_WORDS * 4 + nal * (BB * (uint32_t)sizeof(VT) + 4));
        tma_load_1d(rec, trec, NSB_TILE_WORDS * 4, bar);
        if (nal) {
            tma_load_1d(sv, vals + (size_t)a0 * BB, nal * BB * (uint32_t)sizeof(VT), bar);
            tma_load_1d(sc, col + a0, nal * 4, bar);
        }
    }
    if (a0 + nal < ent0 + nent) {               // <= 3 trailing entries of the matrix: plain loads
        const uint32_t n1 = ent0 + nent - a0;   // entries wanted, counted from a0
        for (uint32_t e = nal * BB + threadIdx.x; e < n1 * BB; e += Cfg::THREADS) sv[e] = vals[(size_t)a0 * BB + e];
        for (uint32_t e = nal + threadIdx.x; e < n1; e += Cfg::THREADS) sc[e] = col[a0 + e];
        __syncthreads();
    }
    // one warp polls the barrier, the others park on the CTA barrier: ncu (r1_j) showed the try_wait polls
    // of all 8 warps (~54 per warp and tile, one shared-memory wavefront each) to be 41 % of the L1
    // wavefronts of this L1-bound kernel.  try_wait acquires at CTA scope and bar.sync orders the rest.
    if (threadIdx.x < 32) mbar_wait(bar, 0);
    __syncthreads();
    const uint16_t* off = reinterpret_cast<const uint16_t*>(rec + 4);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (uint32_t lr = w; lr < nrows; lr += NW) {
        const uint32_t o0 = shift + off[lr], o1 = shift + off[lr + 1];
        XT acc[BS];
#pragma unroll
        for (int b = 0; b < BS; ++b) acc[b] = (XT)0;
        for (uint32_t k = o0 + lane; k < o1; k += 32) {
            const int32_t cn = sc[k];
            const XT* xb = (cn < nown ? x : xg) + (int64_t)cn * PB;
            XT xv[BS];
            if constexpr (X32 && BS == 3) {
                const float4 t = __ldg(reinterpret_cast<const float4*>(xb));
                xv[0] = t.x; xv[1] = t.y; xv[2] = t.z;
            } else if constexpr (X32 && BS == 2) {
                const float2 t = __ldg(reinterpret_cast<const float2*>(xb));
                xv[0] = t.x; xv[1] = t.y;
            } else {
#pragma unroll
                for (int c = 0; c < BS; ++c) xv[c] = __ldg(xb + c);
            }
            const VT* v = sv + (size_t)k * BB;
#pragma unroll
            for (int r = 0; r < BS; ++r)
#pragma unroll
                for (int c = 0; c < BS; ++c) acc[r] += (XT)v[r * BS + c] * xv[c];
        }
#pragma unroll
        for (int b = 0; b < BS; ++b) acc[b] = warp_sum(acc[b]);
        if constexpr (X32) {
            if (lane == 0) {
                const int64_t node = (int64_t)(row0 + lr);
                if (mask != nullptr) {
#pragma unroll
                    for (int b = 0; b < BS; ++b)
                        if (mask[node * BS + b]) acc[b] = x[node * PB + b];
                }
                if constexpr (BS == 3) reinterpret_cast<float4*>(y)[node] = make_float4(acc[0], acc[1], acc[2], 0.0f);
                else reinterpret_cast<float2*>(y)[node] = make_float2(acc[0], acc[1]);
            }
        } else if (lane < BS) {
            const int64_t d = (int64_t)(row0 + lr) * BS + lane;
            XT v = acc[0];
#pragma unroll
            for (int b = 1; b < BS; ++b) v = (lane == b) ? acc[b] : v;
            if (mask != nullptr && mask[d]) v = x[d];
            y[d] = v;
        }
    }
}

template <int BS, typename VT, typename XT = double>
static int launch_spmv_bsr_tma(nsb_ctx* ctx, const nsb_matrix* A, const VT* vals, const XT* x, int32_t nown,
                               const XT* xg, XT* y, cudaStream_t st) {
    using Cfg = SpmvTmaCfg<BS, VT>;
    static bool configured = false;
    if (!configured) {
        NSB_CUDA(ctx, cudaFuncSetAttribute(k_spmv_bsr_tma<BS, VT, XT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Cfg::SMEM));
        NSB_CUDA(ctx, cudaFuncSetAttribute(k_spmv_bsr_tma<BS, VT, XT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        configured = true;
    }
    k_spmv_bsr_tma<BS, VT, XT><<<A->ntiles, Cfg::THREADS, Cfg::SMEM, st>>>(A->tiles, (uint32_t)A->nnz, A->col, vals,
                                                                 A->bc_mask, x, nown, xg, y);
    NSB_LAUNCH_CHECK(ctx, "k_spmv_bsr_tma");
    return 0;
}
