// comm.cu -- multi-GPU plumbing of libnsb200 (K14): owner -> ghost halo exchange (pack kernel +
// grouped ncclSend/ncclRecv, ghosts received in place) and the all-reduce of the Krylov inner
// products.  One rank per GPU; the communicator is created from an id the host side broadcasts
// with torch.distributed.  NCCL is bound at run time (dlopen of the libnccl.so.2 that torch has
// already loaded), so the library has no link-time dependency and still loads on a CPU-only box.
#include <dlfcn.h>
#include "common.cuh"

namespace {
struct nccl_id { char internal[128]; };          // ncclUniqueId
typedef void* nccl_comm_t;
enum { NCCL_SUM = 0, NCCL_FLOAT64 = 8 };         // ncclRedOp_t / ncclDataType_t values (stable ABI)

struct nccl_api {
    void* handle = nullptr;
    int (*GetUniqueId)(nccl_id*) = nullptr;
    int (*CommInitRank)(nccl_comm_t*, int, nccl_id, int) = nullptr;
    int (*CommDestroy)(nccl_comm_t) = nullptr;
    int (*AllReduce)(const void*, void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Send)(const void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*Recv)(void*, size_t, int, int, nccl_comm_t, cudaStream_t) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
} g_nccl;

int load_nccl(nsb_ctx* ctx) {
    if (g_nccl.handle) return 0;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return nsb_fail(ctx, "cannot load libnccl.so.2 (%s): import torch first or set LD_LIBRARY_PATH%s", dlerror());
#define NSB_SYM(name)                                                                   \
    *(void**)(&g_nccl.name) = dlsym(h, "nccl" #name);                                    \
    if (!g_nccl.name) return nsb_fail(ctx, "libnccl lacks symbol nccl%s%s", #name)
    NSB_SYM(GetUniqueId); NSB_SYM(CommInitRank); NSB_SYM(CommDestroy); NSB_SYM(AllReduce);
    NSB_SYM(Send); NSB_SYM(Recv); NSB_SYM(GroupStart); NSB_SYM(GroupEnd); NSB_SYM(GetErrorString);
#undef NSB_SYM
    g_nccl.handle = h;
    return 0;
}
}  // namespace

#define NSB_NCCL(ctx, call)                                                              \
    do {                                                                                 \
        int r_ = (call);                                                                 \
        if (r_ != 0) return nsb_fail(ctx, "NCCL error %s at %s", g_nccl.GetErrorString(r_), #call); \
    } while (0)

extern "C" int nsb_comm_unique_id(uint8_t* out) {
    if (!out) return nsb_fail(nullptr, "nsb_comm_unique_id: null output%s%s");
    if (load_nccl(nullptr)) return -1;
    nccl_id id;
    NSB_NCCL(nullptr, g_nccl.GetUniqueId(&id));
    memcpy(out, id.internal, 128);
    return 0;
}

extern "C" int nsb_comm_init(nsb_ctx* ctx, int32_t rank, int32_t nranks, const uint8_t* id128) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, nranks >= 1 && rank >= 0 && rank < nranks, "bad rank / nranks");
    NSB_REQUIRE(ctx, ctx->nccl_comm == nullptr, "communicator already initialised");
    ctx->rank = rank; ctx->nranks = nranks;
    if (nranks == 1) return 0;
    NSB_REQUIRE(ctx, id128 != nullptr, "null NCCL id");
    if (load_nccl(ctx)) return -1;
    NSB_CUDA(ctx, cudaSetDevice(ctx->device));
    nccl_id id;
    memcpy(id.internal, id128, 128);
    nccl_comm_t comm = nullptr;
    NSB_NCCL(ctx, g_nccl.CommInitRank(&comm, nranks, id, rank));
    ctx->nccl_comm = comm;
    return 0;
}

// nsb_destroy path: the communicator is NOT torn down implicitly.  ncclCommDestroy synchronises with
// the peers, so calling it from object finalisers (ranks exit in any order) can block forever;
// the explicit, collective nsb_comm_finalize does it, process exit reclaims it otherwise.
int nsb_p2p_release(nsb_ctx* ctx);

int nsb_comm_destroy(nsb_ctx* ctx) {
    if (ctx) ctx->nccl_comm = nullptr;
    return 0;
}

extern "C" int nsb_comm_finalize(nsb_ctx* ctx) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_CUDA(ctx, cudaDeviceSynchronize());
    nsb_p2p_release(ctx);            // the host has put a barrier before this call: no peer still pushes here
    if (ctx->nccl_comm && g_nccl.CommDestroy) {
        NSB_NCCL(ctx, g_nccl.CommDestroy((nccl_comm_t)ctx->nccl_comm));
    }
    ctx->nccl_comm = nullptr;
    ctx->nranks = 1; ctx->rank = 0;
    return 0;
}

extern "C" int nsb_set_halo(nsb_ctx* ctx, int32_t which, int32_t n_owned, int32_t n_ghost, int64_t n_global,
                            int32_t npeers, const int32_t* peers, const int32_t* send_ptr,
                            const int32_t* send_idx_dev, const int32_t* recv_ptr, double* sendbuf_dev,
                            double* ghostbuf_dev) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which == NSB_HALO_NODES || which == NSB_HALO_VERTS, "bad halo index space");
    NSB_REQUIRE(ctx, n_owned > 0 && n_ghost >= 0 && n_global >= n_owned && npeers >= 0, "bad halo sizes");
    NSB_REQUIRE(ctx, npeers == 0 || (peers && send_ptr && recv_ptr), "null peer tables");
    nsb_halo& H = ctx->halo[which];
    H.n_owned = n_owned; H.n_ghost = n_ghost; H.n_global = n_global;
    H.peers.assign(peers, peers + npeers);
    H.send_ptr.assign(1, 0); H.recv_ptr.assign(1, 0);
    if (npeers) {
        H.send_ptr.assign(send_ptr, send_ptr + npeers + 1);
        H.recv_ptr.assign(recv_ptr, recv_ptr + npeers + 1);
    }
    for (int i = 0; i < npeers; ++i) {
        NSB_REQUIRE(ctx, peers[i] >= 0 && peers[i] < ctx->nranks && peers[i] != ctx->rank, "bad peer rank");
        NSB_REQUIRE(ctx, H.send_ptr[i + 1] >= H.send_ptr[i] && H.recv_ptr[i + 1] >= H.recv_ptr[i], "peer ranges not monotone");
    }
    NSB_REQUIRE(ctx, H.send_ptr[0] == 0 && H.recv_ptr[0] == 0 && H.recv_ptr.back() == n_ghost, "ghost ranges do not cover the ghost block");
    NSB_REQUIRE(ctx, H.send_ptr.back() == 0 || (send_idx_dev && sendbuf_dev), "null send tables");
    NSB_REQUIRE(ctx, n_ghost == 0 || ghostbuf_dev, "null ghost buffer");
    H.send_idx = send_idx_dev; H.sendbuf = sendbuf_dev; H.ghostbuf = ghostbuf_dev;
    return 0;
}

// buf[bs*i + c] = x[bs*idx[i] + c]
__global__ void __launch_bounds__(256)
k_halo_pack(int64_t n, int bs, const int32_t* __restrict__ idx, const double* __restrict__ x, double* __restrict__ buf) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n * bs) return;
    const int64_t i = e / bs;
    buf[e] = x[(int64_t)idx[i] * bs + (e - i * bs)];
}

static int p2p_exchange(nsb_ctx* ctx, nsb_halo& H, int bs, const double* x, double* ghost_dst, cudaStream_t st);
static int p2p_allreduce(nsb_ctx* ctx, double* buf, int count, cudaStream_t st);

extern "C" int nsb_halo_exchange(nsb_ctx* ctx, int32_t which, int32_t bs, const double* x, double* ghost_dst,
                                 void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which == NSB_HALO_NODES || which == NSB_HALO_VERTS, "bad halo index space");
    NSB_REQUIRE(ctx, bs >= 1 && bs <= 3 && x, "bad arguments");
    if (ctx->nranks <= 1) return 0;
    nsb_halo& H = ctx->halo[which];
    NSB_REQUIRE(ctx, H.n_owned > 0, "halo plan not set");
    NSB_REQUIRE(ctx, ctx->nccl_comm != nullptr, "communicator not initialised");
    cudaStream_t st = (cudaStream_t)stream;
    if (!ghost_dst) ghost_dst = H.ghostbuf;
    if (H.p2p_ready) return p2p_exchange(ctx, H, bs, x, ghost_dst, st);      // NVLink peer-memory path
    const int npeers = (int)H.peers.size();
    const int64_t nsend = H.send_ptr.back();
    NSB_TIMED(ctx, NSB_T_COMM, st);
    if (nsend) {
        k_halo_pack<<<nsb_blocks(nsend * bs, 256), 256, 0, st>>>(nsend, bs, H.send_idx, x, H.sendbuf);
        NSB_LAUNCH_CHECK(ctx, "k_halo_pack");
    }
    nccl_comm_t comm = (nccl_comm_t)ctx->nccl_comm;
    NSB_NCCL(ctx, g_nccl.GroupStart());
    for (int i = 0; i < npeers; ++i) {
        const int64_t ns = H.send_ptr[i + 1] - H.send_ptr[i], nr = H.recv_ptr[i + 1] - H.recv_ptr[i];
        if (ns) NSB_NCCL(ctx, g_nccl.Send(H.sendbuf + (int64_t)H.send_ptr[i] * bs, (size_t)(ns * bs), NCCL_FLOAT64, H.peers[i], comm, st));
        if (nr) NSB_NCCL(ctx, g_nccl.Recv(ghost_dst + (int64_t)H.recv_ptr[i] * bs, (size_t)(nr * bs), NCCL_FLOAT64, H.peers[i], comm, st));
    }
    NSB_NCCL(ctx, g_nccl.GroupEnd());
    return 0;
}

extern "C" int nsb_allreduce(nsb_ctx* ctx, double* buf_dev, int64_t count, void* stream) {
    NSB_REQUIRE(ctx, ctx && buf_dev && count >= 0 && count < (int64_t)1 << 31, "bad arguments");
    return nsb_allreduce_sum(ctx, buf_dev, (int)count, (cudaStream_t)stream);
}

int nsb_allreduce_sum(nsb_ctx* ctx, double* buf, int count, cudaStream_t st) {
    if (ctx->nranks <= 1) return 0;
    if (ctx->ar_ready && count <= NSB_AR_MAX) return p2p_allreduce(ctx, buf, count, st);
    NSB_REQUIRE(ctx, ctx->nccl_comm != nullptr, "communicator not initialised");
    NSB_TIMED(ctx, NSB_T_COMM, st);
    NSB_NCCL(ctx, g_nccl.AllReduce(buf, buf, (size_t)count, NCCL_FLOAT64, NCCL_SUM, (nccl_comm_t)ctx->nccl_comm, st));
    return 0;
}

// ---------------------------------------------------------------------------------------------
// Halo exchange over NVLink peer memory (no NCCL send / recv).  Every rank owns an IPC-shared block
// per index space: acknowledgement words and a ring of NSB_P2P_RING receive slots shaped like its
// ghost block.  ONE kernel per exchange on the caller's stream, no fences, no atomics:
//   push blocks  (one group per neighbour): gather the boundary values out of x and store each
//                double into the neighbour's ring slot as a 16-byte line {lo, epoch, hi, epoch}
//                (flag-in-data, 8-byte store atomicity is all it needs);
//   recv blocks  (one group per neighbour): poll the LOCAL slot lines until both flags carry this
//                exchange's epoch, write the doubles into the ghost block of the destination vector,
//                then store the epoch into the neighbour's acknowledgement word of that block.
// All remote traffic is stores, all polling is on local memory.  Push block b and recv block b of a
// pair cover the same entries (the grid of a pair depends only on its entry count, which both ends
// know), so flow control is per block: a slot is overwritten only after the neighbour acknowledged
// epoch - NSB_P2P_RING.  Every rank issues its exchanges in the same order (SPMD), so the per-plan
// epoch counters agree without a handshake.  Waits are bounded (about 30 s): a lost peer raises the
// context's error word (nsb_p2p_status) instead of hanging the GPU.
// ---------------------------------------------------------------------------------------------
namespace {
constexpr int P2P_MAXB = 128;            // blocks per neighbour pair
constexpr int P2P_PER_BLOCK = 512;      // entries (doubles) per block before the grid-stride loop kicks in

__host__ __device__ inline int p2p_pair_blocks(long long entries) {
    long long b = (entries + P2P_PER_BLOCK - 1) / P2P_PER_BLOCK;
    return (int)(b < 1 ? 1 : (b > P2P_MAXB ? P2P_MAXB : b));
}
__device__ __forceinline__ unsigned long long ld_volatile_u64(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_volatile_u64(unsigned long long* p, unsigned long long v) {
    asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

struct p2p_args {
    int n;                               // neighbours
    unsigned long long* peer_ack[8];     // their acknowledgement words [8][P2P_MAXB]
    uint4* peer_ring[8];                 // their rings (16-byte lines)
    long long peer_cap[8];               // lines per ring slot over there
    int peer_off[8];                     // entries: where my part starts inside their ghost block
    int my_slot[8];                      // my index in their neighbour list
    int send_ptr[9], recv_ptr[9];        // entries, like the NCCL plan
};

// grid (blocks per pair, 2n): y < n pushes to neighbour y, y >= n receives from neighbour y - n.
__global__ void __launch_bounds__(256)
k_p2p_exchange(int bs, unsigned long long epoch, long long cap, p2p_args P,
               const int32_t* __restrict__ send_idx, const double* __restrict__ x, unsigned long long* ack,
               const uint4* ring, double* __restrict__ ghost_dst, int* err) {
    const int y = blockIdx.y;
    const int slot = (int)(epoch % NSB_P2P_RING);
    const unsigned int flag = (unsigned int)epoch;
    if (y < P.n) {                                            // ---- push my boundary values to neighbour y
        const int j = y;
        const int64_t n = (int64_t)(P.send_ptr[j + 1] - P.send_ptr[j]) * bs;
        const int nb = p2p_pair_blocks(n);
        if (n == 0 || (int)blockIdx.x >= nb) return;
        if (epoch > NSB_P2P_RING) {                           // the lines I am about to overwrite were consumed?
            if (threadIdx.x == 0 && *(volatile int*)err == 0) {
                const long long t0 = clock64();
                while (ld_volatile_u64(ack + j * P2P_MAXB + blockIdx.x) + NSB_P2P_RING < epoch) {
                    if (clock64() - t0 > NSB_P2P_SPIN_CYCLES) { *err = 1; break; }
                    __nanosleep(32);
                }
            }
            __syncthreads();
        }
        uint4* dst = P.peer_ring[j] + (int64_t)slot * P.peer_cap[j] + (int64_t)P.peer_off[j] * bs;
        const int32_t* idx = send_idx + P.send_ptr[j];
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)nb * blockDim.x) {
            const int64_t i = e / bs;
            nsb_st_line(dst + e, x[(int64_t)idx[i] * bs + (e - i * bs)], flag);
        }
    } else {                                                  // ---- take what neighbour i pushed
        const int i = y - P.n;
        const int64_t n = (int64_t)(P.recv_ptr[i + 1] - P.recv_ptr[i]) * bs;
        const int nb = p2p_pair_blocks(n);
        if (n == 0 || (int)blockIdx.x >= nb) return;
        const uint4* src = ring + (int64_t)slot * cap + (int64_t)P.recv_ptr[i] * bs;
        double* dst = ghost_dst + (int64_t)P.recv_ptr[i] * bs;
        for (int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (int64_t)nb * blockDim.x)
            dst[e] = nsb_wait_line(src + e, flag, err, 2);
        __syncthreads();                                      // every line of this block has been read
        if (threadIdx.x == 0) st_volatile_u64(P.peer_ack[i] + P.my_slot[i] * P2P_MAXB + blockIdx.x, epoch);
        // exchanges of this plan differ in width (3 doubles per node, 2 for a padded float4), hence in blocks
        // per pair: the blocks this exchange does not use have nothing to consume -- acknowledge for them too
        if (blockIdx.x == 0 && (int)threadIdx.x >= nb && threadIdx.x < P2P_MAXB)
            st_volatile_u64(P.peer_ack[i] + P.my_slot[i] * P2P_MAXB + threadIdx.x, epoch);
    }
}

inline long long p2p_cap_of(long long n_ghost) { return (n_ghost * 3 + 15) & ~15ll; }

// All-reduce (sum) of a few doubles over peer memory: every rank stores its values as flagged lines
// into every rank's slot [epoch % ring][sender][k], then sums the nranks lines of its own slot in rank
// order -- the same order everywhere, so all ranks hold the bit-identical sum.  One block, no fences.
// No acknowledgements: the exchange is all-to-all, so having consumed every rank's epoch e - 1 proves
// that their kernels of epoch e - 2 finished (same stream); a ring of >= 2 slots cannot be overrun.
struct ar_args {
    int nranks, rank;
    uint4* ring[8];
};
__global__ void __launch_bounds__(256)
k_p2p_allreduce(double* __restrict__ buf, int count, unsigned long long epoch, ar_args A, int* err) {
    const int slot = (int)(epoch % NSB_AR_RING);
    const unsigned int flag = (unsigned int)epoch;
    for (int t = threadIdx.x; t < count * A.nranks; t += blockDim.x) {
        const int r = t / count, k = t - r * count;
        nsb_st_line(A.ring[r] + ((int64_t)slot * 8 + A.rank) * NSB_AR_MAX + k, buf[k], flag);
    }
    __syncthreads();
    if ((int)threadIdx.x < count) {
        const uint4* mine = A.ring[A.rank] + (int64_t)slot * 8 * NSB_AR_MAX + threadIdx.x;
        double sum = 0.0;
        for (int r = 0; r < A.nranks; ++r) {
            const double d = nsb_wait_line(mine + r * NSB_AR_MAX, flag, err, 3);
            sum = r == 0 ? d : sum + d;
        }
        buf[threadIdx.x] = sum;
    }
}
}  // namespace

// Allocate this rank's shared block for one index space and hand out its IPC handle (64 bytes).
extern "C" int nsb_p2p_export(nsb_ctx* ctx, int32_t which, uint8_t* handle_out64) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which >= NSB_HALO_NONE && which <= NSB_HALO_VERTS && handle_out64, "bad arguments");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    if (!ctx->d_p2p_err) {
        NSB_CUDA(ctx, cudaMalloc(&ctx->d_p2p_err, sizeof(int)));
        NSB_CUDA(ctx, cudaMemset(ctx->d_p2p_err, 0, sizeof(int)));
    }
    if (which == NSB_HALO_NONE) {                              // the all-reduce block
        NSB_REQUIRE(ctx, ctx->nranks >= 2 && ctx->nranks <= 8, "peer-memory all-reduce supports 2..8 ranks");
        if (!ctx->ar_block) {
            const size_t bytes = sizeof(uint4) * NSB_AR_RING * 8 * NSB_AR_MAX;
            NSB_CUDA(ctx, cudaMalloc(&ctx->ar_block, bytes));
            NSB_CUDA(ctx, cudaMemset(ctx->ar_block, 0, bytes));
            NSB_CUDA(ctx, cudaDeviceSynchronize());
        }
        cudaIpcMemHandle_t h;
        NSB_CUDA(ctx, cudaIpcGetMemHandle(&h, ctx->ar_block));
        memcpy(handle_out64, &h, 64);
        return 0;
    }
    nsb_halo& H = ctx->halo[which];
    NSB_REQUIRE(ctx, H.n_owned > 0, "halo plan not set");
    NSB_REQUIRE(ctx, H.peers.size() <= 8, "peer-memory halo supports at most 8 neighbours");
    if (!H.p2p_block) {
        H.p2p_cap = p2p_cap_of(H.n_ghost);
        static_assert(NSB_P2P_HDR_BYTES >= 8 * P2P_MAXB * sizeof(unsigned long long), "header too small");
        const size_t bytes = NSB_P2P_HDR_BYTES + sizeof(uint4) * (size_t)H.p2p_cap * NSB_P2P_RING;
        NSB_CUDA(ctx, cudaMalloc(&H.p2p_block, bytes));
        NSB_CUDA(ctx, cudaMemset(H.p2p_block, 0, bytes));      // epoch 0 = nothing pushed, nothing acknowledged
        NSB_CUDA(ctx, cudaDeviceSynchronize());
    }
    cudaIpcMemHandle_t h;
    NSB_CUDA(ctx, cudaIpcGetMemHandle(&h, H.p2p_block));
    memcpy(handle_out64, &h, 64);
    return 0;
}

// Map the neighbours' blocks.  Per neighbour i (plan order): handles64[64*i..] its handle,
// peer_n_ghost[i] its ghost count (shapes its ring), peer_off[i] where this rank's part starts inside
// its ghost block (its recv_ptr for this rank), my_slot[i] this rank's index in its neighbour list.
// Every rank must have exported before any rank connects (the host puts a barrier in between).
extern "C" int nsb_p2p_connect(nsb_ctx* ctx, int32_t which, int32_t npeers, const uint8_t* handles64,
                               const int32_t* peer_n_ghost, const int32_t* peer_off, const int32_t* my_slot) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, which >= NSB_HALO_NONE && which <= NSB_HALO_VERTS, "bad halo index space");
    if (which == NSB_HALO_NONE) {                              // all-reduce: handles of ALL ranks, own entry ignored
        NSB_REQUIRE(ctx, ctx->ar_block != nullptr, "call nsb_p2p_export first");
        NSB_REQUIRE(ctx, npeers == ctx->nranks && handles64, "all-reduce block: one handle per rank");
        ctx->ar_peer.assign(npeers, nullptr);
        for (int r = 0; r < npeers; ++r) {
            if (r == ctx->rank) { ctx->ar_peer[r] = ctx->ar_block; continue; }
            cudaIpcMemHandle_t h;
            memcpy(&h, handles64 + 64 * (size_t)r, 64);
            void* p = nullptr;
            cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
            if (e != cudaSuccess) {
                cudaGetLastError();
                return nsb_fail(ctx, "cudaIpcOpenMemHandle failed (%s): peer-memory all-reduce unavailable", cudaGetErrorString(e));
            }
            ctx->ar_peer[r] = p;
        }
        ctx->ar_epoch = 0;
        ctx->ar_ready = true;
        return 0;
    }
    nsb_halo& H = ctx->halo[which];
    NSB_REQUIRE(ctx, H.p2p_block != nullptr, "call nsb_p2p_export first");
    NSB_REQUIRE(ctx, npeers == (int32_t)H.peers.size() && (npeers == 0 || (handles64 && peer_n_ghost && peer_off && my_slot)),
                "bad peer tables");
    H.peer_block.assign(npeers, nullptr);
    H.peer_cap.assign(npeers, 0);
    H.peer_off.assign(peer_off, peer_off + npeers);
    H.my_slot.assign(my_slot, my_slot + npeers);
    for (int i = 0; i < npeers; ++i) {
        NSB_REQUIRE(ctx, my_slot[i] >= 0 && my_slot[i] < 8 && peer_off[i] >= 0 && peer_n_ghost[i] >= 0, "bad peer slot / offset");
        NSB_REQUIRE(ctx, (int64_t)peer_off[i] + (H.send_ptr[i + 1] - H.send_ptr[i]) <= peer_n_ghost[i], "send range leaves the peer's ghost block");
        H.peer_cap[i] = p2p_cap_of(peer_n_ghost[i]);
        cudaIpcMemHandle_t h;
        memcpy(&h, handles64 + 64 * (size_t)i, 64);
        void* p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            return nsb_fail(ctx, "cudaIpcOpenMemHandle failed (%s): peer-memory halo unavailable", cudaGetErrorString(e));
        }
        H.peer_block[i] = p;
    }
    H.p2p_epoch = 0;
    H.p2p_ready = true;
    return 0;
}

extern "C" int nsb_p2p_disconnect(nsb_ctx* ctx) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_CUDA(ctx, cudaDeviceSynchronize());
    return nsb_p2p_release(ctx);
}

extern "C" int nsb_p2p_status(nsb_ctx* ctx, int32_t* error_out) {
    NSB_REQUIRE(ctx, ctx != nullptr && error_out != nullptr, "bad arguments");
    *error_out = 0;
    if (ctx->d_p2p_err) NSB_CUDA(ctx, cudaMemcpy(error_out, ctx->d_p2p_err, sizeof(int), cudaMemcpyDeviceToHost));
    return 0;
}

// unmap the neighbours and free the shared block (after a host barrier: nobody may still push here)
int nsb_p2p_release(nsb_ctx* ctx) {
    for (int w = 0; w < 3; ++w) {
        nsb_halo& H = ctx->halo[w];
        for (void* p : H.peer_block) if (p) cudaIpcCloseMemHandle(p);
        H.peer_block.clear();
        H.p2p_ready = false;
        if (H.p2p_block) cudaFree(H.p2p_block);
        H.p2p_block = nullptr;
    }
    for (size_t r = 0; r < ctx->ar_peer.size(); ++r)
        if (ctx->ar_peer[r] && (int)r != ctx->rank) cudaIpcCloseMemHandle(ctx->ar_peer[r]);
    ctx->ar_peer.clear();
    ctx->ar_ready = false;
    if (ctx->ar_block) cudaFree(ctx->ar_block);
    ctx->ar_block = nullptr;
    if (ctx->d_p2p_err) cudaFree(ctx->d_p2p_err);
    ctx->d_p2p_err = nullptr;
    cudaGetLastError();
    return 0;
}

static int p2p_exchange(nsb_ctx* ctx, nsb_halo& H, int bs, const double* x, double* ghost_dst, cudaStream_t st) {
    const int npeers = (int)H.peers.size();
    if (npeers == 0) return 0;
    const uint64_t epoch = ++H.p2p_epoch;
    p2p_args P;
    memset(&P, 0, sizeof(P));
    P.n = npeers;
    int64_t most = 0;
    for (int i = 0; i < npeers; ++i) {
        P.peer_ack[i] = (unsigned long long*)H.peer_block[i];
        P.peer_ring[i] = (uint4*)((char*)H.peer_block[i] + NSB_P2P_HDR_BYTES);
        P.peer_cap[i] = H.peer_cap[i];
        P.peer_off[i] = H.peer_off[i];
        P.my_slot[i] = H.my_slot[i];
        most = std::max<int64_t>(most, std::max(H.send_ptr[i + 1] - H.send_ptr[i], H.recv_ptr[i + 1] - H.recv_ptr[i]));
    }
    for (int i = 0; i <= npeers; ++i) { P.send_ptr[i] = H.send_ptr[i]; P.recv_ptr[i] = H.recv_ptr[i]; }
    if (most == 0) return 0;
    NSB_TIMED(ctx, NSB_T_COMM, st);
    k_p2p_exchange<<<dim3(p2p_pair_blocks(most * bs), 2 * npeers), 256, 0, st>>>(
        bs, epoch, H.p2p_cap, P, H.send_idx, x, (unsigned long long*)H.p2p_block,
        (const uint4*)((const char*)H.p2p_block + NSB_P2P_HDR_BYTES), ghost_dst, ctx->d_p2p_err);
    NSB_LAUNCH_CHECK(ctx, "k_p2p_exchange");
    return 0;
}

static int p2p_allreduce(nsb_ctx* ctx, double* buf, int count, cudaStream_t st) {
    if (count <= 0) return 0;
    ar_args A;
    memset(&A, 0, sizeof(A));
    A.nranks = ctx->nranks; A.rank = ctx->rank;
    for (int r = 0; r < ctx->nranks; ++r) A.ring[r] = (uint4*)ctx->ar_peer[r];
    const uint64_t epoch = ++ctx->ar_epoch;
    NSB_TIMED(ctx, NSB_T_COMM, st);
    k_p2p_allreduce<<<1, 256, 0, st>>>(buf, count, epoch, A, ctx->d_p2p_err);
    NSB_LAUNCH_CHECK(ctx, "k_p2p_allreduce");
    return 0;
}
