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
int nsb_comm_destroy(nsb_ctx* ctx) {
    if (ctx) ctx->nccl_comm = nullptr;
    return 0;
}

extern "C" int nsb_comm_finalize(nsb_ctx* ctx) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    if (ctx->nccl_comm && g_nccl.CommDestroy) {
        NSB_CUDA(ctx, cudaDeviceSynchronize());
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

int nsb_allreduce_sum(nsb_ctx* ctx, double* buf, int count, cudaStream_t st) {
    if (ctx->nranks <= 1) return 0;
    NSB_REQUIRE(ctx, ctx->nccl_comm != nullptr, "communicator not initialised");
    NSB_TIMED(ctx, NSB_T_COMM, st);
    NSB_NCCL(ctx, g_nccl.AllReduce(buf, buf, (size_t)count, NCCL_FLOAT64, NCCL_SUM, (nccl_comm_t)ctx->nccl_comm, st));
    return 0;
}
