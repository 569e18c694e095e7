// direct.cu -- K16: sparse direct solve of the monolithic Newton system on the GPU.
//
// The reference solves every Newton update of the mixed (velocity, pressure) system with PETSc's
// sparse LU (NonlinearVariationalSolver defaults, source/ns_bdf_solver.py:91-100;
// NewtonSolver(PETScLUSolver), source/ns_solver_base.py:938-949).  The saddle-point matrix is
// indefinite with a zero pressure block, so the Jacobi / Chebyshev Krylov solvers of krylov.cu do
// not apply; BASELINE.json's north star asks for this system to be "solved on the GPU and its
// time reported separately from the hand-written path".  This is that library leg: cuSOLVER's
// sparse QR (cusolverSpDcsrlsvqr, device path, fill-reducing reordering), bound at run time so
// that libnsb200 keeps no link-time dependency.  Seconds spent here are accumulated separately
// (nsb_direct_seconds) and never enter the kernel-class timers of the hand-written kernels.
#include <dlfcn.h>
#include <chrono>
#include "common.cuh"

namespace {
typedef void* solver_handle_t;
typedef void* mat_descr_t;

struct direct_api {
    void* hs = nullptr;   // libcusolver
    void* hp = nullptr;   // libcusparse
    int (*SpCreate)(solver_handle_t*) = nullptr;
    int (*SpDestroy)(solver_handle_t) = nullptr;
    int (*SpSetStream)(solver_handle_t, cudaStream_t) = nullptr;
    int (*SpDcsrlsvqr)(solver_handle_t, int, int, const mat_descr_t, const double*, const int*, const int*,
                       const double*, double, int, double*, int*) = nullptr;
    int (*CreateMatDescr)(mat_descr_t*) = nullptr;
    int (*DestroyMatDescr)(mat_descr_t) = nullptr;
    solver_handle_t handle = nullptr;
    mat_descr_t descr = nullptr;     // defaults: general matrix, zero-based indices
} g_direct;

void* open_any(const char* const* names) {
    for (int i = 0; names[i]; ++i) {
        void* h = dlopen(names[i], RTLD_NOW | RTLD_GLOBAL);
        if (h) return h;
    }
    return nullptr;
}

int load_direct(nsb_ctx* ctx) {
    if (g_direct.handle) return 0;
    static const char* const sparse_names[] = {"libcusparse.so.12", "/usr/local/cuda/lib64/libcusparse.so.12",
                                               "libcusparse.so", nullptr};
    static const char* const solver_names[] = {"libcusolver.so.11", "/usr/local/cuda/lib64/libcusolver.so.11",
                                               "libcusolver.so", nullptr};
    g_direct.hp = open_any(sparse_names);
    if (!g_direct.hp) return nsb_fail(ctx, "cannot load libcusparse (%s)%s", dlerror());
    g_direct.hs = open_any(solver_names);
    if (!g_direct.hs) return nsb_fail(ctx, "cannot load libcusolver (%s)%s", dlerror());
#define NSB_SYM(lib, field, name)                                                   \
    *(void**)(&g_direct.field) = dlsym(g_direct.lib, name);                          \
    if (!g_direct.field) return nsb_fail(ctx, "missing symbol %s%s", name)
    NSB_SYM(hs, SpCreate, "cusolverSpCreate");
    NSB_SYM(hs, SpDestroy, "cusolverSpDestroy");
    NSB_SYM(hs, SpSetStream, "cusolverSpSetStream");
    NSB_SYM(hs, SpDcsrlsvqr, "cusolverSpDcsrlsvqr");
    NSB_SYM(hp, CreateMatDescr, "cusparseCreateMatDescr");
    NSB_SYM(hp, DestroyMatDescr, "cusparseDestroyMatDescr");
#undef NSB_SYM
    if (g_direct.SpCreate(&g_direct.handle) != 0) return nsb_fail(ctx, "cusolverSpCreate failed%s%s");
    if (g_direct.CreateMatDescr(&g_direct.descr) != 0) return nsb_fail(ctx, "cusparseCreateMatDescr failed%s%s");
    return 0;
}
}  // namespace

// x = A^-1 b for a general CSR matrix (int32 indices, ascending columns).  reorder: 0 none,
// 1 symrcm, 2 symamd, 3 csrmetisnd (cuSOLVER's choices).  Blocking: returns after the solve.
extern "C" int nsb_direct_solve(nsb_ctx* ctx, int32_t n, int32_t nnz, const int32_t* rowptr, const int32_t* col,
                                const double* vals, const double* b, double* x, int32_t reorder,
                                double* seconds_out, void* stream) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    NSB_REQUIRE(ctx, n > 0 && nnz > 0 && rowptr && col && vals && b && x, "bad arguments");
    NSB_REQUIRE(ctx, reorder >= 0 && reorder <= 3, "reorder must be 0..3");
    if (load_direct(ctx)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    const auto t0 = std::chrono::steady_clock::now();
    if (g_direct.SpSetStream(g_direct.handle, st) != 0) return nsb_fail(ctx, "cusolverSpSetStream failed%s%s");
    int singularity = -1;
    const int rc = g_direct.SpDcsrlsvqr(g_direct.handle, n, nnz, g_direct.descr, vals, rowptr, col, b, 1.0e-14,
                                        reorder, x, &singularity);
    NSB_CUDA(ctx, cudaStreamSynchronize(st));
    const double sec = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    ctx->direct_seconds += sec;
    ctx->direct_solves += 1;
    if (seconds_out) *seconds_out = sec;
    if (rc != 0) {
        char buf[32];
        snprintf(buf, sizeof buf, "%d", rc);
        return nsb_fail(ctx, "cusolverSpDcsrlsvqr failed with status %s%s", buf);
    }
    if (singularity >= 0) {
        char buf[32];
        snprintf(buf, sizeof buf, "%d", singularity);
        return nsb_fail(ctx, "direct solve: matrix is numerically singular at row %s%s", buf);
    }
    return 0;
}

extern "C" int nsb_direct_seconds(nsb_ctx* ctx, double* seconds_out, int64_t* solves_out, int32_t reset) {
    NSB_REQUIRE(ctx, ctx != nullptr, "null context");
    if (seconds_out) *seconds_out = ctx->direct_seconds;
    if (solves_out) *solves_out = ctx->direct_solves;
    if (reset) { ctx->direct_seconds = 0.0; ctx->direct_solves = 0; }
    return 0;
}
