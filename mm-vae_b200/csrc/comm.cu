// K7 — data-parallel gradient exchange: ncclAllReduce(sum) of the flat gradient
// buffer over NVLink 5 / NVSwitch, one process per GPU (the
// reference has no counterpart: grep nccl|mpi -> none).
//
// libnccl is resolved at run time with dlopen so that the library loads (and
// every single-GPU path works) without it; inside a torchrun process the
// already-loaded torch-bundled libnccl.so.2 is picked up by soname.
#include "common.cuh"
#include "internal.hh"

#include <dlfcn.h>

namespace {

// Minimal NCCL ABI (nccl.h 2.x): opaque comm, 128-byte unique id, enums.
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclFloat32 = 7, ncclFloat64 = 8 };
enum { ncclSum = 0 };

struct nccl_api_t {
    void *lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
};

nccl_api_t g_nccl;

int
load_nccl()
{
    if (g_nccl.lib) return 0;
    const char *names[] = { "libnccl.so.2", "libnccl.so" };
    void *lib = nullptr;
    for (const char *n : names) {
        lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
        if (lib) break;
    }
    MMV_CHECK(lib, "comm: cannot dlopen libnccl.so.2 (%s)", dlerror());
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank))dlsym(lib, "ncclCommInitRank");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy))dlsym(lib, "ncclCommDestroy");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce))dlsym(lib, "ncclAllReduce");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString))dlsym(lib, "ncclGetErrorString");
    MMV_CHECK(g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy && g_nccl.AllReduce,
              "comm: libnccl lacks a required symbol");
    g_nccl.lib = lib;
    return 0;
}

#define MMV_NCCL(expr)                                                                      \
    do {                                                                                    \
        int _r = (expr);                                                                    \
        if (_r != ncclSuccess) {                                                            \
            ::mmv::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,                  \
                             g_nccl.GetErrorString ? g_nccl.GetErrorString(_r) : "nccl error"); \
            return 1;                                                                       \
        }                                                                                   \
    } while (0)

} // namespace

namespace mmv {

int
comm_allreduce_f32(mmv_ctx *ctx, float *buf, size_t n)
{
    if (ctx->nranks <= 1 || n == 0) return 0;
    MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat32, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->stream));
    return 0;
}

int
comm_allreduce_f32_async(mmv_ctx *ctx, float *buf, size_t n)
{
    if (ctx->nranks <= 1 || n == 0) return 0;
    MMV_CUDA(cudaEventRecord(ctx->ev_fork, ctx->stream));
    MMV_CUDA(cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_fork, 0));
    MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat32, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->comm_stream));
    MMV_CUDA(cudaEventRecord(ctx->ev_join, ctx->comm_stream));
    return 0;
}

int
comm_join(mmv_ctx *ctx)
{
    if (ctx->nranks <= 1) return 0;
    MMV_CUDA(cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
    return 0;
}

int
comm_allreduce_f64(mmv_ctx *ctx, double *buf, size_t n)
{
    if (ctx->nranks <= 1 || n == 0) return 0;
    MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat64, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->stream));
    return 0;
}

} // namespace mmv

extern "C" {

int
mmv_comm_unique_id(unsigned char id[128])
{
    MMV_CHECK(id, "mmv_comm_unique_id: NULL");
    MMV_TRY(load_nccl());
    ncclUniqueId u;
    MMV_NCCL(g_nccl.GetUniqueId(&u));
    memcpy(id, u.internal, 128);
    return 0;
}

int
mmv_comm_init(mmv_ctx *ctx, int nranks, int rank, const unsigned char id[128])
{
    MMV_CHECK(ctx && id, "mmv_comm_init: NULL");
    MMV_CHECK(nranks >= 1 && rank >= 0 && rank < nranks, "mmv_comm_init: rank %d / %d", rank, nranks);
    MMV_CHECK(ctx->nccl_comm == nullptr, "mmv_comm_init: already initialised");
    if (nranks == 1) return 0;
    MMV_TRY(load_nccl());
    MMV_CUDA(cudaSetDevice(ctx->device));
    ncclUniqueId u;
    memcpy(u.internal, id, 128);
    ncclComm_t comm = nullptr;
    MMV_NCCL(g_nccl.CommInitRank(&comm, nranks, u, rank));
    ctx->nccl_comm = comm;
    MMV_CUDA(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
    MMV_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    MMV_CUDA(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    MMV_CUDA(cudaEventRecord(ctx->ev_join, ctx->comm_stream));
    ctx->nranks = nranks;
    ctx->rank = rank;
    return 0;
}

int
mmv_comm_destroy(mmv_ctx *ctx)
{
    if (!ctx || !ctx->nccl_comm) return 0;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cudaStreamSynchronize(ctx->comm_stream);
    g_nccl.CommDestroy((ncclComm_t)ctx->nccl_comm);
    cudaStreamDestroy(ctx->comm_stream);
    cudaEventDestroy(ctx->ev_fork);
    cudaEventDestroy(ctx->ev_join);
    ctx->comm_stream = nullptr;
    ctx->nccl_comm = nullptr;
    ctx->nranks = 1;
    ctx->rank = 0;
    return 0;
}

} // extern "C"
