// K7 — data-parallel gradient exchange: ncclAllReduce(sum) of the flat gradient
// buffer over NVLink 5 / NVSwitch, one process per GPU (the
// reference has no counterpart: grep nccl|mpi -> none).
//
// libnccl is resolved at run time with dlopen so that the library loads (and
// every single-GPU path works) without it; inside a torchrun process the
// already-loaded torch-bundled libnccl.so.2 is picked up by soname.
#include "common.cuh"
#include "internal.hh"

#include <cuda_bf16.h>
#include <dlfcn.h>
#include <mutex>

namespace {

// Minimal NCCL ABI (nccl.h 2.x): opaque comm, 128-byte unique id, enums.
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef struct ncclWindow *ncclWindow_t;
enum { ncclSuccess = 0 };
enum { ncclFloat32 = 7, ncclFloat64 = 8, ncclBfloat16 = 9 };
enum { ncclSum = 0 };

struct nccl_api_t {
    void *lib = nullptr;
    int (*GetUniqueId)(ncclUniqueId *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    // optional (NCCL >= 2.27): library-allocated buffers registered as symmetric windows
    int (*MemAlloc)(void **, size_t) = nullptr;
    int (*MemFree)(void *) = nullptr;
    int (*CommWindowRegister)(ncclComm_t, void *, size_t, ncclWindow_t *, int) = nullptr;
    int (*CommWindowDeregister)(ncclComm_t, ncclWindow_t) = nullptr;
};

nccl_api_t g_nccl;

std::mutex g_nccl_mu;   // the ranks of a one-process / multi-thread host initialise concurrently

int
load_nccl()
{
    std::lock_guard<std::mutex> lk(g_nccl_mu);
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
    g_nccl.GroupStart = (decltype(g_nccl.GroupStart))dlsym(lib, "ncclGroupStart");
    g_nccl.GroupEnd = (decltype(g_nccl.GroupEnd))dlsym(lib, "ncclGroupEnd");
    g_nccl.MemAlloc = (decltype(g_nccl.MemAlloc))dlsym(lib, "ncclMemAlloc");
    g_nccl.MemFree = (decltype(g_nccl.MemFree))dlsym(lib, "ncclMemFree");
    g_nccl.CommWindowRegister = (decltype(g_nccl.CommWindowRegister))dlsym(lib, "ncclCommWindowRegister");
    g_nccl.CommWindowDeregister = (decltype(g_nccl.CommWindowDeregister))dlsym(lib, "ncclCommWindowDeregister");
    MMV_CHECK(g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy && g_nccl.AllReduce && g_nccl.GroupStart &&
                  g_nccl.GroupEnd,
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

// fp32 <-> bf16 staging of a bucket that crosses the wire in bf16 (8 elements per thread and trip; n % 8 == 0)
__global__ void
pack_bf16_kernel(const float *__restrict__ x, uint4 *__restrict__ y, size_t n8)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const float4 a = reinterpret_cast<const float4 *>(x)[2 * i], b = reinterpret_cast<const float4 *>(x)[2 * i + 1];
        const __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
        const __nv_bfloat162 p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
        uint4 o;
        o.x = *reinterpret_cast<const uint32_t *>(&p0); o.y = *reinterpret_cast<const uint32_t *>(&p1);
        o.z = *reinterpret_cast<const uint32_t *>(&p2); o.w = *reinterpret_cast<const uint32_t *>(&p3);
        y[i] = o;
    }
}

__global__ void
unpack_bf16_kernel(const uint4 *__restrict__ y, float *__restrict__ x, size_t n8)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = y[i];
        reinterpret_cast<float4 *>(x)[2 * i] = make_float4(__uint_as_float(v.x << 16), __uint_as_float(v.x & 0xFFFF0000u),
                                                          __uint_as_float(v.y << 16), __uint_as_float(v.y & 0xFFFF0000u));
        reinterpret_cast<float4 *>(x)[2 * i + 1] = make_float4(__uint_as_float(v.z << 16), __uint_as_float(v.z & 0xFFFF0000u),
                                                              __uint_as_float(v.w << 16), __uint_as_float(v.w & 0xFFFF0000u));
    }
}

// the same, and sum(x^2) of what it writes: one partial per block, summed later in block order (data-parallel replicas
// must compute the identical clip coefficient, so no arrival-order atomics)
__global__ void __launch_bounds__(256)
unpack_bf16_ss_kernel(const uint4 *__restrict__ y, float *__restrict__ x, size_t n8, float *__restrict__ partial)
{
    float s = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
        const uint4 v = y[i];
        const float f[8] = { __uint_as_float(v.x << 16), __uint_as_float(v.x & 0xFFFF0000u), __uint_as_float(v.y << 16),
                             __uint_as_float(v.y & 0xFFFF0000u), __uint_as_float(v.z << 16), __uint_as_float(v.z & 0xFFFF0000u),
                             __uint_as_float(v.w << 16), __uint_as_float(v.w & 0xFFFF0000u) };
        reinterpret_cast<float4 *>(x)[2 * i] = make_float4(f[0], f[1], f[2], f[3]);
        reinterpret_cast<float4 *>(x)[2 * i + 1] = make_float4(f[4], f[5], f[6], f[7]);
#pragma unroll
        for (int q = 0; q < 8; ++q) s = fmaf(f[q], f[q], s);
    }
    __shared__ float scratch[32];
    float v1[1] = { s };
    mmv::block_sum<1>(v1, scratch);
    if (threadIdx.x == 0) partial[blockIdx.x] = v1[0];
}

} // namespace

namespace mmv {

int
comm_allreduce_f32(mmv_ctx *ctx, float *buf, size_t n)
{
    if (ctx->nranks <= 1 || n == 0) return 0;
    MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat32, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->stream));
    return 0;
}

void
trace_begin(mmv_ctx *ctx)
{
    if (ctx->trace_step < 0) {
        const char *e = getenv("MMV_COMM_TRACE");
        ctx->trace_step = e ? atoi(e) : 0;
    }
    ctx->trace_on = ctx->trace_step > 0 && ++ctx->trace_seen == ctx->trace_step;
    if (ctx->trace_on) trace_mark(ctx, "step_begin");
}

void
trace_mark(mmv_ctx *ctx, const char *name, bool on_comm)
{
    if (!ctx->trace_on) return;
    mmv_ctx::mark_t k;
    k.name = name; k.comm = on_comm ? 1 : 0;
    if (cudaEventCreate(&k.ev) != cudaSuccess) return;
    cudaEventRecord(k.ev, on_comm ? ctx->comm_stream : ctx->stream);
    ctx->marks.push_back(k);
}

void
trace_end(mmv_ctx *ctx)
{
    if (!ctx->trace_on) return;
    trace_mark(ctx, "step_end");
    cudaStreamSynchronize(ctx->stream);
    if (ctx->comm_stream) cudaStreamSynchronize(ctx->comm_stream);
    for (auto &k : ctx->marks) {
        float ms = 0.f;
        cudaEventElapsedTime(&ms, ctx->marks[0].ev, k.ev);
        if (ctx->rank == 0) fprintf(stderr, "[trace] %-8s %8.3f ms  %s\n", k.comm ? "comm" : "compute", ms, k.name.c_str());
    }
    for (auto &k : ctx->marks) cudaEventDestroy(k.ev);
    ctx->marks.clear();
    ctx->trace_on = false;
}

// The flat gradient buffer of a data-parallel model: allocated by NCCL and registered on the communicator as a
// symmetric window (every rank registers the same size in the same order — model creation is then collective),
// so the all-reduces run NCCL's symmetric-memory kernels straight on it.  OPT-IN with MMV_COMM_SYMMETRIC=1:
// measured on 2 and 8 B200s the kernels that WRITE the gradients run slower on that memory than the
// reductions gain (3.64 vs 3.13 ms/step at N=2, 3.52 vs 3.41 at N=8); plain cudaMalloc otherwise.
int
comm_alloc_grads(mmv_ctx *ctx, size_t bytes, float **ptr, void **win, bool *nccl_owned)
{
    *win = nullptr; *nccl_owned = false;
    const char *e = getenv("MMV_COMM_SYMMETRIC");
    const bool want = ctx->nranks > 1 && ctx->nccl_comm && e && atoi(e) != 0 && g_nccl.MemAlloc && g_nccl.MemFree &&
                      g_nccl.CommWindowRegister && g_nccl.CommWindowDeregister;
    if (want) {
        const size_t gran = (size_t)2 << 20, rounded = (bytes + gran - 1) / gran * gran;
        void *p = nullptr;
        if (g_nccl.MemAlloc(&p, rounded) == ncclSuccess && p) {
            ncclWindow_t w = nullptr;
            if (g_nccl.CommWindowRegister((ncclComm_t)ctx->nccl_comm, p, rounded, &w, /*NCCL_WIN_COLL_SYMMETRIC*/ 1) != ncclSuccess)
                w = nullptr;   // still a valid buffer for the ordinary kernels
            *ptr = reinterpret_cast<float *>(p); *win = w; *nccl_owned = true;
            if (getenv("MMV_COMM_TRACE") && ctx->rank == 0)
                fprintf(stderr, "[trace] gradient buffer: ncclMemAlloc %zu MB, symmetric window %s\n", rounded >> 20, w ? "registered" : "NOT registered");
            return 0;
        }
        cudaGetLastError();
    }
    MMV_CUDA(cudaMalloc(ptr, bytes));
    return 0;
}

void
comm_free_grads(mmv_ctx *ctx, float *ptr, void *win, bool nccl_owned)
{
    if (!ptr) return;
    if (win && ctx->nccl_comm && g_nccl.CommWindowDeregister)
        g_nccl.CommWindowDeregister((ncclComm_t)ctx->nccl_comm, (ncclWindow_t)win);
    if (nccl_owned && g_nccl.MemFree) g_nccl.MemFree(ptr);
    else cudaFree(ptr);
}

// Several reductions as ONE NCCL launch on the comm stream (the step's small leftovers: each would
// otherwise pay its own launch + rendezvous latency in the exposed tail).
int
comm_group_begin(mmv_ctx *ctx)
{
    if (ctx->nranks <= 1) return 0;
    MMV_CUDA(cudaEventRecord(ctx->ev_fork, ctx->stream));
    MMV_CUDA(cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_fork, 0));
    trace_mark(ctx, "grouped allreduce begin", true);
    MMV_NCCL(g_nccl.GroupStart());
    ctx->in_group = true;
    return 0;
}

int
comm_group_end(mmv_ctx *ctx)
{
    if (ctx->nranks <= 1) return 0;
    ctx->in_group = false;
    MMV_NCCL(g_nccl.GroupEnd());
    trace_mark(ctx, "grouped allreduce end", true);
    MMV_CUDA(cudaEventRecord(ctx->ev_join, ctx->comm_stream));
    return 0;
}

int
comm_allreduce_f64_async(mmv_ctx *ctx, double *buf, size_t n)
{
    if (ctx->nranks <= 1 || n == 0) return 0;
    MMV_CHECK(ctx->in_group, "comm_allreduce_f64_async: only inside a group");
    MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat64, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->comm_stream));
    return 0;
}

int
comm_allreduce_f32_async(mmv_ctx *ctx, float *buf, size_t n, bool as_bf16, const float **ss_partial, int *ss_count)
{
    if (ss_count) *ss_count = 0;
    if (ctx->nranks <= 1 || n == 0) return 0;
    if (ctx->in_group) {
        MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat32, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->comm_stream));
        return 0;
    }
    // bf16 on the wire only where it pays (large buckets) and where the layout allows the 32-byte vector casts
    as_bf16 = as_bf16 && n >= ((size_t)1 << 20) && n % 8 == 0 && (reinterpret_cast<uintptr_t>(buf) & 31) == 0;
    if (as_bf16 && ctx->comm_stage_bytes < n * 2) {
        // (grows to the largest bucket during the first step; freed with the communicator)
        MMV_CUDA(cudaStreamSynchronize(ctx->comm_stream));
        if (ctx->comm_stage) cudaFree(ctx->comm_stage);
        ctx->comm_stage = nullptr; ctx->comm_stage_bytes = 0;
        MMV_CUDA(cudaMalloc(&ctx->comm_stage, n * 2));
        ctx->comm_stage_bytes = n * 2;
    }
    MMV_CUDA(cudaEventRecord(ctx->ev_fork, ctx->stream));
    MMV_CUDA(cudaStreamWaitEvent(ctx->comm_stream, ctx->ev_fork, 0));
    if (ctx->trace_on) {
        char nm[64];
        snprintf(nm, sizeof nm, "allreduce %.1f MB%s begin", n * 4e-6, as_bf16 ? " (bf16 on the wire)" : "");
        trace_mark(ctx, nm, true);
    }
    if (as_bf16) {
        // few CTAs on purpose: the casts share the SMs and HBM with the backward kernels they run under
        const int grid = (int)min_sz((n / 8 + 255) / 256, (size_t)ctx->num_sms * 2);
        pack_bf16_kernel<<<grid, 256, 0, ctx->comm_stream>>>(buf, reinterpret_cast<uint4 *>(ctx->comm_stage), n / 8);
        MMV_CUDA(cudaGetLastError());
        MMV_NCCL(g_nccl.AllReduce(ctx->comm_stage, ctx->comm_stage, n, ncclBfloat16, ncclSum, (ncclComm_t)ctx->nccl_comm,
                                  ctx->comm_stream));
        if (ss_partial && ss_count && ctx->comm_ss && ctx->comm_ss_used + grid <= ctx->comm_ss_cap) {
            float *part = ctx->comm_ss + ctx->comm_ss_used;
            unpack_bf16_ss_kernel<<<grid, 256, 0, ctx->comm_stream>>>(reinterpret_cast<const uint4 *>(ctx->comm_stage), buf, n / 8, part);
            ctx->comm_ss_used += grid;
            *ss_partial = part; *ss_count = grid;
        } else {
            unpack_bf16_kernel<<<grid, 256, 0, ctx->comm_stream>>>(reinterpret_cast<const uint4 *>(ctx->comm_stage), buf, n / 8);
        }
        MMV_CUDA(cudaGetLastError());
    } else {
        MMV_NCCL(g_nccl.AllReduce(buf, buf, n, ncclFloat32, ncclSum, (ncclComm_t)ctx->nccl_comm, ctx->comm_stream));
    }
    trace_mark(ctx, "allreduce end", true);
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
    // under backward (SMs and HBM shared with the GEMMs) the ring all-reduce finishes the 123 MB buckets sooner than
    // the NVLS choice NCCL makes on its own: 3.41 vs 3.59 ms/step on 8 B200s.  The caller's setting wins.
    setenv("NCCL_ALGO", "Ring", 0);
    MMV_NCCL(g_nccl.CommInitRank(&comm, nranks, u, rank));
    ctx->nccl_comm = comm;
    // highest priority: the reductions run UNDER backward, whose GEMMs keep every SM busy; at default priority NCCL's
    // CTAs queue behind them and the first 123 MB bucket takes 0.64 ms instead of 0.37 (profiles/r02n8_*_trace.txt)
    int prio_least = 0, prio_greatest = 0;
    MMV_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    {
        const char *e = getenv("MMV_COMM_PRIORITY");
        const bool high = e ? atoi(e) != 0 : true;
        MMV_CUDA(cudaStreamCreateWithPriority(&ctx->comm_stream, cudaStreamNonBlocking, high ? prio_greatest : prio_least));
    }
    MMV_CUDA(cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    MMV_CUDA(cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    MMV_CUDA(cudaEventRecord(ctx->ev_join, ctx->comm_stream));
    ctx->comm_ss_cap = 16 * 1024;
    MMV_CUDA(cudaMalloc(&ctx->comm_ss, sizeof(float) * ctx->comm_ss_cap));
    ctx->comm_ss_used = 0;
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
    if (ctx->comm_stage) cudaFree(ctx->comm_stage);
    ctx->comm_stage = nullptr; ctx->comm_stage_bytes = 0;
    if (ctx->comm_ss) cudaFree(ctx->comm_ss);
    ctx->comm_ss = nullptr; ctx->comm_ss_cap = ctx->comm_ss_used = 0;
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
