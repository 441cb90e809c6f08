// Device context, resident CSR storage, timers, error string — the plumbing
// half of the C ABI declared in include/mmvae_b200.h.
#include "common.cuh"
#include "internal.hh"

#include <atomic>
#include <cstdarg>
#include <map>
#include <mutex>
#include <vector>

namespace mmv {

static thread_local char g_err[1024] = "";

void
set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

const char *
get_error()
{
    return g_err;
}

static std::atomic<long long> g_launches{0};   // host threads of a multi-GPU process launch concurrently

void
count_launch()
{
    g_launches.fetch_add(1, std::memory_order_relaxed);
}

bool
pdl_enabled()
{
    static const bool on = [] { const char *e = getenv("MMV_PDL"); return e ? atoi(e) != 0 : true; }();
    return on;
}

long long
launch_count()
{
    return g_launches.load(std::memory_order_relaxed);
}

static std::mutex g_dev_mu;
static std::map<int, int> g_dev_sms;
static std::map<std::pair<int, const void *>, int> g_dev_smem;

int
dev_num_sms(int *out)
{
    int dev = 0;
    MMV_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_dev_mu);
    auto it = g_dev_sms.find(dev);
    if (it == g_dev_sms.end()) {
        int n = 0;
        MMV_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
        it = g_dev_sms.emplace(dev, n).first;
    }
    *out = it->second;
    return 0;
}

int
dev_set_smem(const void *func, int bytes)
{
    int dev = 0;
    MMV_CUDA(cudaGetDevice(&dev));
    std::lock_guard<std::mutex> lk(g_dev_mu);
    int &have = g_dev_smem[std::make_pair(dev, func)];
    if (have < bytes) {
        MMV_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
        have = bytes;
    }
    return 0;
}

} // namespace mmv

using namespace mmv;

extern "C" {

const char *
mmv_last_error(void)
{
    return get_error();
}

int
mmv_device_count(int *out)
{
    MMV_CHECK(out != nullptr, "mmv_device_count: out is NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    MMV_CHECK(e == cudaSuccess, "mmv_device_count: %s", cudaGetErrorString(e));
    *out = ndev;
    return 0;
}

int
mmv_ctx_create(int device, mmv_ctx **out)
{
    MMV_CHECK(out != nullptr, "mmv_ctx_create: out is NULL");
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    MMV_CHECK(e == cudaSuccess && ndev > 0,
              "mmv_ctx_create: no CUDA device (%s); this library has no CPU fallback",
              e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    MMV_CHECK(device >= 0 && device < ndev, "mmv_ctx_create: device %d out of range [0,%d)", device, ndev);
    MMV_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    MMV_CUDA(cudaGetDeviceProperties(&prop, device));
    MMV_CHECK(prop.major == 10, "mmv_ctx_create: built for sm_100a (B200), found sm_%d%d", prop.major, prop.minor);
    mmv_ctx *c = new mmv_ctx();
    c->device = device;
    c->num_sms = prop.multiProcessorCount;
    // the step's stream sits one level above the lowest priority, the side stream on the lowest: where both have CTAs
    // pending, the step's dependency chain goes first (the comm stream, comm.cu, takes the highest)
    int prio_least = 0, prio_greatest = 0;
    MMV_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    MMV_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_greatest < prio_least ? prio_least - 1 : prio_least));
    MMV_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    MMV_CUDA(cudaStreamCreateWithPriority(&c->side_stream, cudaStreamNonBlocking, prio_least));
    MMV_CUDA(cudaEventCreateWithFlags(&c->ev_side_fork, cudaEventDisableTiming));
    MMV_CUDA(cudaEventCreateWithFlags(&c->ev_side_join, cudaEventDisableTiming));
    MMV_CUDA(cudaEventCreate(&c->ev0));
    MMV_CUDA(cudaEventCreate(&c->ev1));
    *out = c;
    return 0;
}

int
mmv_ctx_destroy(mmv_ctx *c)
{
    if (!c) return 0;
    cudaSetDevice(c->device);
    mmv_comm_destroy(c);
    cudaStreamSynchronize(c->stream);
    cudaStreamSynchronize(c->side_stream);
    cudaEventDestroy(c->ev0);
    cudaEventDestroy(c->ev1);
    cudaEventDestroy(c->ev_side_fork);
    cudaEventDestroy(c->ev_side_join);
    cudaStreamDestroy(c->stream);
    cudaStreamDestroy(c->copy_stream);
    cudaStreamDestroy(c->side_stream);
    delete c;
    return 0;
}

int
mmv_launch_count(int64_t *out)
{
    MMV_CHECK(out, "mmv_launch_count: NULL");
    *out = launch_count();
    return 0;
}

int
mmv_ctx_sync(mmv_ctx *c)
{
    MMV_CHECK(c, "mmv_ctx_sync: NULL ctx");
    MMV_CUDA(cudaSetDevice(c->device));
    MMV_CUDA(cudaStreamSynchronize(c->copy_stream));
    MMV_CUDA(cudaStreamSynchronize(c->stream));
    return 0;
}

int
mmv_ctx_num_sms(mmv_ctx *c, int *out)
{
    MMV_CHECK(c && out, "mmv_ctx_num_sms: NULL");
    *out = c->num_sms;
    return 0;
}

int
mmv_timer_start(mmv_ctx *c)
{
    MMV_CHECK(c, "mmv_timer_start: NULL ctx");
    MMV_CUDA(cudaSetDevice(c->device));
    MMV_CUDA(cudaEventRecord(c->ev0, c->stream));
    return 0;
}

int
mmv_timer_stop(mmv_ctx *c, float *ms)
{
    MMV_CHECK(c && ms, "mmv_timer_stop: NULL");
    MMV_CUDA(cudaSetDevice(c->device));
    MMV_CUDA(cudaEventRecord(c->ev1, c->stream));
    MMV_CUDA(cudaEventSynchronize(c->ev1));
    MMV_CUDA(cudaEventElapsedTime(ms, c->ev0, c->ev1));
    return 0;
}

int
mmv_host_alloc(void **out, int64_t bytes)
{
    MMV_CHECK(out && bytes >= 0, "mmv_host_alloc: bad arguments");
    MMV_CUDA(cudaMallocHost(out, (size_t)(bytes > 0 ? bytes : 1)));
    return 0;
}

int
mmv_host_free(void *p)
{
    if (p) MMV_CUDA(cudaFreeHost(p));
    return 0;
}

// ------------------------------------------------------------------ CSR
// One warp per cell.  flag[0] is raised when gene ids are not strictly increasing inside a cell (the K4/K5
// kernels bisect a cell's nonzeros by gene slice when it stays 0); flag[1] when a gene id lies outside
// [0, n_genes) — K1/K4/K5 index weight rows, shared-memory slices and column sums with it.  With `sanitize`
// (arrays owned by the library: the streamed-minibatch slot) such an entry is neutralised in place (gene 0,
// value 0: it contributes nothing) so no kernel can go out of bounds; the flag still fails the step.
__global__ void
csr_check_sorted_kernel(const int64_t *indptr, int32_t *indices, float *vals, int64_t n_cells, int32_t n_genes, int sanitize,
                        int *flag)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    bool bad = false, oob = false;
    for (int64_t i = warp; i < n_cells; i += nwarp) {
        const int64_t b = indptr[i], e = indptr[i + 1];
        if (lane == 0 && e > b) {                          // the row's first entry has no predecessor
            const int j = indices[b];
            if (j < 0 || j >= n_genes) {
                oob = true;
                if (sanitize) { indices[b] = 0; vals[b] = 0.f; }
            }
        }
        for (int64_t q0 = b + 1; q0 < e; q0 += 32 * 8) {   // 8 independent loads in flight per lane
            int cur[8], prev[8];
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int64_t q = q0 + lane + 32 * u;
                cur[u] = q < e ? indices[q] : 0x7fffffff;
                prev[u] = q < e ? indices[q - 1] : -1;
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int64_t q = q0 + lane + 32 * u;
                bad |= cur[u] <= prev[u];
                if (q < e && (cur[u] < 0 || cur[u] >= n_genes)) {
                    oob = true;
                    if (sanitize) { indices[q] = 0; vals[q] = 0.f; }
                }
            }
        }
    }
    if (__any_sync(0xffffffffu, bad) && lane == 0) flag[0] = 1;
    if (__any_sync(0xffffffffu, oob) && lane == 0) { flag[0] = 1; flag[1] = 1; }   // (a neutralised row is no longer sorted)
}

static int
csr_check_sorted(mmv_csr *s, cudaStream_t st)
{
    if (!s->d_unsorted) MMV_CUDA(cudaMalloc(&s->d_unsorted, 2 * sizeof(int)));
    MMV_CUDA(cudaMemsetAsync(s->d_unsorted, 0, 2 * sizeof(int), st));
    if (s->nnz == 0) return 0;
    const int64_t want = (s->n_cells * 32 + 255) / 256;
    const int grid = (int)(want < 1 ? 1 : (want > 148 * 16 ? 148 * 16 : want));
    csr_check_sorted_kernel<<<grid, 256, 0, st>>>(s->indptr, s->indices, s->vals, s->n_cells, (int32_t)s->n_genes,
                                                 s->borrowed ? 0 : 1, s->d_unsorted);
    ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

static int
csr_validate(int64_t n_cells, int64_t n_genes, const int64_t *indptr, const int32_t *indices, int64_t *max_row,
             bool check_indices)
{
    MMV_CHECK(indptr[0] == 0, "csr: indptr[0] must be 0");
    int64_t mx = 0;
    for (int64_t i = 0; i < n_cells; ++i) {
        const int64_t len = indptr[i + 1] - indptr[i];
        MMV_CHECK(len >= 0, "csr: indptr not monotone at cell %ld", (long)i);
        if (len > mx) mx = len;
    }
    const int64_t nnz = indptr[n_cells];
    for (int64_t k = 0; check_indices && k < nnz; ++k)   // the reference does NOT check rows; we do
        MMV_CHECK(indices[k] >= 0 && indices[k] < n_genes, "csr: gene index %d out of range [0,%ld)",
                  indices[k], (long)n_genes);
    *max_row = mx;
    return 0;
}

int
mmv_csr_create(mmv_ctx *ctx, int64_t max_cells, int64_t max_nnz, int64_t n_genes, int64_t c_dim, mmv_csr **out)
{
    MMV_CHECK(ctx && out, "mmv_csr_create: NULL");
    MMV_CHECK(max_cells > 0 && n_genes > 0 && max_nnz >= 0 && c_dim >= 0, "mmv_csr_create: bad sizes");
    MMV_CHECK(n_genes < (1LL << 31), "mmv_csr_create: n_genes must fit int32");
    MMV_CUDA(cudaSetDevice(ctx->device));
    mmv_csr *s = new mmv_csr();
    s->ctx = ctx;
    s->cap_cells = max_cells;
    s->cap_nnz = max_nnz;
    s->n_genes = n_genes;
    s->c_dim = c_dim;
    MMV_CUDA(cudaMalloc(&s->indptr, sizeof(int64_t) * (max_cells + 1)));
    MMV_CUDA(cudaMalloc(&s->indices, sizeof(int32_t) * (max_nnz > 0 ? max_nnz : 1)));
    MMV_CUDA(cudaMalloc(&s->vals, sizeof(float) * (max_nnz > 0 ? max_nnz : 1)));
    if (c_dim > 0) MMV_CUDA(cudaMalloc(&s->cov, sizeof(float) * max_cells * c_dim));
    *out = s;
    return 0;
}

static int
csr_update_impl(mmv_csr *s, int64_t n_cells, const int64_t *indptr, const int32_t *indices, const float *values,
                const float *covar, bool check_indices)
{
    MMV_CHECK(s && indptr && (indices || indptr[n_cells] == 0) && (values || indptr[n_cells] == 0),
              "mmv_csr_update: NULL");
    MMV_CHECK(n_cells > 0 && n_cells <= s->cap_cells, "mmv_csr_update: %ld cells exceed capacity %ld",
              (long)n_cells, (long)s->cap_cells);
    const int64_t nnz = indptr[n_cells];
    MMV_CHECK(nnz <= s->cap_nnz, "mmv_csr_update: %ld nonzeros exceed capacity %ld", (long)nnz, (long)s->cap_nnz);
    int64_t mx = 0;
    MMV_TRY(csr_validate(n_cells, s->n_genes, indptr, indices, &mx, check_indices));
    // streaming updates (mmv_csr_update) go through the copy stream: they wait for the last step that
    // read this slot and overlap whatever the ctx stream is running on another slot
    const bool streaming = !check_indices;
    cudaStream_t st = streaming ? s->ctx->copy_stream : s->ctx->stream;
    MMV_CUDA(cudaSetDevice(s->ctx->device));
    if (streaming) {
        if (!s->ev_ready) {
            MMV_CUDA(cudaEventCreateWithFlags(&s->ev_ready, cudaEventDisableTiming));
            MMV_CUDA(cudaEventCreateWithFlags(&s->ev_consumed, cudaEventDisableTiming));
            MMV_CUDA(cudaEventRecord(s->ev_consumed, s->ctx->stream));
        }
        MMV_CUDA(cudaStreamWaitEvent(st, s->ev_consumed, 0));
    }
    MMV_CUDA(cudaMemcpyAsync(s->indptr, indptr, sizeof(int64_t) * (n_cells + 1), cudaMemcpyHostToDevice, st));
    if (nnz > 0) {
        MMV_CUDA(cudaMemcpyAsync(s->indices, indices, sizeof(int32_t) * nnz, cudaMemcpyHostToDevice, st));
        MMV_CUDA(cudaMemcpyAsync(s->vals, values, sizeof(float) * nnz, cudaMemcpyHostToDevice, st));
    }
    s->has_cov = covar != nullptr;
    if (covar) {
        MMV_CHECK(s->c_dim > 0, "mmv_csr_update: covariates given but c_dim = 0");
        MMV_CUDA(cudaMemcpyAsync(s->cov, covar, sizeof(float) * n_cells * s->c_dim, cudaMemcpyHostToDevice, st));
    }
    s->n_cells = n_cells;
    s->nnz = nnz;
    if (mx > s->max_row_nnz) s->max_row_nnz = mx;
    s->h2d_bytes = sizeof(int64_t) * (n_cells + 1) + 8 * nnz + (covar ? sizeof(float) * n_cells * s->c_dim : 0);
    MMV_TRY(csr_check_sorted(s, st));
    if (streaming) MMV_CUDA(cudaEventRecord(s->ev_ready, st));
    return 0;
}

// Streaming path: no per-minibatch host scan of the nonzeros (it would cost more than the step itself) — the
// device-side check that already reads every gene id for the sortedness flag also range-checks them,
// neutralises offenders in the slot and raises a flag that fails the step which consumes the slot.
int
mmv_csr_update(mmv_csr *s, int64_t n_cells, const int64_t *indptr, const int32_t *indices, const float *values,
               const float *covar)
{
    return csr_update_impl(s, n_cells, indptr, indices, values, covar, false);
}

int
mmv_csr_upload(mmv_ctx *ctx, int64_t n_cells, int64_t n_genes, const int64_t *indptr, const int32_t *indices,
               const float *values, const float *covar, int64_t c_dim, mmv_csr **out)
{
    MMV_CHECK(ctx && indptr && out, "mmv_csr_upload: NULL");
    MMV_CHECK(n_cells > 0, "mmv_csr_upload: empty matrix");
    mmv_csr *s = nullptr;
    MMV_TRY(mmv_csr_create(ctx, n_cells, indptr[n_cells], n_genes, covar ? c_dim : 0, &s));
    if (!covar) s->c_dim = c_dim > 0 ? c_dim : 1;
    int rc = csr_update_impl(s, n_cells, indptr, indices, values, covar, true);
    if (rc == 0) rc = mmv_ctx_sync(ctx);   // host arrays are borrowed for the call only
    if (rc != 0) {
        mmv_csr_destroy(s);
        return rc;
    }
    *out = s;
    return 0;
}

int
mmv_csr_wrap_device(mmv_ctx *ctx, int64_t n_cells, int64_t n_genes, int64_t nnz, int64_t max_row_nnz,
                    const int64_t *d_indptr, const int32_t *d_indices, const float *d_values, const float *d_covar,
                    int64_t c_dim, mmv_csr **out)
{
    MMV_CHECK(ctx && d_indptr && d_indices && d_values && out, "mmv_csr_wrap_device: NULL");
    MMV_CHECK(n_cells > 0 && n_genes > 0 && n_genes < (1LL << 31) && nnz >= 0 && max_row_nnz >= 0,
              "mmv_csr_wrap_device: bad sizes");
    mmv_csr *s = new mmv_csr();
    s->ctx = ctx;
    s->borrowed = true;
    s->n_cells = s->cap_cells = n_cells;
    s->n_genes = n_genes;
    s->nnz = s->cap_nnz = nnz;
    s->max_row_nnz = max_row_nnz;
    s->indptr = const_cast<int64_t *>(d_indptr);
    s->indices = const_cast<int32_t *>(d_indices);
    s->vals = const_cast<float *>(d_values);
    s->cov = const_cast<float *>(d_covar);
    s->has_cov = d_covar != nullptr;
    s->c_dim = d_covar ? c_dim : (c_dim > 0 ? c_dim : 1);
    MMV_CUDA(cudaSetDevice(ctx->device));
    int flags[2] = { 0, 0 };   // adopted arrays cannot be repaired: a one-off synchronous range check instead
    if (csr_check_sorted(s, ctx->stream) != 0 ||
        cudaMemcpyAsync(flags, s->d_unsorted, sizeof(flags), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess ||
        cudaStreamSynchronize(ctx->stream) != cudaSuccess || flags[1] != 0) {
        if (flags[1]) set_error("csr: gene index out of range [0,%ld) in the adopted device arrays", (long)n_genes);
        cudaFree(s->d_unsorted);
        delete s;
        return 1;
    }
    *out = s;
    return 0;
}

int
mmv_csr_destroy(mmv_csr *s)
{
    if (!s) return 0;
    cudaSetDevice(s->ctx->device);
    cudaStreamSynchronize(s->ctx->copy_stream);
    cudaStreamSynchronize(s->ctx->stream);
    if (!s->borrowed) {
        cudaFree(s->indptr);
        cudaFree(s->indices);
        cudaFree(s->vals);
        cudaFree(s->cov);
    }
    cudaFree(s->d_unsorted);
    if (s->ev_ready) cudaEventDestroy(s->ev_ready);
    if (s->ev_consumed) cudaEventDestroy(s->ev_consumed);
    delete s;
    return 0;
}

} // extern "C"

// ---------------------------------------------------------------- unit-test entry
extern "C" int
mmv_test_gemm(mmv_ctx *ctx, int use_tc, int M, int N, int K, int a_mn, int b_mn, const float *A, const float *B,
              const float *bias, float beta, int relu, float *C)
{
    MMV_CHECK(ctx && A && B && C && M > 0 && N > 0 && K > 0, "mmv_test_gemm: bad arguments");
    MMV_CUDA(cudaSetDevice(ctx->device));
    cudaStream_t st = ctx->stream;
    const size_t na = (size_t)M * K, nb = (size_t)N * K, nc = (size_t)M * N;
    float *dA = nullptr, *dB = nullptr, *dC = nullptr, *dbias = nullptr;
    void *bA = nullptr, *bB = nullptr;
    MMV_CUDA(cudaMalloc(&dA, na * 4)); MMV_CUDA(cudaMalloc(&dB, nb * 4)); MMV_CUDA(cudaMalloc(&dC, nc * 4));
    MMV_CUDA(cudaMemcpyAsync(dA, A, na * 4, cudaMemcpyHostToDevice, st));
    MMV_CUDA(cudaMemcpyAsync(dB, B, nb * 4, cudaMemcpyHostToDevice, st));
    MMV_CUDA(cudaMemcpyAsync(dC, C, nc * 4, cudaMemcpyHostToDevice, st));
    if (bias) {
        MMV_CUDA(cudaMalloc(&dbias, (size_t)N * 4));
        MMV_CUDA(cudaMemcpyAsync(dbias, bias, (size_t)N * 4, cudaMemcpyHostToDevice, st));
    }
    int rc = 0;
    if (use_tc) {
        MMV_CUDA(cudaMalloc(&bA, na * 2)); MMV_CUDA(cudaMalloc(&bB, nb * 2));
        rc = launch_to_bf16(dA, bA, na, st);
        if (!rc) rc = launch_to_bf16(dB, bB, nb, st);
        gemm_tc_args_t g;
        g.A = bA; g.B = bB; g.a_mn = a_mn; g.b_mn = b_mn; g.lda = a_mn ? M : K; g.ldb = b_mn ? N : K;
        g.C = dC; g.ldc = N; g.bias = dbias; g.M = M; g.N = N; g.K = K; g.beta = beta; g.relu = relu;
        if (!rc) rc = launch_gemm_tc(g, st);
    } else {
        gemm_args_t g;
        g.A = dA; g.B = dB; g.transA = a_mn; g.transB = !b_mn; g.lda = a_mn ? M : K; g.ldb = b_mn ? N : K;
        g.C = dC; g.ldc = N; g.bias = dbias; g.M = M; g.N = N; g.K = K; g.beta = beta; g.relu = relu;
        rc = launch_gemm_f32(g, st);
    }
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(C, dC, nc * 4, cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { set_error("mmv_test_gemm: %s", cudaGetErrorString(e)); rc = 1; }
    }
    cudaFree(dA); cudaFree(dB); cudaFree(dC); cudaFree(dbias); cudaFree(bA); cudaFree(bB);
    return rc;
}
