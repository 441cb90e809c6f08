// mmv_gemm_bf16_noalloc — the tensor-core GEMM of mmv_test_gemm for callers that cannot allocate device memory
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the
// reference.
//
// The tcgen05 GEMM (gemm_tc.cu) multiplies bf16 operands, so fp32 inputs need bf16 copies: scratch memory.
// mmv_test_gemm cudaMallocs them per call; here they live in ONE file-scope static __device__ array of
// MMV_NOALLOC_SCRATCH_BYTES (4 MiB), which every GPU that loads this library holds until the process exits.
//   * The caller passes DEVICE pointers for the inputs (A, B, bias) and the output (C); nothing is copied from or to
//     the host, and the call makes no cudaMalloc / cudaFree / allocating library call.
//   * In host code the array's NAME is a host-side stand-in, not a device address (a kernel or a TMA descriptor given
//     it would fault, or silently use the wrong memory where the GPU can reach host memory): the device address comes
//     from cudaGetSymbolAddress, looked up once per device with that device current (scratch_ptr below).
//   * Every call on a device shares the array, and it keeps the last call's contents.  Nothing in it has to start at
//     zero: each call overwrites exactly the bytes it then reads (the bf16 copies of ITS operands).
//   * Calls that overlap on the GPU (two contexts / streams on one device) RACE on the scratch: serialise them.
//     Calls through one mmv_ctx are ordered by its stream.
//   * Sizes beyond the maximum are rejected with an error, never truncated.
#include "common.cuh"
#include "internal.hh"

#include <mutex>

namespace {

static __device__ __align__(256) unsigned char g_noalloc_scratch[MMV_NOALLOC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char *g_scratch_dev[kMaxDevices] = {};   // device address of g_noalloc_scratch on each ordinal (nullptr: not set up)

// once per device, with that device current
int
scratch_ptr(int device, unsigned char **out)
{
    MMV_CHECK(device >= 0 && device < kMaxDevices, "mmv_gemm_bf16_noalloc: device ordinal %d not in [0,%d)", device, kMaxDevices);
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_dev[device]) {
        int cur = -1;
        MMV_CUDA(cudaGetDevice(&cur));
        MMV_CHECK(cur == device, "mmv_gemm_bf16_noalloc: set-up needs device %d current, found %d", device, cur);
        void *p = nullptr;
        MMV_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        MMV_CHECK(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, "mmv_gemm_bf16_noalloc: bad scratch address %p", p);
        g_scratch_dev[device] = reinterpret_cast<unsigned char *>(p);
    }
    *out = g_scratch_dev[device];
    return 0;
}

} // namespace

using namespace mmv;

extern "C" int
mmv_gemm_bf16_noalloc(mmv_ctx *ctx, int M, int N, int K, int a_mn, int b_mn, const float *d_A, const float *d_B,
                      const float *d_bias, float beta, int relu, float *d_C)
{
    MMV_CHECK(ctx && d_A && d_B && d_C && M > 0 && N > 0 && K > 0, "mmv_gemm_bf16_noalloc: bad arguments");
    const size_t na = (size_t)M * K, nb = (size_t)N * K;
    const size_t off_b = (na * 2 + 255) & ~(size_t)255;          // B's copy starts on a 256-byte boundary
    const size_t need = off_b + nb * 2;
    MMV_CHECK(need <= (size_t)MMV_NOALLOC_SCRATCH_BYTES,
              "mmv_gemm_bf16_noalloc: operands of %d x %d x %d need %zu bytes of bf16 scratch, the static array holds %d "
              "(MMV_NOALLOC_SCRATCH_BYTES); use mmv_test_gemm or split the product",
              M, N, K, need, (int)MMV_NOALLOC_SCRATCH_BYTES);
    const int lda = a_mn ? M : K, ldb = b_mn ? N : K;
    MMV_CHECK(lda % 8 == 0 && ldb % 8 == 0, "mmv_gemm_bf16_noalloc: leading dimensions (%d, %d) must be multiples of 8", lda, ldb);
    MMV_CUDA(cudaSetDevice(ctx->device));
    unsigned char *scratch = nullptr;
    MMV_TRY(scratch_ptr(ctx->device, &scratch));
    cudaStream_t st = ctx->stream;
    void *bA = scratch, *bB = scratch + off_b;
    MMV_TRY(launch_to_bf16(d_A, bA, na, st));
    MMV_TRY(launch_to_bf16(d_B, bB, nb, st));
    gemm_tc_args_t g;
    g.A = bA; g.B = bB; g.a_mn = a_mn; g.b_mn = b_mn; g.lda = lda; g.ldb = ldb;
    g.C = d_C; g.ldc = N; g.bias = d_bias; g.M = M; g.N = N; g.K = K; g.beta = beta; g.relu = relu;
    return launch_gemm_tc(g, st);   // asynchronous on the context's stream: mmv_ctx_sync (or a later call on it) orders after it
}
