// Shared device/host helpers for the mm-vae B200 kernels (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <string>

namespace mmv {

constexpr int kWarp = 32;

// ---------------------------------------------------------------------------
// error plumbing: every C-ABI entry returns 0 (EXIT_SUCCESS) or nonzero and
// leaves a message retrievable through mmv_last_error (util.hh:35-80 keeps the
// reference's "ELOG + EXIT_FAILURE" convention on the host side).
// ---------------------------------------------------------------------------
void set_error(const char *fmt, ...);
void count_launch();          // one call per kernel launch (mmv_launch_count)
long long launch_count();
const char *get_error();
// per-DEVICE launch state (function attributes and the SM count belong to a device, not to the process: one
// process may hold a context per GPU, mmv_ctx_create(device)).  Both act on the CURRENT device.
int dev_num_sms(int *out);                          // SM count, cached per ordinal
int dev_set_smem(const void *func, int bytes);      // MaxDynamicSharedMemorySize once per (device, kernel)

#define MMV_CUDA(expr)                                                           \
    do {                                                                         \
        cudaError_t _e = (expr);                                                 \
        if (_e != cudaSuccess) {                                                 \
            ::mmv::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr,       \
                             cudaGetErrorString(_e));                            \
            return 1;                                                            \
        }                                                                        \
    } while (0)

#define MMV_CHECK(cond, ...)                 \
    do {                                     \
        if (!(cond)) {                       \
            ::mmv::set_error(__VA_ARGS__);   \
            return 1;                        \
        }                                    \
    } while (0)

#define MMV_TRY(expr)            \
    do {                         \
        int _rc = (expr);        \
        if (_rc != 0) return _rc; \
    } while (0)

inline size_t
min_sz(size_t a, size_t b)
{
    return a < b ? a : b;
}

inline int
ceil_div(int64_t a, int64_t b)
{
    return (int)((a + b - 1) / b);
}

// ---------------------------------------------------------------------------
// Programmatic dependent launch (PDL).  A step is ~30 kernels back to back on one stream; between two of them the GPU
// drains, the next grid is launched and its CTAs ramp up (2-4 us each time).  Kernels launched through launch_k()
// carry cudaLaunchAttributeProgrammaticStreamSerialization: their CTAs may become resident while the previous kernel's
// last CTAs still run, and park at pdl_sync() — griddepcontrol.wait returns only when the previous grid has COMPLETED
// and its memory is visible, so nothing before that point may touch global memory.  pdl_sync() is the first statement
// of every kernel launched this way (a no-op under an ordinary launch).  MMV_PDL=0 turns the attribute off.
// ---------------------------------------------------------------------------
bool pdl_enabled();

__device__ __forceinline__ void
pdl_sync()
{
#if defined(__CUDA_ARCH__)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}

template <typename... P, typename... A>
inline cudaError_t
launch_k(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, A &&...args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, static_cast<P>(args)...);
}

// ---------------------------------------------------------------------------
// device math.  Accurate libdevice versions on purpose: parity with the
// reference's fp32 ATen ops is 1e-4 relative (BASELINE.json north_star).
// ---------------------------------------------------------------------------
__device__ __forceinline__ float
softplus_f(float x)
{   // libtorch softplus, beta=1, threshold=20
    return x > 20.f ? x : log1pf(expf(x));
}

__device__ __forceinline__ float
sigmoid_f(float x)
{
    return 1.f / (1.f + expf(-x));
}

// digamma for x > 0: recurrence up to x >= 6, then the asymptotic series.
__device__ __forceinline__ float
digamma_f(float x)
{
    float acc = 0.f;
    while (x < 6.f) {
        acc -= 1.f / x;
        x += 1.f;
    }
    const float r = 1.f / x, r2 = r * r;
    // ln x - 1/(2x) - 1/(12x^2) + 1/(120x^4) - 1/(252x^6) + 1/(240x^8)
    const float s = r2 * (1.f / 12.f - r2 * (1.f / 120.f - r2 * (1.f / 252.f - r2 * (1.f / 240.f))));
    return acc + logf(x) - 0.5f * r - s;
}

// counter-based N(0,1) of the library's own reparameterisation noise (MMV_EPS_PHILOX): element i of draw `step` under
// `seed` — splitmix64 finaliser, Box-Muller.  Any element is reproducible on its own, so the kernel that consumes the
// noise can draw it in place.
__device__ __forceinline__ uint32_t
mix32(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    z = z ^ (z >> 31);
    return (uint32_t)(z >> 16);
}

__device__ __forceinline__ float
normal_draw(uint64_t seed, uint64_t step, uint64_t i)
{
    const uint64_t key = seed * 0x9E3779B97F4A7C15ull + step * 0xD1B54A32D192ED03ull + i * 2;
    const float u1 = ((float)mix32(key) + 1.f) * (1.f / 4294967808.f);          // (0, 1)
    const float u2 = (float)mix32(key + 1) * (1.f / 4294967296.f);
    return sqrtf(-2.f * logf(u1)) * cospif(2.f * u2);                            // Box-Muller
}

// ---------------------------------------------------------------------------
// warp / block reductions
// ---------------------------------------------------------------------------
__device__ __forceinline__ float
warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float
warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum of N values per thread; result valid in ALL threads.
// `scratch` must hold N * 32 floats. Contains two __syncthreads.
template <int N>
__device__ __forceinline__ void
block_sum(float (&v)[N], float *scratch)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarp = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int k = 0; k < N; ++k) v[k] = warp_sum(v[k]);
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) scratch[k * 32 + warp] = v[k];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < N; ++k) {
        float t = lane < nwarp ? scratch[k * 32 + lane] : 0.f;
        v[k] = warp_sum(t);
    }
    __syncthreads();
}

__device__ __forceinline__ float
block_max(float v, float *scratch)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int nwarp = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    if (lane == 0) scratch[warp] = v;
    __syncthreads();
    float t = lane < nwarp ? scratch[lane] : -INFINITY;
    v = warp_max(t);
    __syncthreads();
    return v;
}

} // namespace mmv
