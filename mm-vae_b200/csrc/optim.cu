// K6 — fused global-norm clip + Adam (L2 weight decay) over the flat parameter
// buffer.  Replaces torch::nn::utils::clip_grad_norm_ + torch::optim::Adam
// (mmvae_alg.hh:234-236,306-310): Adam beta=(0.9,0.999), eps=1e-8, bias
// corrected, weight_decay added to the gradient (L2, not AdamW);
// clip scale = min(1, max_norm / (||g||_2 + 1e-6)).
#include "common.cuh"
#include "kernels.hh"

#include <cuda_bf16.h>

#include <algorithm>

namespace mmv {

// Order-deterministic: every block leaves its partial in a slot, the last block to finish (ticket) adds the slots
// up in index order.  Data-parallel replicas compute the clip coefficient from this sum: an atomicAdd of the
// partials in arrival order could differ in the last bit from rank to rank and let the replicas drift apart.
constexpr int kSumsqMaxBlocks = 148 * 16;
__device__ float g_sumsq_partial[kSumsqMaxBlocks];
__device__ unsigned int g_sumsq_ticket = 0;

// up to four ranges per launch: the flat gradient buffer minus the tensors whose producers already added their share
struct sumsq_ranges_t {
    const float *g[4];
    size_t n[4];
    int count;
    const float *pre[8];   // already-squared partial sums (added as they are): the unpack passes of bf16 gradient buckets
    int npre[8];
    int pre_count;
};

__global__ void
sumsq_kernel(sumsq_ranges_t rg, double *out)
{
    pdl_sync();
    float s = 0.f;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < rg.count; ++r) {
        const float *g = rg.g[r];
        const size_t n = rg.n[r];
        const size_t n4 = ((reinterpret_cast<uintptr_t>(g) & 15) == 0) ? n / 4 : 0;   // 16-byte loads, 4 in flight
        const float4 *g4 = reinterpret_cast<const float4 *>(g);
        size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
        for (; i + 3 * stride < n4; i += 4 * stride) {
            const float4 a = g4[i], b = g4[i + stride], c = g4[i + 2 * stride], d = g4[i + 3 * stride];
            s += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w + b.x * b.x + b.y * b.y + b.z * b.z + b.w * b.w;
            s += c.x * c.x + c.y * c.y + c.z * c.z + c.w * c.w + d.x * d.x + d.y * d.y + d.z * d.z + d.w * d.w;
        }
        for (; i < n4; i += stride) {
            const float4 a = g4[i];
            s += a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w;
        }
        for (size_t k = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n; k += stride) s = fmaf(g[k], g[k], s);
    }
    for (int r = 0; r < rg.pre_count; ++r)
        for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < (size_t)rg.npre[r]; k += stride) s += rg.pre[r][k];
    __shared__ float scratch[32];
    __shared__ bool last;
    float v[1] = { s };
    block_sum<1>(v, scratch);
    if (threadIdx.x == 0) {
        g_sumsq_partial[blockIdx.x] = v[0];
        __threadfence();
        last = atomicAdd(&g_sumsq_ticket, 1u) == gridDim.x - 1;
    }
    __syncthreads();
    if (!last) return;
    __threadfence();
    if (threadIdx.x < 32) {
        double t = 0.;
        for (unsigned b = threadIdx.x; b < gridDim.x; b += 32) t += (double)__ldcg(&g_sumsq_partial[b]);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) {
            *out += t;              // launches on one stream: the only writer at this point
            g_sumsq_ticket = 0;
        }
    }
}

int
launch_sumsq_ranges(const float *const *g, const size_t *n, int count, double *out, int num_sms, cudaStream_t st,
                    const float *const *pre, const int *npre, int pre_count)
{
    MMV_CHECK(pre_count <= 8, "sumsq: at most 8 pre-summed ranges");
    for (int i0 = 0; i0 < std::max(count, 1); i0 += 4) {
        sumsq_ranges_t rg;
        rg.count = 0;
        rg.pre_count = 0;
        size_t total = 0;
        if (i0 == 0)
            for (int i = 0; i < pre_count; ++i)
                if (npre[i] > 0) { rg.pre[rg.pre_count] = pre[i]; rg.npre[rg.pre_count] = npre[i]; total += (size_t)npre[i]; ++rg.pre_count; }
        for (int i = i0; i < count && i < i0 + 4; ++i) {
            if (n[i] == 0) continue;
            rg.g[rg.count] = g[i];
            rg.n[rg.count] = n[i];
            total += n[i];
            ++rg.count;
        }
        if (rg.count == 0 && rg.pre_count == 0) continue;
        const int grid = (int)min_sz(min_sz((total + 1023) / 1024, (size_t)num_sms * 8), (size_t)kSumsqMaxBlocks);
        MMV_CUDA(launch_k(sumsq_kernel, dim3(grid), dim3(256), 0, st, rg, out)); ::mmv::count_launch();
    }
    return 0;
}

int
launch_sumsq(const float *g, size_t n, double *out, int num_sms, cudaStream_t st)
{
    return launch_sumsq_ranges(&g, &n, 1, out, num_sms, st, nullptr, nullptr, 0);
}

__global__ void
adam_kernel(adam_args_t a)
{
    pdl_sync();
    const float total = (float)sqrt(*a.sumsq);
    float coef = a.max_norm / (total + 1e-6f);
    coef = fminf(coef, 1.f);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (a.norm_out) *a.norm_out = total;
        if (a.coef_out) *a.coef_out = coef;
    }
    // 4 elements per thread per trip (the flat buffers are 16-byte aligned and n is a multiple of 8)
    const bool vec = ((reinterpret_cast<uintptr_t>(a.grad) | reinterpret_cast<uintptr_t>(a.param) |
                       reinterpret_cast<uintptr_t>(a.m) | reinterpret_cast<uintptr_t>(a.v)) & 15) == 0 && a.n % 4 == 0;
    const size_t n4 = vec ? a.n / 4 : 0;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 g4 = reinterpret_cast<float4 *>(a.grad)[i];
        const float4 p4 = reinterpret_cast<const float4 *>(a.param)[i];
        float4 m4 = reinterpret_cast<float4 *>(a.m)[i], v4 = reinterpret_cast<float4 *>(a.v)[i];
        float g[4] = { g4.x * coef, g4.y * coef, g4.z * coef, g4.w * coef };
        const float p[4] = { p4.x, p4.y, p4.z, p4.w };
        float m[4] = { m4.x, m4.y, m4.z, m4.w }, v[4] = { v4.x, v4.y, v4.z, v4.w }, pn[4];
        if (a.write_grad) reinterpret_cast<float4 *>(a.grad)[i] = make_float4(g[0], g[1], g[2], g[3]);   // clip scales .grad in place
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float gw = fmaf(a.weight_decay, p[q], g[q]);
            m[q] = a.beta1 * m[q] + a.one_minus_beta1 * gw;
            v[q] = a.beta2 * v[q] + (a.one_minus_beta2 * gw) * gw;
            const float denom = sqrtf(v[q]) / a.bc2_sqrt + a.eps;
            pn[q] = p[q] + (a.neg_step * m[q]) / denom;
        }
        reinterpret_cast<float4 *>(a.m)[i] = make_float4(m[0], m[1], m[2], m[3]);
        reinterpret_cast<float4 *>(a.v)[i] = make_float4(v[0], v[1], v[2], v[3]);
        reinterpret_cast<float4 *>(a.param)[i] = make_float4(pn[0], pn[1], pn[2], pn[3]);
        if (a.param_bf16) {
            const __nv_bfloat162 lo = __floats2bfloat162_rn(pn[0], pn[1]), hi = __floats2bfloat162_rn(pn[2], pn[3]);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t *>(&lo);
            pk.y = *reinterpret_cast<const uint32_t *>(&hi);
            reinterpret_cast<uint2 *>(a.param_bf16)[i] = pk;
            if (a.param_lo) {   // what the bf16 mirror lost, again in bf16: mirror + residual carry ~16 mantissa bits
                const __nv_bfloat162 rl = __floats2bfloat162_rn(pn[0] - __low2float(lo), pn[1] - __high2float(lo));
                const __nv_bfloat162 rh = __floats2bfloat162_rn(pn[2] - __low2float(hi), pn[3] - __high2float(hi));
                uint2 pr;
                pr.x = *reinterpret_cast<const uint32_t *>(&rl);
                pr.y = *reinterpret_cast<const uint32_t *>(&rh);
                reinterpret_cast<uint2 *>(a.param_lo)[i] = pr;
            }
        }
    }
    for (size_t i = n4 * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (size_t)gridDim.x * blockDim.x) {
        float g = a.grad[i] * coef;
        if (a.write_grad) a.grad[i] = g;     // clip_grad_norm_ scales .grad in place
        const float p = a.param[i];
        g = fmaf(a.weight_decay, p, g);
        const float m = a.beta1 * a.m[i] + a.one_minus_beta1 * g;
        const float v = a.beta2 * a.v[i] + (a.one_minus_beta2 * g) * g;
        a.m[i] = m;
        a.v[i] = v;
        const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
        const float pn = p + (a.neg_step * m) / denom;
        a.param[i] = pn;
        if (a.param_bf16) {
            const __nv_bfloat16 h = __float2bfloat16(pn);
            reinterpret_cast<__nv_bfloat16 *>(a.param_bf16)[i] = h;
            if (a.param_lo) reinterpret_cast<__nv_bfloat16 *>(a.param_lo)[i] = __float2bfloat16(pn - __bfloat162float(h));
        }
    }
}

int
launch_adam(const adam_args_t &a, int num_sms, cudaStream_t st)
{
    if (a.n == 0) return 0;
    const int grid = (int)min_sz((a.n / 4 + 255) / 256 + 1, (size_t)num_sms * 8);
    MMV_CUDA(launch_k(adam_kernel, dim3(grid), dim3(256), 0, st, a)); ::mmv::count_launch();
    return 0;
}

__global__ void
scale_by_kernel(float *x, size_t n, const float *coef)
{
    const float c = *coef;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) x[i] *= c;
}

// materialise a clip scale that launch_adam (write_grad = 0) left pending
int
launch_scale_by(float *x, size_t n, const float *coef, cudaStream_t st)
{
    if (n == 0) return 0;
    scale_by_kernel<<<(int)min_sz((n + 255) / 256, 2048), 256, 0, st>>>(x, n, coef); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
