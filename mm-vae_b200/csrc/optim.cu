// K6 — fused global-norm clip + Adam (L2 weight decay) over the flat parameter
// buffer.  Replaces torch::nn::utils::clip_grad_norm_ + torch::optim::Adam
// (mmvae_alg.hh:234-236,306-310): Adam beta=(0.9,0.999), eps=1e-8, bias
// corrected, weight_decay added to the gradient (L2, not AdamW);
// clip scale = min(1, max_norm / (||g||_2 + 1e-6)).
#include "common.cuh"
#include "kernels.hh"

#include <cuda_bf16.h>

namespace mmv {

__global__ void
sumsq_kernel(const float *g, size_t n, double *out)
{
    float s = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        s = fmaf(g[i], g[i], s);
    __shared__ float scratch[32];
    float v[1] = { s };
    block_sum<1>(v, scratch);
    if (threadIdx.x == 0) atomicAdd(out, (double)v[0]);
}

int
launch_sumsq(const float *g, size_t n, double *out, int num_sms, cudaStream_t st)
{
    if (n == 0) return 0;
    const int grid = (int)min_sz((n + 1023) / 1024, (size_t)num_sms * 8);
    sumsq_kernel<<<grid, 256, 0, st>>>(g, n, out); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

__global__ void
adam_kernel(adam_args_t a)
{
    const float total = (float)sqrt(*a.sumsq);
    float coef = a.max_norm / (total + 1e-6f);
    coef = fminf(coef, 1.f);
    if (blockIdx.x == 0 && threadIdx.x == 0 && a.norm_out) *a.norm_out = total;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += (size_t)gridDim.x * blockDim.x) {
        float g = a.grad[i] * coef;
        a.grad[i] = g;                       // clip_grad_norm_ scales .grad in place
        const float p = a.param[i];
        g = fmaf(a.weight_decay, p, g);
        const float m = a.beta1 * a.m[i] + a.one_minus_beta1 * g;
        const float v = a.beta2 * a.v[i] + (a.one_minus_beta2 * g) * g;
        a.m[i] = m;
        a.v[i] = v;
        const float denom = sqrtf(v) / a.bc2_sqrt + a.eps;
        const float pn = p + (a.neg_step * m) / denom;
        a.param[i] = pn;
        if (a.param_bf16) reinterpret_cast<__nv_bfloat16 *>(a.param_bf16)[i] = __float2bfloat16(pn);
    }
}

int
launch_adam(const adam_args_t &a, int num_sms, cudaStream_t st)
{
    if (a.n == 0) return 0;
    const int grid = (int)min_sz((a.n + 255) / 256, (size_t)num_sms * 8);
    adam_kernel<<<grid, 256, 0, st>>>(a); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
