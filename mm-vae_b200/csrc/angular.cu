// Angular module (modules/angular.hh:34-42): Linear without bias whose weight is
//   W'_k = v_k / max(||v_k||, 1e-12),  v_k = relu(W_k) + 1e-4      (per output unit k)
// backward:
//   dv_k = (dW'_k - W'_k (W'_k . dW'_k)) / ||v_k|| ;  dW_k = dv_k * 1[W_k > 0]
// A weight is addressed as W[k * us + j * is] (unit k, input j): the first
// encoder layer is stored transposed ([D x h1]: us = 1, is = h1), deeper layers
// in libtorch layout ([out x in]: us = in, is = 1).
#include "common.cuh"
#include "kernels.hh"

namespace mmv {

__device__ __forceinline__ float
ang_v(float w)
{
    return fmaxf(w, 0.f) + 1e-4f;
}

// partial[k] += sum over an input chunk of f(k, j); threads run over units (us == 1 layout)
template <int MODE>   // 0: v^2 ; 1: W' * dW'
__global__ void
ang_reduce_units_kernel(const float *W, const float *dWp, const float *scale, float *out,
                        int units, int inputs, size_t is, int chunk)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= units) return;
    const int j0 = blockIdx.y * chunk, j1 = min(inputs, j0 + chunk);
    float s = 0.f;
    for (int j = j0; j < j1; ++j) {
        const float v = ang_v(W[(size_t)j * is + k]);
        if (MODE == 0) s = fmaf(v, v, s);
        else s = fmaf(v * scale[k], dWp[(size_t)j * is + k], s);
    }
    atomicAdd(out + k, s);
}

// one warp per unit (is == 1 layout)
template <int MODE>
__global__ void
ang_reduce_inputs_kernel(const float *W, const float *dWp, const float *scale, float *out,
                         int units, int inputs, size_t us)
{
    const int k = blockIdx.x * (blockDim.x / 32) + threadIdx.x / 32;
    if (k >= units) return;
    float s = 0.f;
    for (int j = threadIdx.x % 32; j < inputs; j += 32) {
        const float v = ang_v(W[(size_t)k * us + j]);
        if (MODE == 0) s = fmaf(v, v, s);
        else s = fmaf(v * scale[k], dWp[(size_t)k * us + j], s);
    }
    s = warp_sum(s);
    if (threadIdx.x % 32 == 0) out[k] += s;
}

__global__ void
ang_finish_scale_kernel(float *scale, int units)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < units) scale[k] = 1.f / fmaxf(sqrtf(scale[k]), 1e-12f);   // F::normalize eps
}

// scale[k] = 1 / max(||v_k||, 1e-12)
int
launch_angular_scale(const float *W, float *scale, int units, int inputs, size_t us, size_t is, cudaStream_t st)
{
    MMV_CUDA(cudaMemsetAsync(scale, 0, sizeof(float) * units, st));
    if (us == 1) {
        const int chunk = 256;
        dim3 grid(ceil_div(units, 128), ceil_div(inputs, chunk));
        ang_reduce_units_kernel<0><<<grid, 128, 0, st>>>(W, nullptr, nullptr, scale, units, inputs, is, chunk); ::mmv::count_launch();
    } else {
        MMV_CHECK(is == 1, "angular: unsupported weight strides");
        ang_reduce_inputs_kernel<0><<<ceil_div(units, 4), 128, 0, st>>>(W, nullptr, nullptr, scale, units, inputs, us); ::mmv::count_launch();
    }
    ang_finish_scale_kernel<<<ceil_div(units, 128), 128, 0, st>>>(scale, units); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// Wp[k, j] = v(W[k, j]) * scale[k]   (libtorch layout only; used by the dense layers)
__global__ void
ang_materialize_kernel(const float *W, const float *scale, float *Wp, int units, int inputs)
{
    const size_t n = (size_t)units * inputs;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        Wp[i] = ang_v(W[i]) * scale[i / inputs];
}

int
launch_angular_materialize(const float *W, const float *scale, float *Wp, int units, int inputs, cudaStream_t st)
{
    const size_t n = (size_t)units * inputs;
    ang_materialize_kernel<<<(int)min_sz((n + 255) / 256, 2048), 256, 0, st>>>(W, scale, Wp, units, inputs); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

__global__ void
ang_bwd_apply_kernel(const float *W, const float *dWp, const float *scale, const float *dot, float *dW,
                     int units, int inputs, size_t us, size_t is)
{
    const size_t n = (size_t)units * inputs;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        // i enumerates memory order; recover (k, j)
        int k, j;
        if (us == 1) { j = (int)(i / is); k = (int)(i % is); } else { k = (int)(i / us); j = (int)(i % us); }
        const size_t addr = (size_t)k * us + (size_t)j * is;
        const float w = W[addr];
        if (w > 0.f) {
            const float wp = ang_v(w) * scale[k];
            dW[addr] += (dWp[addr] - wp * dot[k]) * scale[k];
        }
    }
}

// dW += angular_backward(W, dWp); `dot` is scratch [units]
int
launch_angular_bwd(const float *W, const float *dWp, const float *scale, float *dot, float *dW,
                   int units, int inputs, size_t us, size_t is, cudaStream_t st)
{
    MMV_CUDA(cudaMemsetAsync(dot, 0, sizeof(float) * units, st));
    if (us == 1) {
        const int chunk = 256;
        dim3 grid(ceil_div(units, 128), ceil_div(inputs, chunk));
        ang_reduce_units_kernel<1><<<grid, 128, 0, st>>>(W, dWp, scale, dot, units, inputs, is, chunk); ::mmv::count_launch();
    } else {
        MMV_CHECK(is == 1, "angular: unsupported weight strides");
        ang_reduce_inputs_kernel<1><<<ceil_div(units, 4), 128, 0, st>>>(W, dWp, scale, dot, units, inputs, us); ::mmv::count_launch();
    }
    const size_t n = (size_t)units * inputs;
    ang_bwd_apply_kernel<<<(int)min_sz((n + 255) / 256, 2048), 256, 0, st>>>(W, dWp, scale, dot, dW, units,
                                                                               inputs, us, is); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
