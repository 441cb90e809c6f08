// K3 and small helpers: per-gene parameter prep, Gaussian reparameterisation +
// KL (forward and backward), column sums, ReLU mask, Angular weight
// normalisation (forward/backward).
//
// Reference call sites: reparameterize nb.hh:462-472 / vmf.hh:394-404,
// kl_loss nb.hh:533-537 / vmf.hh:410-414, clamp(lnvar,-4,4) nb.hh:414,449 /
// vmf.hh:262, covar_enc nb.hh:412 / vmf.hh:261, softplus(ln_x_sd)+eps
// nb.hh:408-410 / vmf.hh:258, Angular angular.hh:34-42.
#include "common.cuh"
#include "kernels.hh"

#include <cuda_bf16.h>

namespace mmv {

// inv_t[j] = 1 / (softplus(ln_x_sd_j) + eps);  out = a + b + c (nullable) element-wise
__global__ void
gene_prep_kernel(const float *ln_x_sd, float eps, float *inv_t, int D)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < D) inv_t[j] = 1.f / (softplus_f(ln_x_sd[j]) + eps);
}

__global__ void
vec_combine_kernel(const float *a, const float *b, const float *c, float sb, float sc, float *out, int n)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n) out[j] = (a ? a[j] : 0.f) + (b ? sb * b[j] : 0.f) + (c ? sc * c[j] : 0.f);
}

// per-gene constants of K4 in one launch: abias = mu_decoding.bias + covar_decoding.bias + mu_bias (+ sum_c Wcd[j, c] when the
// covariates are the implicit all-ones block and the bias travels with the logits), ubias = nu_decoding.bias - nu_bias
__global__ void
nb_gene_bias_kernel(const float *blast, const float *bcd, const float *mu_bias, const float *Wcd, int C, const float *bnd,
                    const float *nu_bias, float *abias, float *ubias, int D)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= D) return;
    float ab = (blast ? blast[j] : 0.f) + bcd[j] + mu_bias[j];
    if (Wcd)
        for (int c = 0; c < C; ++c) ab += Wcd[(size_t)j * C + c];
    abias[j] = ab;
    ubias[j] = bnd[j] - nu_bias[j];
}

int
launch_nb_gene_bias(const float *blast, const float *bcd, const float *mu_bias, const float *Wcd, int C, const float *bnd,
                    const float *nu_bias, float *abias, float *ubias, int D, cudaStream_t st)
{
    nb_gene_bias_kernel<<<ceil_div(D, 256), 256, 0, st>>>(blast, bcd, mu_bias, Wcd, C, bnd, nu_bias, abias, ubias, D);
    ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_gene_prep(const float *ln_x_sd, float eps, float *inv_t, int D, cudaStream_t st)
{
    gene_prep_kernel<<<ceil_div(D, 256), 256, 0, st>>>(ln_x_sd, eps, inv_t, D); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_vec_combine(const float *a, const float *b, const float *c, float sb, float sc, float *out, int n,
                   cudaStream_t st)
{
    vec_combine_kernel<<<ceil_div(n, 256), 256, 0, st>>>(a, b, c, sb, sc, out, n); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// ---------------------------------------------------------------- reparam + KL
// head = [mean_raw | lnvar_raw]  (B x 2R, from the stacked representation heads)
__global__ void
reparam_fwd_kernel(reparam_args_t a)
{
    const int R = a.R, n = a.B * R;
    float kl = 0.f;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
        const int i = idx / R, r = idx % R;
        float mean = a.head[(size_t)i * 2 * R + r];
        if (a.Wce) {   // + covar_enc(c)  (nb.hh:412,416)
            const int64_t cell = a.cells ? (int64_t)a.cells[i] : (int64_t)i;
            float hc = a.bce[r];
            for (int c = 0; c < a.C; ++c)
                hc = fmaf(a.cov ? a.cov[cell * a.C + c] : 1.f, a.Wce[r * a.C + c], hc);
            mean += hc;
        }
        const float raw = a.head[(size_t)i * 2 * R + R + r];
        const float lv = fminf(fmaxf(raw, -4.f), 4.f);
        const float ev = expf(lv);
        float z = mean;
        if (a.eps) z = fmaf(a.eps[idx], expf(0.5f * lv), mean);   // nb.hh:466-468
        a.mean[idx] = mean;
        a.lnvar[idx] = lv;
        a.z[idx] = z;
        kl += 1.f + lv - mean * mean - ev;                        // nb.hh:536
    }
    __shared__ float scratch[32];
    float k1[1] = { kl };
    block_sum<1>(k1, scratch);
    if (threadIdx.x == 0 && a.kl_acc) atomicAdd(a.kl_acc, (double)(-0.5f * k1[0]));
}

// dhead = [dmean | dlnvar_raw] given dz; loss = (... + w KL) / B
__global__ void
reparam_bwd_kernel(reparam_args_t a)
{
    const int R = a.R, n = a.B * R;
    const float wB = a.kl_weight * a.inv_B;
    for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += gridDim.x * blockDim.x) {
        const int i = idx / R, r = idx % R;
        const float dz = a.dz[idx];
        const float mean = a.mean[idx], lv = a.lnvar[idx];
        const float raw = a.head[(size_t)i * 2 * R + R + r];
        const float dmean = fmaf(wB, mean, dz);
        float dlv = -0.5f * wB * (1.f - expf(lv));
        if (a.eps) dlv = fmaf(dz * 0.5f * a.eps[idx], expf(0.5f * lv), dlv);
        if (!(raw >= -4.f && raw <= 4.f)) dlv = 0.f;              // clamp backward (inclusive)
        a.dhead[(size_t)i * 2 * R + r] = dmean;
        a.dhead[(size_t)i * 2 * R + R + r] = dlv;
    }
}

int
launch_reparam_fwd(const reparam_args_t &a, cudaStream_t st)
{
    const int n = a.B * a.R;
    reparam_fwd_kernel<<<min(ceil_div(n, 256), 1024), 256, 0, st>>>(a); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_reparam_bwd(const reparam_args_t &a, cudaStream_t st)
{
    const int n = a.B * a.R;
    reparam_bwd_kernel<<<min(ceil_div(n, 256), 1024), 256, 0, st>>>(a); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------ column sums
// out[col * (1 + C) + 0]     += sum_i X[i, col]
// out[col * (1 + C) + 1 + c] += sum_i X[i, col] * cov[cell_i, c]      (C may be 0)
// The two destinations are separate arrays: out_sum[N] and out_cov[N x C].
__global__ void
colsum_kernel(colsum_args_t a)
{
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    const int r0 = blockIdx.y * a.rows_per_cta, r1 = min(a.M, r0 + a.rows_per_cta);
    if (col >= a.N) return;
    float s = 0.f;
    float sc[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) sc[c] = 0.f;
    const bool gather = a.out_cov != nullptr && a.cov != nullptr;
    for (int ib = r0; ib < r1; ib += 8) {      // 8 rows in flight
        float v[8];
        int64_t cell[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int i = ib + u;
            v[u] = 0.f;
            cell[u] = 0;
            if (i < r1) {
                v[u] = a.Xb ? __bfloat162float(reinterpret_cast<const __nv_bfloat16 *>(a.Xb)[(size_t)i * a.ldx + col])
                            : a.X[(size_t)i * a.ldx + col];
                if (gather) cell[u] = a.cells ? (int64_t)a.cells[i] : (int64_t)i;
            }
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            s += v[u];
            if (gather) {
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    if (c < a.C) sc[c] = fmaf(v[u], a.cov[cell[u] * a.C + c], sc[c]);
            }
        }
    }
    if (a.out_sum) atomicAdd(a.out_sum + col, s);
    if (a.out_sum2) atomicAdd(a.out_sum2 + col, s);
    if (a.out_sum3) atomicAdd(a.out_sum3 + col, s);
    if (a.out_cov)   // all-ones covariates (cov == nullptr): every covariate column sees the plain sum
        for (int c = 0; c < a.C; ++c) atomicAdd(a.out_cov + (size_t)col * a.C + c, gather ? sc[c] : s);
}

// bf16 input, all-ones covariates, N % 8 == 0: each thread owns 8 adjacent columns (one 16-byte load
// per row), 4 rows in flight; a CTA covers 256 columns x rows_per_cta rows.
__global__ void __launch_bounds__(256)
colsum_bf16x8_kernel(colsum_args_t a)
{
    __shared__ float s_part[8][32 * 8 + 1];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int col = (blockIdx.x * 32 + lane) * 8;
    const int r0 = blockIdx.y * a.rows_per_cta, r1 = min(a.M, r0 + a.rows_per_cta);
    float s[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) s[q] = 0.f;
    if (col < a.N) {
        const __nv_bfloat16 *X = reinterpret_cast<const __nv_bfloat16 *>(a.Xb) + col;
        for (int i = r0 + warp; i < r1; i += 32) {
            uint4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                v[u] = i + 8 * u < r1 ? *reinterpret_cast<const uint4 *>(X + (size_t)(i + 8 * u) * a.ldx)
                                      : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const uint32_t w[4] = { v[u].x, v[u].y, v[u].z, v[u].w };
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    s[2 * q] += __uint_as_float(w[q] << 16);
                    s[2 * q + 1] += __uint_as_float(w[q] & 0xFFFF0000u);
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) s_part[warp][lane * 8 + q] = s[q];
    __syncthreads();
    const int c = threadIdx.x;   // 256 threads = 256 columns of the CTA
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += s_part[w][c];
    const int gc = blockIdx.x * 256 + c;
    if (gc < a.N) {
        if (a.out_sum) atomicAdd(a.out_sum + gc, tot);
        if (a.out_sum2) atomicAdd(a.out_sum2 + gc, tot);
        if (a.out_sum3) atomicAdd(a.out_sum3 + gc, tot);
        if (a.out_cov)
            for (int k = 0; k < a.C; ++k) atomicAdd(a.out_cov + (size_t)gc * a.C + k, tot);
    }
}

int
launch_colsum(colsum_args_t a, cudaStream_t st)
{
    if (a.M <= 0 || a.N <= 0) return 0;
    MMV_CHECK(a.C <= 8, "colsum: covariate dim %d > 8 not supported", a.C);
    if (a.Xb && a.cov == nullptr && a.N % 8 == 0 && a.ldx % 8 == 0 && (reinterpret_cast<uintptr_t>(a.Xb) & 15) == 0 &&
        a.N >= 256) {
        a.rows_per_cta = 256;
        dim3 grid(ceil_div(a.N, 256), ceil_div(a.M, a.rows_per_cta));
        colsum_bf16x8_kernel<<<grid, 256, 0, st>>>(a); ::mmv::count_launch();
        MMV_CUDA(cudaGetLastError());
        return 0;
    }
    // narrow matrices (bias gradients of the hidden layers: N = 64 ... 1024): the pass is a chain of dependent row
    // loads, so fewer rows per CTA (more CTAs, more atomics) until the grid covers the machine about four times
    a.rows_per_cta = 64;
    while (a.rows_per_cta > 8 && (int64_t)ceil_div(a.N, 128) * ceil_div(a.M, a.rows_per_cta) < 592) a.rows_per_cta /= 2;
    dim3 grid(ceil_div(a.N, 128), ceil_div(a.M, a.rows_per_cta));
    colsum_kernel<<<grid, 128, 0, st>>>(a); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// ------------------------------------------------------------------- ReLU mask
__global__ void
relu_bwd_kernel(float *dX, const float *Y, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        if (!(Y[i] > 0.f)) dX[i] = 0.f;
}

int
launch_relu_bwd(float *dX, const float *Y, size_t n, cudaStream_t st)
{
    if (n == 0) return 0;
    relu_bwd_kernel<<<(int)min_sz((n + 255) / 256, 4096), 256, 0, st>>>(dX, Y, n); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
