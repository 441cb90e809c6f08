// K0/K1 — minibatch CSR slice + fused log1p / standardisation + sparse x dense
// first encoder layer, forward and backward.
//
// Replaces the reference's dense path
//   mtx_data_block_t::read / torch_tensor  mmvae_io.hh:208-245,81-88  (densify B x D)
//   index_select bootstrap                 mmvae_alg.hh:295-301        (copy rows)
//   encode_mu first Linear                 nb.hh:406-411  (log1p, -x_mean, /sd, Linear(D,h1))
//   encode first Angular                   vmf.hh:253-260, angular.hh:34-42
//   nu_enc(x), depth(x) on raw counts      nb.hh:448,498
// The matrix stays resident as CSR (cells x genes); a minibatch (and a
// bootstrap resample of it) is just a list of cell ids, so nothing is copied.
//   h1_i = sum_{j: x_ij != 0} (log1p(x_ij) / t_j) W1[:, j]  +  (b1 - W1 (x_mean / t))
// W1 is stored transposed, [D x h1], so each nonzero gathers one contiguous row.
#include "common.cuh"
#include "kernels.hh"

#include <cuda_bf16.h>

#include <algorithm>

namespace mmv {

constexpr int kMaxAcc = 16;

// ------------------------------------------------------------------ forward
template <int NACC, bool ANGULAR>
__global__ void
senc_fwd_kernel(senc_args_t a)
{
    extern __shared__ __align__(16) float smem[];
    const int T = blockDim.x, tid = threadIdx.x;
    int *s_g = reinterpret_cast<int *>(smem);      // [T]
    float *s_v = smem + T;                          // [T]
    float *scratch = s_v + T;                       // [32]

    const int row = blockIdx.x;
    const int64_t cell = a.cells ? (int64_t)a.cells[row] : (int64_t)row;
    const int64_t beg = a.indptr[cell], end = a.indptr[cell + 1];
    const int h1 = a.h1;

    float acc[NACC];
#pragma unroll
    for (int q = 0; q < NACC; ++q) acc[q] = 0.f;
    float side_depth = 0.f, side_sq = 0.f;

    for (int64_t k0 = beg; k0 < end; k0 += T) {
        const int cnt = (int)min((int64_t)T, end - k0);
        if (tid < cnt) {
            const int g = a.indices[k0 + tid];
            const float x = a.vals[k0 + tid];
            const float l = log1pf(x);                        // nb.hh:410 / vmf.hh:255
            s_g[tid] = g;
            s_v[tid] = l * a.inv_t[g];
            side_sq = fmaf(l, l, side_sq);
            if (a.wdepth) side_depth = fmaf(x, a.wdepth[g], side_depth);   // nb.hh:498 (raw x)
        }
        __syncthreads();
        int e = 0;
        for (; e + 4 <= cnt; e += 4) {
            const int g0 = s_g[e], g1 = s_g[e + 1], g2 = s_g[e + 2], g3 = s_g[e + 3];
            const float v0 = s_v[e], v1 = s_v[e + 1], v2 = s_v[e + 2], v3 = s_v[e + 3];
#pragma unroll
            for (int q = 0; q < NACC; ++q) {
                const int col = tid + q * T;
                if (col < h1) {
                    float w0 = a.W1T[(size_t)g0 * h1 + col], w1 = a.W1T[(size_t)g1 * h1 + col];
                    float w2 = a.W1T[(size_t)g2 * h1 + col], w3 = a.W1T[(size_t)g3 * h1 + col];
                    if (ANGULAR) {   // angular.hh:38-40: relu(W) + 1e-4 (row norm applied at the end)
                        w0 = fmaxf(w0, 0.f) + 1e-4f; w1 = fmaxf(w1, 0.f) + 1e-4f;
                        w2 = fmaxf(w2, 0.f) + 1e-4f; w3 = fmaxf(w3, 0.f) + 1e-4f;
                    }
                    acc[q] = fmaf(v0, w0, acc[q]);
                    acc[q] = fmaf(v1, w1, acc[q]);
                    acc[q] = fmaf(v2, w2, acc[q]);
                    acc[q] = fmaf(v3, w3, acc[q]);
                }
            }
        }
        for (; e < cnt; ++e) {
            const int g = s_g[e];
            const float v = s_v[e];
#pragma unroll
            for (int q = 0; q < NACC; ++q) {
                const int col = tid + q * T;
                if (col < h1) {
                    float w = a.W1T[(size_t)g * h1 + col];
                    if (ANGULAR) w = fmaxf(w, 0.f) + 1e-4f;
                    acc[q] = fmaf(v, w, acc[q]);
                }
            }
        }
        __syncthreads();
    }

    // per-row scalars: depth pre-activation, ||log1p(x)||, nu-encoder hidden units
    float sd[2] = { side_depth, side_sq };
    block_sum<2>(sd, scratch);
    float row_scale = 1.f;
    if (ANGULAR) {
        const float nrm = fmaxf(sqrtf(sd[1]), 1e-12f);     // F::normalize eps (vmf.hh:255)
        row_scale = 1.f / nrm;
        if (tid == 0 && a.rownorm) a.rownorm[row] = nrm;
    }
    if (tid == 0 && a.dpre) a.dpre[row] = sd[0] + a.bdepth[0];
    for (int h = 0; h < a.H; ++h) {                         // nb.hh:448 (raw x); H is 1 by default
        float s[1] = { 0.f };
        for (int64_t k = beg + tid; k < end; k += T)
            s[0] = fmaf(a.vals[k], a.WnuT[(size_t)a.indices[k] * a.H + h], s[0]);
        block_sum<1>(s, scratch);
        if (tid == 0) a.Hnu[(size_t)row * a.H + h] = s[0] + a.bnu[h];
    }

#pragma unroll
    for (int q = 0; q < NACC; ++q) {
        const int col = tid + q * T;
        if (col < h1) {
            float v = acc[q] * row_scale;
            if (ANGULAR) v *= a.col_scale[col];
            v += a.bias_eff[col];
            if (a.relu) v = fmaxf(v, 0.f);
            a.H1[(size_t)row * h1 + col] = v;
        }
    }
}

static int
pick_threads(int h1, int *nacc)
{
    int T = 32;
    while (T < 256 && T < h1) T *= 2;
    int n = (h1 + T - 1) / T;
    int r = 1;
    while (r < n) r *= 2;
    *nacc = r;
    return T;
}

template <bool ANG>
static int
launch_fwd_t(const senc_args_t &a, int T, int nacc, cudaStream_t st)
{
    const size_t smem = sizeof(float) * (2 * T + 64);
    switch (nacc) {
    case 1: senc_fwd_kernel<1, ANG><<<a.B, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 2: senc_fwd_kernel<2, ANG><<<a.B, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 4: senc_fwd_kernel<4, ANG><<<a.B, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 8: senc_fwd_kernel<8, ANG><<<a.B, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 16: senc_fwd_kernel<16, ANG><<<a.B, T, smem, st>>>(a); ::mmv::count_launch(); break;
    default: MMV_CHECK(false, "sparse encoder: first layer width %d too large (max %d)", a.h1, 256 * kMaxAcc);
    }
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_senc_fwd(const senc_args_t &a, cudaStream_t st)
{
    if (a.B <= 0) return 0;
    int nacc, T = pick_threads(a.h1, &nacc);
    return a.angular ? launch_fwd_t<true>(a, T, nacc, st) : launch_fwd_t<false>(a, T, nacc, st);
}

// ------------------------------------------------- bias correction (mat-vec)
// bias_eff[k] (pre-initialised with b1[k] or 0) -= sum_j (x_mean_j / t_j) Weff[j, k]
__global__ void
senc_bias_kernel(const float *W1T, const float *x_mean, const float *inv_t, const float *col_scale,
                 float *bias_eff, int D, int h1, int genes_per_cta, int angular)
{
    const int j0 = blockIdx.x * genes_per_cta, j1 = min(D, j0 + genes_per_cta);
    if (h1 % 4 == 0 && (reinterpret_cast<uintptr_t>(W1T) & 15) == 0) {
        // 16-byte loads, 8 genes in flight per thread: the [D x h1] weight is read once at HBM speed
        for (int c4 = threadIdx.x; c4 < h1 / 4; c4 += blockDim.x) {
            float s[4] = { 0.f, 0.f, 0.f, 0.f };
            for (int jb = j0; jb < j1; jb += 8) {
                float4 w[8];
                float f[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int j = min(jb + u, j1 - 1);
                    w[u] = *reinterpret_cast<const float4 *>(W1T + (size_t)j * h1 + 4 * c4);
                    f[u] = jb + u < j1 ? x_mean[j] * inv_t[j] : 0.f;
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    float v[4] = { w[u].x, w[u].y, w[u].z, w[u].w };
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        if (angular) v[q] = fmaxf(v[q], 0.f) + 1e-4f;
                        s[q] = fmaf(f[u], v[q], s[q]);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                if (angular) s[q] *= col_scale[4 * c4 + q];
                if (s[q] != 0.f) atomicAdd(bias_eff + 4 * c4 + q, -s[q]);
            }
        }
        return;
    }
    for (int col = threadIdx.x; col < h1; col += blockDim.x) {
        float s = 0.f;
        for (int j = j0; j < j1; ++j) {
            float w = W1T[(size_t)j * h1 + col];
            if (angular) w = fmaxf(w, 0.f) + 1e-4f;
            s = fmaf(x_mean[j] * inv_t[j], w, s);
        }
        if (angular) s *= col_scale[col];
        if (s != 0.f) atomicAdd(bias_eff + col, -s);
    }
}

// The same mat-vec over the bf16 EFFECTIVE weight the tensor-core GEMM multiplies with (parameter mirror for nb, the
// materialised (relu(W) + 1e-4) * col_scale for vmf): half the bytes, and the bias then cancels exactly what the GEMM
// adds for the mean shift.  A thread owns 8 adjacent columns (one 16-byte load per gene); 256 / (h1 / 8) gene groups per CTA.
constexpr int kSbT = 512, kSbU = 8;   // threads per CTA, 16-byte loads in flight per thread

__global__ void __launch_bounds__(kSbT)
senc_bias_bf16_kernel(const __nv_bfloat16 *W, const float *x_mean, const float *inv_t, float *bias_eff, int D, int h1,
                      int genes_per_cta)
{
    __shared__ float s_part[kSbT * 8];
    pdl_sync();
    const int nc8 = h1 / 8, ngrp = kSbT / nc8;
    const int cg = threadIdx.x % nc8, grp = threadIdx.x / nc8;
    const int j0 = blockIdx.x * genes_per_cta, j1 = min(D, j0 + genes_per_cta);
    float s[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) s[q] = 0.f;
    if (grp < ngrp) {
        for (int jb = j0 + grp; jb < j1; jb += kSbU * ngrp) {
            uint4 w[kSbU];
            float f[kSbU];
#pragma unroll
            for (int u = 0; u < kSbU; ++u) {
                const int j = jb + u * ngrp;
                const bool ok = j < j1;
                w[u] = ok ? __ldg(reinterpret_cast<const uint4 *>(W + (size_t)j * h1 + 8 * cg)) : make_uint4(0u, 0u, 0u, 0u);
                f[u] = ok ? x_mean[j] * inv_t[j] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < kSbU; ++u) {
                const uint32_t v[4] = { w[u].x, w[u].y, w[u].z, w[u].w };
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    s[2 * q] = fmaf(f[u], __uint_as_float(v[q] << 16), s[2 * q]);
                    s[2 * q + 1] = fmaf(f[u], __uint_as_float(v[q] & 0xFFFF0000u), s[2 * q + 1]);
                }
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) s_part[threadIdx.x * 8 + q] = s[q];
    __syncthreads();
    if ((int)threadIdx.x < nc8) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            float t = 0.f;
            for (int g = 0; g < ngrp; ++g) t += s_part[(g * nc8 + threadIdx.x) * 8 + q];
            if (t != 0.f) atomicAdd(bias_eff + 8 * threadIdx.x + q, -t);
        }
    }
}

bool
senc_bias_bf16_supported(int h1)
{
    return h1 % 8 == 0 && h1 / 8 <= kSbT && kSbT % (h1 / 8) == 0;
}

int
launch_senc_bias_bf16(const void *W1b, const float *x_mean, const float *inv_t, float *bias_eff, int D, int h1,
                      cudaStream_t st)
{
    MMV_CHECK(senc_bias_bf16_supported(h1) && (reinterpret_cast<uintptr_t>(W1b) & 15) == 0,
              "senc_bias (bf16): first width %d not supported", h1);
    // two CTAs per SM, every thread with eight independent 16-byte loads in flight: the [D x h1] bf16 matrix streams once
    const int gpc = std::max(8, (ceil_div(D, 296) + 7) / 8 * 8);
    MMV_CUDA(launch_k(senc_bias_bf16_kernel, dim3(ceil_div(D, gpc)), dim3(kSbT), 0, st, reinterpret_cast<const __nv_bfloat16 *>(W1b),
                      x_mean, inv_t, bias_eff, D, h1, gpc));
    ::mmv::count_launch();
    return 0;
}

int
launch_senc_bias(const float *W1T, const float *x_mean, const float *inv_t, const float *col_scale,
                 float *bias_eff, int D, int h1, int angular, cudaStream_t st)
{
    // ~2 CTAs per SM, each walking a contiguous gene range with register accumulators: the per-column
    // atomics at the end (h1 per CTA, all CTAs on the same h1 addresses) stay a small part of the pass
    const int gpc = std::max(8, (ceil_div(D, 296) + 7) / 8 * 8);
    senc_bias_kernel<<<ceil_div(D, gpc), 256, 0, st>>>(W1T, x_mean, inv_t, col_scale, bias_eff, D, h1, gpc, angular); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// --------------------------------------------- minibatch transpose (gene-major)
// The backward of the first layer is dW1[:, j] = sum_{i: x_ij != 0} log1p(x_ij) delta_i:
// a scatter in cell-major order.  Re-bucketing the minibatch's nonzeros by gene
// turns it into a gather with one writer per output row (no atomics on D x h1).
__global__ void
csc_count_kernel(const int64_t *indptr, const int32_t *indices, const int32_t *cells, int B, int *counts)
{
    const int row = blockIdx.x;
    const int64_t cell = cells ? (int64_t)cells[row] : (int64_t)row;
    for (int64_t k = indptr[cell] + threadIdx.x; k < indptr[cell + 1]; k += blockDim.x)
        atomicAdd(counts + indices[k], 1);
}

// single-block exclusive scan of counts[D] -> ptr[D+1], cursor[D] = ptr[0..D)
__global__ void
csc_scan_kernel(const int *counts, int *ptr, int *cursor, int D)
{
    __shared__ int s_part[1024];
    const int tid = threadIdx.x, T = blockDim.x;
    const int per = (D + T - 1) / T;
    const int lo = min(D, tid * per), hi = min(D, lo + per);
    int s = 0;
    for (int j = lo; j < hi; ++j) s += counts[j];
    s_part[tid] = s;
    __syncthreads();
    if (tid == 0) {
        int run = 0;
        for (int t = 0; t < T; ++t) { const int v = s_part[t]; s_part[t] = run; run += v; }
        ptr[D] = run;
    }
    __syncthreads();
    int run = s_part[tid];
    for (int j = lo; j < hi; ++j) { ptr[j] = run; cursor[j] = run; run += counts[j]; }
}

__global__ void
csc_fill_kernel(const int64_t *indptr, const int32_t *indices, const float *vals, const int32_t *cells,
                const float *rownorm, int B, int *cursor, int *ent_row, float *ent_l, float *ent_x)
{
    const int row = blockIdx.x;
    const int64_t cell = cells ? (int64_t)cells[row] : (int64_t)row;
    const float rs = rownorm ? 1.f / rownorm[row] : 1.f;
    for (int64_t k = indptr[cell] + threadIdx.x; k < indptr[cell + 1]; k += blockDim.x) {
        const int pos = atomicAdd(cursor + indices[k], 1);
        const float x = vals[k];
        ent_row[pos] = row;
        ent_l[pos] = log1pf(x) * rs;
        ent_x[pos] = x;
    }
}

int
launch_batch_csc(const batch_csc_args_t &c, cudaStream_t st)
{
    MMV_CUDA(cudaMemsetAsync(c.counts, 0, sizeof(int) * c.D, st));
    csc_count_kernel<<<c.B, 128, 0, st>>>(c.indptr, c.indices, c.cells, c.B, c.counts); ::mmv::count_launch();
    csc_scan_kernel<<<1, 1024, 0, st>>>(c.counts, c.gene_ptr, c.cursor, c.D); ::mmv::count_launch();
    csc_fill_kernel<<<c.B, 128, 0, st>>>(c.indptr, c.indices, c.vals, c.cells, c.rownorm, c.B, c.cursor,
                                         c.ent_row, c.ent_l, c.ent_x); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// ----------------------------------------------------------------- backward
// One CTA per gene j.  S[:] = sum_e l_e * delta[row_e, :]  (gather from the L2-resident delta)
//   dWeff[j, :]  = (S - x_mean_j * Delta) / t_j                      
//   d x_mean_j   = -(Weff_j . Delta) / t_j
//   d t_j        = -(Weff_j . S - x_mean_j Weff_j . Delta) / t_j^2 ; d ln_x_sd_j = d t_j * sigmoid(ln_x_sd_j)
//   d WnuT[j, h] = sum_e x_e dHnu[row_e, h] ;  d wdepth_j = sum_e x_e ddpre[row_e]
template <int NACC>
__global__ void
senc_bwd_kernel(senc_bwd_args_t a)
{
    extern __shared__ __align__(16) float smem[];
    const int T = blockDim.x, tid = threadIdx.x;
    int *s_r = reinterpret_cast<int *>(smem);
    float *s_l = smem + T;
    float *scratch = s_l + T;   // [64]
    const int j = blockIdx.x, h1 = a.h1;
    const int beg = a.gene_ptr[j], end = a.gene_ptr[j + 1];

    float acc[NACC];
#pragma unroll
    for (int q = 0; q < NACC; ++q) acc[q] = 0.f;
    float side_depth = 0.f;

    for (int k0 = beg; k0 < end; k0 += T) {
        const int cnt = min(T, end - k0);
        if (tid < cnt) {
            const int r = a.ent_row[k0 + tid];
            s_r[tid] = r;
            s_l[tid] = a.ent_l[k0 + tid];
            if (a.ddpre) side_depth = fmaf(a.ent_x[k0 + tid], a.ddpre[r], side_depth);
        }
        __syncthreads();
        int e = 0;
        for (; e + 4 <= cnt; e += 4) {
            const int r0 = s_r[e], r1 = s_r[e + 1], r2 = s_r[e + 2], r3 = s_r[e + 3];
            const float l0 = s_l[e], l1 = s_l[e + 1], l2 = s_l[e + 2], l3 = s_l[e + 3];
#pragma unroll
            for (int q = 0; q < NACC; ++q) {
                const int col = tid + q * T;
                if (col < h1) {
                    const float d0 = a.delta[(size_t)r0 * h1 + col], d1 = a.delta[(size_t)r1 * h1 + col];
                    const float d2 = a.delta[(size_t)r2 * h1 + col], d3 = a.delta[(size_t)r3 * h1 + col];
                    acc[q] = fmaf(l0, d0, acc[q]);
                    acc[q] = fmaf(l1, d1, acc[q]);
                    acc[q] = fmaf(l2, d2, acc[q]);
                    acc[q] = fmaf(l3, d3, acc[q]);
                }
            }
        }
        for (; e < cnt; ++e) {
            const int r = s_r[e];
            const float l = s_l[e];
#pragma unroll
            for (int q = 0; q < NACC; ++q) {
                const int col = tid + q * T;
                if (col < h1) acc[q] = fmaf(l, a.delta[(size_t)r * h1 + col], acc[q]);
            }
        }
        __syncthreads();
    }

    const float it = a.inv_t[j], xm = a.x_mean[j];
    float dots[3] = { 0.f, 0.f, side_depth };   // W.S, W.Delta, depth grad
#pragma unroll
    for (int q = 0; q < NACC; ++q) {
        const int col = tid + q * T;
        if (col < h1) {
            float w = a.W1T[(size_t)j * h1 + col];
            if (a.angular) w = (fmaxf(w, 0.f) + 1e-4f) * a.col_scale[col];
            const float dl = a.Delta[col];
            dots[0] = fmaf(w, acc[q], dots[0]);
            dots[1] = fmaf(w, dl, dots[1]);
            const float dw = (acc[q] - xm * dl) * it;
            if (a.dW1T) {
                float *dst = a.dW1T + (size_t)j * h1 + col;
                *dst = a.accumulate ? *dst + dw : dw;
            }
        }
    }
    block_sum<3>(dots, scratch);
    if (tid == 0) {
        a.dx_mean[j] += -dots[1] * it;
        const float dt = -(dots[0] - xm * dots[1]) * it * it;
        a.dln_x_sd[j] += dt * sigmoid_f(a.ln_x_sd[j]);   // d softplus
        if (a.dwdepth) a.dwdepth[j] += dots[2];
    }
    for (int h = 0; h < a.H; ++h) {
        float s[1] = { 0.f };
        for (int k = beg + tid; k < end; k += T)
            s[0] = fmaf(a.ent_x[k], a.dHnu[(size_t)a.ent_row[k] * a.H + h], s[0]);
        block_sum<1>(s, scratch);
        if (tid == 0) a.dWnuT[(size_t)j * a.H + h] += s[0];
    }
}

int
launch_senc_bwd(const senc_bwd_args_t &a, cudaStream_t st)
{
    int nacc, T = pick_threads(a.h1, &nacc);
    const size_t smem = sizeof(float) * (2 * T + 128);
    switch (nacc) {
    case 1: senc_bwd_kernel<1><<<a.D, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 2: senc_bwd_kernel<2><<<a.D, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 4: senc_bwd_kernel<4><<<a.D, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 8: senc_bwd_kernel<8><<<a.D, T, smem, st>>>(a); ::mmv::count_launch(); break;
    case 16: senc_bwd_kernel<16><<<a.D, T, smem, st>>>(a); ::mmv::count_launch(); break;
    default: MMV_CHECK(false, "sparse encoder bwd: width %d too large", a.h1);
    }
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
