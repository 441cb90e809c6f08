// K2 (fp32 parity mode) — plain FFMA tiled GEMM.
//
// The reference's dense layers are torch::nn::Linear (nb.hh:335-394,
// vmf.hh:357-388) = libtorch CPU sgemm in true fp32; tcgen05 has no fp32 MMA
// (kind::tf32 keeps 10 mantissa bits), so the 1e-4 parity mode runs on the
// FMA pipe.  The bf16 throughput mode is gemm_tc.cu (tcgen05/TMEM/TMA).
//
//   C[M x N] = alpha * op(A) * op(B) + beta * C  (+ bias[N]) (ReLU optional)
//   op(A) is M x K: A row-major (lda) if !TA, else A stored K x M (lda)
//   op(B) is K x N: B row-major K x N (ldb) if !TB, else B stored N x K (ldb)
// Linear forward  Y = X W^T + b : TA=0, TB=1   (W is [out x in])
// Linear dX = dY W              : TA=0, TB=0
// Linear dW = dY^T X            : TA=1, TB=0
#include "common.cuh"
#include "kernels.hh"

namespace mmv {

constexpr int BM = 128, BN = 128, BK = 16, GT = 256, PAD = 4;

template <bool TA, bool TB>
__global__ void __launch_bounds__(GT)
gemm_f32_kernel(gemm_args_t g)
{
    __shared__ __align__(16) float As[BK][BM + PAD];
    __shared__ __align__(16) float Bs[BK][BN + PAD];

    const int tid = threadIdx.x;
    const int tx = tid % 16, ty = tid / 16;         // 16 x 16 threads, 8 x 8 each
    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    const int M = g.M, N = g.N, K = g.K;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    // split-K: blockIdx.z owns a K range and adds its partial product atomically (dW-type products
    // with few output tiles and a long batch dimension would otherwise run on a handful of SMs)
    const int k_per = ((K + (int)gridDim.z - 1) / (int)gridDim.z + BK - 1) / BK * BK;
    const int k_beg = blockIdx.z * k_per, k_end = min(K, k_beg + k_per);
    for (int k0 = k_beg; k0 < k_end; k0 += BK) {
        // ---- stage A tile (BM x BK) as As[k][m]
        for (int idx = tid; idx < BM * BK; idx += GT) {
            int m, k;
            if (TA) { k = idx / BM; m = idx % BM; } else { m = idx / BK; k = idx % BK; }
            const int gm = m0 + m, gk = k0 + k;
            float v = 0.f;
            if (gm < M && gk < k_end)
                v = TA ? g.A[(size_t)gk * g.lda + gm] : g.A[(size_t)gm * g.lda + gk];
            As[k][m] = v;
        }
        // ---- stage B tile (BK x BN) as Bs[k][n]
        for (int idx = tid; idx < BN * BK; idx += GT) {
            int n, k;
            if (TB) { n = idx / BK; k = idx % BK; } else { k = idx / BN; n = idx % BN; }
            const int gn = n0 + n, gk = k0 + k;
            float v = 0.f;
            if (gn < N && gk < k_end)
                v = TB ? g.B[(size_t)gn * g.ldb + gk] : g.B[(size_t)gk * g.ldb + gn];
            Bs[k][n] = v;
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            float av[8], bv[8];
            const float4 a0 = *reinterpret_cast<const float4 *>(&As[k][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4 *>(&As[k][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4 *>(&Bs[k][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4 *>(&Bs[k][64 + tx * 4]);
            av[0] = a0.x; av[1] = a0.y; av[2] = a0.z; av[3] = a0.w;
            av[4] = a1.x; av[5] = a1.y; av[6] = a1.z; av[7] = a1.w;
            bv[0] = b0.x; bv[1] = b0.y; bv[2] = b0.z; bv[3] = b0.w;
            bv[4] = b1.x; bv[5] = b1.y; bv[6] = b1.z; bv[7] = b1.w;
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }

    // ---- epilogue: rows {ty*4..+3, 64+ty*4..+3}, cols {tx*4..+3, 64+tx*4..+3}
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int gm = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int gn = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
            if (gn >= N) continue;
            float v = g.alpha * acc[i][j];
            float *c = g.C + (size_t)gm * g.ldc + gn;
            if (gridDim.z > 1) {   // C already holds beta * C (launcher); bias once
                if (g.bias && blockIdx.z == 0) v += g.bias[gn];
                atomicAdd(c, v);
                continue;
            }
            if (g.bias) v += g.bias[gn];
            if (g.beta != 0.f) v = fmaf(g.beta, *c, v);
            if (g.relu) v = fmaxf(v, 0.f);
            *c = v;
        }
    }
}

int
launch_gemm_f32(const gemm_args_t &g, cudaStream_t stream)
{
    if (g.M <= 0 || g.N <= 0) return 0;
    MMV_CHECK(g.K > 0, "gemm_f32: K=%d", g.K);
    dim3 grid(ceil_div(g.N, BN), ceil_div(g.M, BM));
    const int tiles = (int)(grid.x * grid.y);
    if (tiles <= 32 && g.K >= 1024 && !g.relu && (g.beta == 1.f || (g.beta == 0.f && g.ldc == g.N))) {
        int splits = g.K / 64;                      // few output tiles, long K: spread K over the machine
        if (splits > 256 / tiles) splits = 256 / tiles;
        if (splits > 1) {
            const int per = ceil_div(ceil_div(g.K, splits), BK) * BK;
            splits = ceil_div(g.K, per);
        }
        if (splits > 1) {
            grid.z = splits;
            if (g.beta == 0.f) MMV_CUDA(cudaMemsetAsync(g.C, 0, sizeof(float) * (size_t)g.M * g.N, stream));
        }
    }
    if (!g.transA && !g.transB) gemm_f32_kernel<false, false><<<grid, GT, 0, stream>>>(g);
    else if (!g.transA && g.transB) gemm_f32_kernel<false, true><<<grid, GT, 0, stream>>>(g);
    else if (g.transA && !g.transB) gemm_f32_kernel<true, false><<<grid, GT, 0, stream>>>(g);
    else gemm_f32_kernel<true, true><<<grid, GT, 0, stream>>>(g);
    ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
