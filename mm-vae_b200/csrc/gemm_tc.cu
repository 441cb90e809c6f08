// K2 (throughput mode) — bf16 GEMM on the 5th-gen tensor cores.
//
//   C[M x N] (fp32, + optional bf16 copy) = op(A) op(B) (+ bias[N]) (+ beta C) (ReLU optional)
// A and B are bf16 and may each be K-major or MN-major, so the three products
// of a Linear layer (nb.hh:335-394, vmf.hh:357-388 + autograd) need no
// transposed copies:
//   forward  Y  = X W^T : A = X  [M x K]  K-major , B = W [N x K]  K-major
//   dX       dX = dY W  : A = dY [M x K]  K-major , B = W [K x N]  MN-major
//   dW       dW = dY^T X: A = dY [K x M]  MN-major, B = X [K x N]  MN-major
//
// One CTA per 128 x 128 output tile, warp-specialised:
//   warp 0   TMA producer  (cp.async.bulk.tensor.2d, SWIZZLE_128B, mbarrier tx-count)
//   warp 1   MMA issuer    (one elected thread: tcgen05.mma.cta_group::1.kind::f16, D in TMEM)
//   warp 2   TMEM allocator
//   warps 4-7 epilogue     (tcgen05.ld 32x32b -> registers -> shared transpose -> coalesced stores)
// 3-stage shared-memory ring (full/empty mbarriers); two CTAs are co-resident per
// SM so one tile's epilogue overlaps the other's MMAs.  Out-of-range parts of
// edge tiles are zero-filled by TMA, so no padding of M, N or K is needed
// (leading dimensions must be multiples of 8 elements = 16 bytes).
#include "common.cuh"
#include "kernels.hh"
#include "tc_ptx.cuh"

#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <atomic>
#include <cstdlib>

namespace mmv {

namespace tc {

constexpr int BLOCK_M = 128, BLOCK_N = 128, BLOCK_K = 64, UMMA_K = 16;
constexpr int THREADS = 256;
constexpr int A_BYTES = BLOCK_M * BLOCK_K * 2, B_BYTES = BLOCK_N * BLOCK_K * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int TMEM_COLS = 128;
constexpr int smem_bytes(int stages) { return stages * STAGE_BYTES + 1024 /*align slack*/ + 256 /*barriers*/; }

struct params_t {
    float *C;
    __nv_bfloat16 *Cb;
    const float *bias;
    int M, N, K, ldc, ldcb;
    float beta;
    int relu;
    int tma_epi;   // 0: register epilogue ; 1: TMA tile store ; 2: TMA reduce-add (beta = 1 or split-K)
    double *sumsq; // persistent store epilogue: += sum of squares of the product
    float2 *row_stats;   // persistent store epilogue: [tiles_n x M] softmax partials (max log2(e), sum exp2) per row and column tile
    int n_fast;    // persistent tile order: 1 = consecutive tiles walk N (A tile reused from L2), 0 = walk M (B tile reused)
    int segs;      // 1, or 3 = fp32-tolerance mode: the K axis is walked three times (A_hi B_lo, A_lo B_hi, A_hi B_hi) into one accumulator
};

// STAGES = 3: two CTAs per SM (many-tile grids) ; STAGES = 6: one CTA per SM with a deeper ring, for
// single-wave grids whose short K loop is bound by the TMA round trip
template <bool A_MN, bool B_MN, int STAGES>
__global__ void __launch_bounds__(THREADS, STAGES <= 3 ? 2 : 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmAlo,
               const __grid_constant__ CUtensorMap tmBlo, params_t p)
{
    extern __shared__ uint8_t smem_raw[];
    // SWIZZLE_128B operands need 1024-byte alignment
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 1);
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + STAGES), tfull = smem_u32(bars + 2 * STAGES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.y * BLOCK_M, n0 = blockIdx.x * BLOCK_N;
    // split-K (gridDim.z > 1): this CTA owns k-blocks [kb0, kb0 + num_kb) and adds its partial tile atomically
    const int base_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
    const int total_kb = base_kb * p.segs;
    const int kb_per = (total_kb + (int)gridDim.z - 1) / (int)gridDim.z;
    const int kb0 = blockIdx.z * kb_per;
    const int num_kb = max(0, min(total_kb - kb0, kb_per));

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        if (p.tma_epi) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmC) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    pdl_sync();   // (descriptor prefetch, barrier init and the TMEM allocation above overlap the previous kernel's tail)

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                mbar_wait(empty0 + 8 * s, ph ^ 1);
                const uint32_t sa = smem_u32(smem + s * STAGE_BYTES), sb = sa + A_BYTES;
                const uint32_t fb = full0 + 8 * s;
                mbar_expect_tx(fb, STAGE_BYTES);
                // segment of the K walk: 0 = A_hi B_lo, 1 = A_lo B_hi, 2 (and the only one of a plain product) = A_hi B_hi
                const int g = kb0 + kb, seg = p.segs == 3 ? g / base_kb : 2;
                const int k0 = (p.segs == 3 ? g - seg * base_kb : g) * BLOCK_K;
                const CUtensorMap *mA = seg == 1 ? &tmAlo : &tmA, *mB = seg == 0 ? &tmBlo : &tmB;
                if (A_MN) {     // tensor dims {MN, K}: two [64 k x 64 m] boxes
                    tma_load_2d(sa, mA, fb, m0, k0);
                    tma_load_2d(sa + A_BYTES / 2, mA, fb, m0 + 64, k0);
                } else {        // tensor dims {K, MN}: one [128 m x 64 k] box
                    tma_load_2d(sa, mA, fb, k0, m0);
                }
                if (B_MN) {
                    tma_load_2d(sb, mB, fb, n0, k0);
                    tma_load_2d(sb + B_BYTES / 2, mB, fb, n0 + 64, k0);
                } else {
                    tma_load_2d(sb, mB, fb, k0, n0);
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(A_MN, B_MN, BLOCK_M, BLOCK_N);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STAGES;
                const uint32_t ph = (kb / STAGES) & 1;
                mbar_wait(full0 + 8 * s, ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = smem_u32(smem + s * STAGE_BYTES), sb = sa + A_BYTES;
#pragma unroll
                for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                    // K-major  : rows of 128 B, 8-row groups 1024 B apart; a K step of 16 = +32 B
                    // MN-major : [k][64 mn] rows of 128 B, 8-k groups 1024 B apart (SBO), 64-mn
                    //            chunks 8192 B apart (LBO); a K step of 16 = 2 groups = +2048 B
                    const uint64_t ad = A_MN ? make_desc(sa + k * 2048, A_BYTES / 2, 1024)
                                             : make_desc(sa + k * 32, 16, 1024);
                    const uint64_t bd = B_MN ? make_desc(sb + k * 2048, B_BYTES / 2, 1024)
                                             : make_desc(sb + k * 32, 16, 1024);
                    umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0);
                }
                umma_commit(empty0 + 8 * s);          // frees the smem stage when these MMAs finish
            }
            umma_commit(tfull);                        // accumulator complete
        }
    } else if (warp >= 4) {
        // ================= epilogue =================
        mbar_wait(tfull, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int e = warp - 4;                         // TMEM lanes [32e, 32e+32) = tile rows
        if (p.tma_epi) {
            // TMEM -> registers -> [32 x 32] fp32 boxes in shared memory (SWIZZLE_128B: conflict-free
            // 16-byte stores) -> TMA tile store, or TMA reduce-add when the product accumulates
            // (beta = 1, split-K): the read-modify-write happens in L2, no thread waits on a load.
            uint8_t *boxes = smem + e * (BLOCK_N / 32) * 4096;      // operand stages are free now
            for (int c = 0; c < BLOCK_N / 32; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(32 * e) << 16) + (uint32_t)(32 * c), v);
                uint8_t *box = boxes + c * 4096;
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    float f[4];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const int col = n0 + 32 * c + 4 * j + q;
                        f[q] = __uint_as_float(v[4 * j + q]);
                        if (p.bias && blockIdx.z == 0 && col < p.N) f[q] += p.bias[col];
                        if (p.relu) f[q] = fmaxf(f[q], 0.f);
                    }
                    *reinterpret_cast<float4 *>(box + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_float4(f[0], f[1], f[2], f[3]);
                    if (p.Cb) {   // bf16 copy for the next tensor-core GEMM, straight from the registers
                        const int row = m0 + 32 * e + lane, col = n0 + 32 * c + 4 * j;
                        if (row < p.M && col < p.N) {   // N % 4 == 0 (launcher)
                            const __nv_bfloat162 lo = __floats2bfloat162_rn(f[0], f[1]), hi = __floats2bfloat162_rn(f[2], f[3]);
                            uint2 pk;
                            pk.x = *reinterpret_cast<const uint32_t *>(&lo);
                            pk.y = *reinterpret_cast<const uint32_t *>(&hi);
                            *reinterpret_cast<uint2 *>(p.Cb + (size_t)row * p.ldcb + col) = pk;
                        }
                    }
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                for (int c = 0; c < BLOCK_N / 32; ++c) {
                    const uint32_t src = smem_u32(boxes + c * 4096);
                    const int c0 = n0 + 32 * c, c1 = m0 + 32 * e;
                    if (c0 >= p.N || c1 >= p.M) continue;
                    if (p.tma_epi == 2)
                        asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
                                     ::"l"(&tmC), "r"(c0), "r"(c1), "r"(src) : "memory");
                    else
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];"
                                     ::"l"(&tmC), "r"(c0), "r"(c1), "r"(src) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncwarp();
        } else {
        float *tile = reinterpret_cast<float *>(smem) + e * (32 * 33);   // operand stages are free now
        for (int c = 0; c < BLOCK_N / 32; ++c) {
            uint32_t v[32];
            tmem_ld32(tmem_base + ((uint32_t)(32 * e) << 16) + (uint32_t)(32 * c), v);
#pragma unroll
            for (int j = 0; j < 32; ++j) tile[lane * 33 + j] = __uint_as_float(v[j]);
            __syncwarp();
            const int col = n0 + 32 * c + lane;
            const float bv = (p.bias && col < p.N && blockIdx.z == 0) ? p.bias[col] : 0.f;
#pragma unroll 4
            for (int r = 0; r < 32; ++r) {
                const int row = m0 + 32 * e + r;
                if (row < p.M && col < p.N) {
                    float x = tile[r * 33 + lane] + bv;
                    if (gridDim.z > 1) {   // C already holds beta * C (launcher)
                        atomicAdd(p.C + (size_t)row * p.ldc + col, x);
                        continue;
                    }
                    if (p.beta != 0.f) x = fmaf(p.beta, p.C[(size_t)row * p.ldc + col], x);
                    if (p.relu) x = fmaxf(x, 0.f);
                    if (p.C) p.C[(size_t)row * p.ldc + col] = x;
                    if (p.Cb) p.Cb[(size_t)row * p.ldcb + col] = __float2bfloat16(x);
                }
            }
            __syncwarp();
        }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }

    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
    }
}

// ---------------------------------------------------------------------------
// Persistent variant for many-tile products (the decoder's last layer: 4096 x 30000 x 1024).
// One CTA per SM walks 128 x 256 tiles (half the operand traffic per flop of a 128 x 128 tile);
// the TMA producer and the MMA issuer run ahead across tile boundaries (the 4-stage ring never
// drains), and the accumulator is double-buffered in TMEM (2 x 256 columns) so the epilogue warps
// drain tile i through TMA tile stores while tile i + 1 is being multiplied.
// ---------------------------------------------------------------------------
constexpr int P_BLOCK_N = 256, P_STAGES = 4;
constexpr int P_B_BYTES = P_BLOCK_N * BLOCK_K * 2, P_STAGE_BYTES = A_BYTES + P_B_BYTES;
constexpr int P_BOX_BYTES = 4 * 2 * 4096;        // 4 epilogue warps x 2 staging boxes of [32 x 32] fp32
constexpr int P_SMEM_BYTES = P_STAGES * P_STAGE_BYTES + P_BOX_BYTES + 1024 + 256;

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(THREADS, 1)
gemm_tc_persist_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                       const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmAlo,
                       const __grid_constant__ CUtensorMap tmBlo, params_t p)
{
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint8_t *boxes_all = smem + P_STAGES * P_STAGE_BYTES;
    uint64_t *bars = reinterpret_cast<uint64_t *>(boxes_all + P_BOX_BYTES);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * P_STAGES + 4);
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + P_STAGES);
    const uint32_t tfull0 = smem_u32(bars + 2 * P_STAGES), tempty0 = tfull0 + 16;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tiles_m = (p.M + BLOCK_M - 1) / BLOCK_M, tiles_n = (p.N + P_BLOCK_N - 1) / P_BLOCK_N;
    const int num_tiles = tiles_m * tiles_n;
    const int base_kb = (p.K + BLOCK_K - 1) / BLOCK_K;
    const int num_kb = base_kb * p.segs;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmC) : "memory");
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < P_STAGES; ++s) {
            mbar_init(full0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(tfull0 + 8 * b, 1);
            mbar_init(tempty0 + 8 * b, 4);            // one arrival per epilogue warp
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    pdl_sync();   // (descriptor prefetch, barrier init and the TMEM allocation above overlap the previous kernel's tail)

    if (warp == 0) {
        // ================= TMA producer (runs ahead across tiles) =================
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x) {
                // concurrent CTAs share the operand that is walked slowly; the LARGER operand must be the one reused
                // back to back, or it is streamed from HBM once per tile of the other dimension
                const int m0 = (p.n_fast ? t / tiles_n : t % tiles_m) * BLOCK_M, n0 = (p.n_fast ? t % tiles_n : t / tiles_m) * P_BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const uint32_t s = it % P_STAGES, ph = (it / P_STAGES) & 1;
                    mbar_wait(empty0 + 8 * s, ph ^ 1);
                    const uint32_t sa = smem_u32(smem + s * P_STAGE_BYTES), sb = sa + A_BYTES;
                    const uint32_t fb = full0 + 8 * s;
                    mbar_expect_tx(fb, P_STAGE_BYTES);
                    const int seg = p.segs == 3 ? kb / base_kb : 2;       // (see gemm_tc_kernel)
                    const int k0 = (p.segs == 3 ? kb - seg * base_kb : kb) * BLOCK_K;
                    const CUtensorMap *mA = seg == 1 ? &tmAlo : &tmA, *mB = seg == 0 ? &tmBlo : &tmB;
                    if (A_MN) {
                        tma_load_2d(sa, mA, fb, m0, k0);
                        tma_load_2d(sa + A_BYTES / 2, mA, fb, m0 + 64, k0);
                    } else {
                        tma_load_2d(sa, mA, fb, k0, m0);
                    }
                    if (B_MN) {
#pragma unroll
                        for (int q = 0; q < P_BLOCK_N / 64; ++q) tma_load_2d(sb + q * 8192, mB, fb, n0 + 64 * q, k0);
                    } else {
                        tma_load_2d(sb, mB, fb, k0, n0);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(A_MN, B_MN, BLOCK_M, P_BLOCK_N);
            uint32_t it = 0, lt = 0;
            for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++lt) {
                const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
                mbar_wait(tempty0 + 8 * acc, aph ^ 1);              // epilogue has drained this accumulator
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d_tmem = tmem_base + acc * P_BLOCK_N;
                for (int kb = 0; kb < num_kb; ++kb, ++it) {
                    const uint32_t s = it % P_STAGES, ph = (it / P_STAGES) & 1;
                    mbar_wait(full0 + 8 * s, ph);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = smem_u32(smem + s * P_STAGE_BYTES), sb = sa + A_BYTES;
#pragma unroll
                    for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
                        const uint64_t ad = A_MN ? make_desc(sa + k * 2048, A_BYTES / 2, 1024)
                                                 : make_desc(sa + k * 32, 16, 1024);
                        const uint64_t bd = B_MN ? make_desc(sb + k * 2048, 8192, 1024)
                                                 : make_desc(sb + k * 32, 16, 1024);
                        umma_bf16(d_tmem, ad, bd, idesc, (kb | k) != 0);
                    }
                    umma_commit(empty0 + 8 * s);
                }
                umma_commit(tfull0 + 8 * acc);
            }
        }
    } else if (warp >= 4) {
        // ================= epilogue (overlaps the next tile's MMAs) =================
        const int e = warp - 4;
        uint8_t *boxes = boxes_all + e * 2 * 4096;
        uint32_t lt = 0, nstore = 0;
        float ss = 0.f;
        for (int t = blockIdx.x; t < num_tiles; t += gridDim.x, ++lt) {
            const int m0 = (p.n_fast ? t / tiles_n : t % tiles_m) * BLOCK_M, n0 = (p.n_fast ? t % tiles_n : t / tiles_m) * P_BLOCK_N;
            const uint32_t acc = lt & 1, aph = (lt >> 1) & 1;
            mbar_wait(tfull0 + 8 * acc, aph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            float rmax = -INFINITY, rsum = 0.f;      // softmax partials of this lane's row over the tile's columns (log2 domain)
            for (int c = 0; c < P_BLOCK_N / 32; ++c) {
                uint32_t v[32];
                float bl = 0.f;
                if (p.bias) {
                    const int col = n0 + 32 * c + lane;
                    if (col < p.N) bl = __ldg(p.bias + col);
                }
                tmem_ld32(tmem_base + ((uint32_t)(32 * e) << 16) + acc * P_BLOCK_N + (uint32_t)(32 * c), v);
                if (c == P_BLOCK_N / 32 - 1) {   // accumulator fully read: hand it back to the MMA warp
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(tempty0 + 8 * acc);
                }
                uint8_t *box = boxes + (nstore & 1) * 4096;
                if (nstore >= 2) {               // the store issued two chunks ago has finished reading this box
                    if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    __syncwarp();
                }
                // The four epilogue warps sit alone on their sub-partitions: every instruction here is on the critical
                // path of the tile, so no per-element branches or loads.  The chunk's 32 bias values come in with ONE
                // coalesced load (issued before the TMEM read above) and are handed round by shuffles.
                if (p.bias) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) + __shfl_sync(0xffffffffu, bl, i));
                }
                if (p.relu) {
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(fmaxf(__uint_as_float(v[i]), 0.f));
                }
                if (p.row_stats) {
                    const int ncol = p.N - (n0 + 32 * c);             // columns of this chunk inside the matrix
                    if (ncol < 32) {                                  // edge tile: -inf drops a column from max and sum
                        // (the TMA store clips such a column, the -inf never reaches memory)
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (i >= ncol) v[i] = __float_as_uint(-INFINITY);
                    }
                    float cm = __uint_as_float(v[0]);
#pragma unroll
                    for (int i = 1; i < 32; ++i) cm = fmaxf(cm, __uint_as_float(v[i]));
                    if (cm > -INFINITY) {                             // online merge into the tile's running (max, sum), log2 domain
                        constexpr float kL2E = 1.4426950408889634f;
                        const float nm = fmaxf(rmax, cm * kL2E);
                        float a0 = rmax > -INFINITY ? rsum * exp2f(rmax - nm) : 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
                        for (int i = 0; i < 32; i += 4) {
                            float t0, t1, t2, t3;
                            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t0) : "f"(fmaf(__uint_as_float(v[i]), kL2E, -nm)));
                            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t1) : "f"(fmaf(__uint_as_float(v[i + 1]), kL2E, -nm)));
                            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t2) : "f"(fmaf(__uint_as_float(v[i + 2]), kL2E, -nm)));
                            asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t3) : "f"(fmaf(__uint_as_float(v[i + 3]), kL2E, -nm)));
                            a0 += t0; a1 += t1; a2 += t2; a3 += t3;
                        }
                        rsum = (a0 + a1) + (a2 + a3);
                        rmax = nm;
                    }
                }
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const float f0 = __uint_as_float(v[4 * j]), f1 = __uint_as_float(v[4 * j + 1]), f2 = __uint_as_float(v[4 * j + 2]),
                                f3 = __uint_as_float(v[4 * j + 3]);
                    if (p.sumsq) ss += f0 * f0 + f1 * f1 + f2 * f2 + f3 * f3;   // out-of-range parts of edge tiles are zero (no bias: launcher)
                    *reinterpret_cast<float4 *>(box + lane * 128 + ((j ^ (lane & 7)) << 4)) = make_float4(f0, f1, f2, f3);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    const int c0 = n0 + 32 * c, c1 = m0 + 32 * e;
                    if (c0 < p.N && c1 < p.M) {
                        if (p.tma_epi == 2)
                            asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%1, %2}], [%3];"
                                         ::"l"(&tmC), "r"(c0), "r"(c1), "r"(smem_u32(box)) : "memory");
                        else
                            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];"
                                         ::"l"(&tmC), "r"(c0), "r"(c1), "r"(smem_u32(box)) : "memory");
                    }
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");   // (possibly empty) group: keeps the count in step
                }
                ++nstore;
            }
            if (p.row_stats) {
                const int row = m0 + 32 * e + lane;
                if (row < p.M) p.row_stats[(size_t)(n0 / P_BLOCK_N) * p.M + row] = make_float2(rmax, rsum);   // coalesced over rows
            }
        }
        if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
        __syncwarp();
        if (p.sumsq) {
            ss = warp_sum(ss);
            if (lane == 0) atomicAdd(p.sumsq, (double)ss);
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }

    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512) : "memory");
    }
}

} // namespace tc

// ---------------------------------------------------------------------------
// host side: tensor maps + launch
// ---------------------------------------------------------------------------
namespace {

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

encode_tiled_fn
get_encode()
{
    static const encode_tiled_fn fn = [] {   // thread-safe one-time lookup
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            return (encode_tiled_fn)p;
        return (encode_tiled_fn) nullptr;
    }();
    return fn;
}

// bf16 matrix stored row-major [outer x inner] with leading dimension ld (elements)
int
make_map(CUtensorMap *map, const void *base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
         uint32_t box_outer, bool f32 = false)
{
    encode_tiled_fn enc = get_encode();
    MMV_CHECK(enc, "gemm_tc: cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t dims[2] = { inner, outer };
    cuuint64_t strides[1] = { ld * (f32 ? 4 : 2) };
    cuuint32_t box[2] = { box_inner, box_outer };
    cuuint32_t estr[2] = { 1, 1 };
    CUresult r = enc(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    MMV_CHECK(r == CUDA_SUCCESS, "gemm_tc: cuTensorMapEncodeTiled failed (%d): inner=%lu outer=%lu ld=%lu", (int)r,
              (unsigned long)inner, (unsigned long)outer, (unsigned long)ld);
    return 0;
}

} // namespace

int
make_tma_map_bf16(void *map, const void *base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                  uint32_t box_outer)
{
    return make_map(reinterpret_cast<CUtensorMap *>(map), base, inner, outer, ld, box_inner, box_outer);
}

int
make_tma_map_f32(void *map, const void *base, uint64_t inner, uint64_t outer, uint64_t ld, uint32_t box_inner,
                 uint32_t box_outer)
{
    return make_map(reinterpret_cast<CUtensorMap *>(map), base, inner, outer, ld, box_inner, box_outer, true);
}

bool
gemm_tc_supported(const gemm_tc_args_t &g)
{
    auto al16 = [](const void *p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; };
    return g.M > 0 && g.N > 0 && g.K > 0 && g.lda % 8 == 0 && g.ldb % 8 == 0 && al16(g.A) && al16(g.B);
}

int
launch_gemm_tc(const gemm_tc_args_t &g, cudaStream_t stream)
{
    using namespace tc;
    MMV_CHECK(gemm_tc_supported(g), "gemm_tc: operands must be 16-byte aligned with leading dimensions %% 8 == 0");
    CUtensorMap ta, tb;
    if (g.a_mn) MMV_TRY(make_map(&ta, g.A, g.M, g.K, g.lda, 64, BLOCK_K));
    else MMV_TRY(make_map(&ta, g.A, g.K, g.M, g.lda, BLOCK_K, BLOCK_M));
    if (g.b_mn) MMV_TRY(make_map(&tb, g.B, g.N, g.K, g.ldb, 64, BLOCK_K));
    else MMV_TRY(make_map(&tb, g.B, g.K, g.N, g.ldb, BLOCK_K, BLOCK_N));
    // fp32-tolerance mode: the residual copies share the layout of their hi halves
    MMV_CHECK((g.A_lo == nullptr) == (g.B_lo == nullptr), "gemm_tc: the three-product mode needs both residual operands");
    auto al16 = [](const void *q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    MMV_CHECK(!g.A_lo || (al16(g.A_lo) && al16(g.B_lo)), "gemm_tc: residual operands must be 16-byte aligned");
    CUtensorMap tal = ta, tbl = tb;
    if (g.A_lo) {
        if (g.a_mn) MMV_TRY(make_map(&tal, g.A_lo, g.M, g.K, g.lda, 64, BLOCK_K));
        else MMV_TRY(make_map(&tal, g.A_lo, g.K, g.M, g.lda, BLOCK_K, BLOCK_M));
        if (g.b_mn) MMV_TRY(make_map(&tbl, g.B_lo, g.N, g.K, g.ldb, 64, BLOCK_K));
        else MMV_TRY(make_map(&tbl, g.B_lo, g.K, g.N, g.ldb, BLOCK_K, BLOCK_N));
    }
    params_t p;
    p.segs = g.A_lo ? 3 : 1;
    p.C = g.C; p.Cb = reinterpret_cast<__nv_bfloat16 *>(g.Cb); p.bias = g.bias;
    p.M = g.M; p.N = g.N; p.K = g.K; p.ldc = g.ldc; p.ldcb = g.ldcb; p.beta = g.beta; p.relu = g.relu;
    p.sumsq = nullptr;
    p.row_stats = nullptr;
    p.n_fast = g.M > g.N ? 1 : 0;
    if (g.sumsq_done) *g.sumsq_done = 0;
    if (g.row_stats_tiles) *g.row_stats_tiles = 0;
    dim3 grid(ceil_div(g.N, BLOCK_N), ceil_div(g.M, BLOCK_M));
    const int tiles = (int)(grid.x * grid.y), total_kb = ceil_div(g.K, BLOCK_K) * (g.A_lo ? 3 : 1);
    if (tiles <= 64 && total_kb >= 16 && !g.relu && !g.Cb && g.C && (g.beta == 1.f || (g.beta == 0.f && g.ldc == g.N))) {
        int splits = total_kb / 4;                 // at least 4 k-blocks per CTA
        if (splits > 256 / tiles) splits = 256 / tiles;
        if (splits > 1) {
            const int per = ceil_div(total_kb, splits);
            splits = ceil_div(total_kb, per);      // every split owns at least one k-block
        }
        if (splits > 1) {
            grid.z = splits;
            if (g.beta == 0.f) MMV_CUDA(cudaMemsetAsync(g.C, 0, sizeof(float) * (size_t)g.M * g.N, stream));
        }
    }
    // epilogue through TMA when C can be described by a tensor map and no bf16 copy is wanted
    CUtensorMap tc_map = ta;
    p.tma_epi = 0;
    const bool cb_ok = !g.Cb || (g.beta == 0.f && grid.z == 1 && g.N % 4 == 0 && g.ldcb % 4 == 0 &&
                                 (reinterpret_cast<uintptr_t>(g.Cb) & 7) == 0);
    if (g.C && cb_ok && g.ldc % 4 == 0 && (reinterpret_cast<uintptr_t>(g.C) & 15) == 0 &&
        (g.beta == 0.f || g.beta == 1.f) && !(g.relu && (g.beta == 1.f || grid.z > 1))) {
        MMV_TRY(make_map(&tc_map, g.C, g.N, g.M, g.ldc, 32, 32, true));
        p.tma_epi = (g.beta == 1.f || grid.z > 1) ? 2 : 1;
    }
    // many-tile products: persistent 128 x 256 kernel (operand ring never drains, epilogue overlapped)
    int num_sms = 0;
    MMV_TRY(dev_num_sms(&num_sms));
    static const int persist_mode = [] { const char *e = getenv("MMV_GEMM_PERSIST"); return e ? atoi(e) : 1; }();
    static std::atomic<int> log_left{ [] { const char *e = getenv("MMV_GEMM_LOG"); return e ? atoi(e) : 0; }() };   // MMV_GEMM_LOG=<n>: print the first n launches' shapes
    if (log_left.load(std::memory_order_relaxed) > 0 && log_left.fetch_sub(1) > 0) {
        fprintf(stderr, "[gemm_tc] M=%d N=%d K=%d a_mn=%d b_mn=%d lda=%d ldb=%d ldc=%d beta=%g bias=%d relu=%d Cb=%d epi=%d grid=(%u,%u,%u) A=%p B=%p C=%p\n",
                g.M, g.N, g.K, g.a_mn, g.b_mn, g.lda, g.ldb, g.ldc, g.beta, g.bias != nullptr, g.relu, g.Cb != nullptr, p.tma_epi,
                grid.x, grid.y, grid.z, g.A, g.B, (void *)g.C);
    }
    const int ptiles = ceil_div(g.M, BLOCK_M) * ceil_div(g.N, P_BLOCK_N);
    if (persist_mode > 0 && p.tma_epi && !g.Cb && grid.z == 1 && ptiles >= 2 * num_sms) {
        if (!g.b_mn) MMV_TRY(make_map(&tb, g.B, g.K, g.N, g.ldb, BLOCK_K, P_BLOCK_N));
        if (!g.b_mn) { tbl = tb; if (g.B_lo) MMV_TRY(make_map(&tbl, g.B_lo, g.K, g.N, g.ldb, BLOCK_K, P_BLOCK_N)); }
        if (g.sumsq && p.tma_epi == 1 && !g.bias) {
            p.sumsq = g.sumsq;
            if (g.sumsq_done) *g.sumsq_done = 1;
        }
        if (g.row_stats && g.row_stats_tiles && p.tma_epi == 1 && !g.relu) {
            p.row_stats = reinterpret_cast<float2 *>(g.row_stats);
            *g.row_stats_tiles = ceil_div(g.N, P_BLOCK_N);
        }
        void (*pk)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, params_t) =
            g.a_mn ? (g.b_mn ? gemm_tc_persist_kernel<true, true> : gemm_tc_persist_kernel<true, false>)
                   : (g.b_mn ? gemm_tc_persist_kernel<false, true> : gemm_tc_persist_kernel<false, false>);
        MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(pk), P_SMEM_BYTES));
        MMV_CUDA(launch_k(pk, dim3(std::min(ptiles, num_sms)), dim3(THREADS), P_SMEM_BYTES, stream, ta, tb, tc_map, tal, tbl, p));
        ::mmv::count_launch();
        return 0;
    }
    const bool deep = (int)(grid.x * grid.y * grid.z) <= num_sms && ceil_div(total_kb, (int)grid.z) >= 6;
    void (*kern)(const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, const CUtensorMap, params_t);
    if (deep)
        kern = g.a_mn ? (g.b_mn ? gemm_tc_kernel<true, true, 6> : gemm_tc_kernel<true, false, 6>)
                      : (g.b_mn ? gemm_tc_kernel<false, true, 6> : gemm_tc_kernel<false, false, 6>);
    else
        kern = g.a_mn ? (g.b_mn ? gemm_tc_kernel<true, true, 3> : gemm_tc_kernel<true, false, 3>)
                      : (g.b_mn ? gemm_tc_kernel<false, true, 3> : gemm_tc_kernel<false, false, 3>);
    const int smem = smem_bytes(deep ? 6 : 3);
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(kern), smem));
    MMV_CUDA(launch_k(kern, grid, dim3(THREADS), smem, stream, ta, tb, tc_map, tal, tbl, p));
    ::mmv::count_launch();
    return 0;
}

// fp32 -> (hi, lo) bf16 pair, x = hi + lo up to 2^-17 relative: the operands of the error-compensated three-product GEMM
// of the fp32-tolerance tensor-core mode ("dtype" 2: A B ~ A_hi B_lo + A_lo B_hi + A_hi B_hi, fp32 accumulation)
__global__ void
split_bf16_kernel(const float *x, __nv_bfloat16 *hi, __nv_bfloat16 *lo, size_t n)
{
    pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
        const float v = x[i];
        const __nv_bfloat16 h = __float2bfloat16(v);
        hi[i] = h;
        lo[i] = __float2bfloat16(v - __bfloat162float(h));
    }
}

int
launch_split_bf16(const float *x, void *hi, void *lo, size_t n, cudaStream_t st)
{
    if (n == 0) return 0;
    MMV_CUDA(launch_k(split_bf16_kernel, dim3((unsigned)min_sz((n + 255) / 256, 8192)), dim3(256), 0, st, x,
                      reinterpret_cast<__nv_bfloat16 *>(hi), reinterpret_cast<__nv_bfloat16 *>(lo), n));
    ::mmv::count_launch();
    return 0;
}

// fp32 -> bf16 (activations entering a tensor-core GEMM)
__global__ void
to_bf16_kernel(const float *x, __nv_bfloat16 *y, size_t n)
{
    pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        y[i] = __float2bfloat16(x[i]);
}

int
launch_to_bf16(const float *x, void *y, size_t n, cudaStream_t st)
{
    if (n == 0) return 0;
    MMV_CUDA(launch_k(to_bf16_kernel, dim3((unsigned)min_sz((n + 255) / 256, 4096)), dim3(256), 0, st, x, reinterpret_cast<__nv_bfloat16 *>(y), n));
    ::mmv::count_launch();
    return 0;
}

} // namespace mmv
