// K1, tensor-core path — sparse x dense first encoder layer and its weight gradient on tcgen05.
//
// Replaces (same math as sparse_encoder.cu, which stays as the fp32 parity path):
//   encode_mu first Linear            nb.hh:406-411   h1 = ((log1p x - x_mean) / t) W1^T + b1
//   encode first Angular              vmf.hh:253-260, angular.hh:34-42
//   nu_enc(x), depth(x) on raw counts nb.hh:448,498
//   their autograd backward (dW1, d x_mean, d ln_x_sd, dW_nu, dw_depth)
//
// At 5 % density a gather formulation moves nnz x h1 weight elements through L2 (12.6 GB per
// minibatch at cfg 4) and runs on the FMA pipe; the dense product is 20x more flops but the
// tensor pipe is >20x faster and the weight tile is read once per 128 cells.  So the minibatch is
// re-bucketed once per step into a tiled COO "panel" (tiles of 64 cells x 64 genes, packed 32-bit
// entries), and the GEMM kernels DENSIFY one [128 x 64] bf16 operand tile per pipeline stage
// directly in shared memory (SWIZZLE_128B K-major layout, written with st.shared, published to the
// async proxy with fence.proxy.async) while TMA streams the dense operand; tcgen05.mma accumulates
// in TMEM.  The same panel serves forward (rows = cells, K = genes) and backward (rows = genes,
// K = cells) — only the scatter coordinates swap.
#include "common.cuh"
#include "kernels.hh"
#include "tc_ptx.cuh"

#include <cuda.h>
#include <cuda_bf16.h>

#include <algorithm>
#include <cstdlib>

namespace mmv {

namespace {

// ===========================================================================
// panel build
// ===========================================================================
#ifndef MMV_PANEL_PREFETCH
#define MMV_PANEL_PREFETCH 1
#endif
constexpr int kPanelThreads = 512;
constexpr int kUn = 4;

// One CTA per (cell tile T, gene split s): the tile's 64 rows are cut at the split's chunk range
// [k_lo, k_hi) by bisection of the sorted gene ids, so every CTA owns its buckets outright and the
// grid covers the machine even at 64 cell tiles.  Unsorted rows: the whole run is scanned and filtered.
__global__ void __launch_bounds__(kPanelThreads)
panel_build_kernel(panel_args_t p, int nK, int nS)
{
    extern __shared__ int s_int[];
    __shared__ int s_scan[kPanelThreads];
    __shared__ int64_t s_qlo[64], s_qhi[64];
    __shared__ int s_before;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int NW = kPanelThreads / 32;
    pdl_sync();
    const int T = blockIdx.x, sp = blockIdx.y, r0 = T * 64, nrows = min(64, p.B - r0);
    const int k_lo = (int)((int64_t)nK * sp / nS), k_hi = (int)((int64_t)nK * (sp + 1) / nS), nKl = k_hi - k_lo;
    int *s_off = s_int;               // [nKl + 1] exclusive offsets of this CTA's buckets
    int *s_cur = s_int + nKl + 1;     // [nKl] fill cursors
    const bool sorted = p.unsorted != nullptr && *p.unsorted == 0;
    const bool filter = !sorted && nS > 1;

    for (int k = tid; k < nKl; k += kPanelThreads) s_cur[k] = 0;
    if (tid == 0) s_before = 0;
    // minibatch-local CSR offset of the tile's first row = run lengths of all earlier rows.  Every CTA adds them up
    // itself (<= B gathered loads from L2, 8 per thread at B = 4096): cheaper than a one-CTA scan kernel in front
    int prefix = 0;
    for (int r = tid; r < r0; r += kPanelThreads) {
        const int64_t c = p.cells ? (int64_t)p.cells[r] : (int64_t)r;
        prefix += (int)(p.indptr[c + 1] - p.indptr[c]);
    }
    __syncthreads();
    if (tid < 128) {                  // per row: where the split's gene range starts / ends in the run
        const int r = tid >> 1, which = tid & 1;
        if (r < nrows) {
            const int64_t c = p.cells ? (int64_t)p.cells[r0 + r] : (int64_t)(r0 + r);
            const int64_t beg = p.indptr[c], end = p.indptr[c + 1];
            const int kk = which ? k_hi : k_lo;
            int64_t v = which ? end : beg;
            if (sorted && kk > 0 && kk < nK) {
                const int key = kk * 64;
                int64_t lo = beg, hi = end;
                while (lo < hi) {
                    const int64_t mid = (lo + hi) >> 1;
                    if (p.indices[mid] < key) lo = mid + 1; else hi = mid;
                }
                v = lo;
            }
            if (which) s_qhi[r] = v;
            else {
                s_qlo[r] = v;
                if (sorted && v > beg) atomicAdd(&s_before, (int)(v - beg));   // entries of earlier splits
            }
        }
    }
    __syncthreads();
#if MMV_PANEL_PREFETCH
    // pull every row's slice of the CSR arrays into L2 before the two passes walk them: with a resident shard the rows of
    // a minibatch are scattered over gigabytes, each run a cold DRAM + TLB miss behind only kUn loads in flight
    for (int r = warp; r < nrows; r += NW) {
        const int64_t beg = s_qlo[r], end = s_qhi[r];
        for (int64_t q = beg + lane * 32; q < end; q += 32 * 32) {
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.indices + q));
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p.vals + q));
        }
    }
#endif
    // ---- pass 1: bucket sizes
    int before_unsorted = 0;          // unsorted rows: count the entries that belong to earlier splits
    for (int r = warp; r < nrows; r += NW) {
        const int64_t beg = s_qlo[r], end = s_qhi[r];
        for (int64_t q0 = beg; q0 < end; q0 += 32 * kUn) {   // kUn independent loads in flight per lane
            int j[kUn];
#pragma unroll
            for (int u = 0; u < kUn; ++u) {
                const int64_t q = q0 + lane + 32 * u;
                j[u] = q < end ? p.indices[q] : -1;
            }
#pragma unroll
            for (int u = 0; u < kUn; ++u) {
                const int k = (j[u] >> 6) - k_lo;
                if (j[u] >= 0 && k >= 0 && k < nKl) atomicAdd(&s_cur[k], 1);
                else if (filter && j[u] >= 0 && k < 0) ++before_unsorted;
            }
        }
    }
    if (filter) {
        before_unsorted = warp_sum(before_unsorted);
        if (lane == 0 && before_unsorted) atomicAdd(&s_before, before_unsorted);
    }
    __syncthreads();
    {   // block sum of the row prefix (s_scan is free until the scan below)
        prefix = warp_sum(prefix);
        if (lane == 0) s_scan[warp] = prefix;
        __syncthreads();
        prefix = 0;
#pragma unroll
        for (int w = 0; w < NW; ++w) prefix += s_scan[w];
        __syncthreads();
    }
    const int base = prefix + s_before;
    // ---- exclusive scan over the chunks (each thread owns a contiguous span)
    const int per = (nKl + kPanelThreads - 1) / kPanelThreads;
    const int k0 = min(nKl, tid * per), k1 = min(nKl, k0 + per);
    int sum = 0;
    for (int k = k0; k < k1; ++k) sum += s_cur[k];
    s_scan[tid] = sum;
    __syncthreads();
    for (int off = 1; off < kPanelThreads; off <<= 1) {
        const int v = tid >= off ? s_scan[tid - off] : 0;
        __syncthreads();
        s_scan[tid] += v;
        __syncthreads();
    }
    int run = s_scan[tid] - sum;
    for (int k = k0; k < k1; ++k) {
        const int n = s_cur[k];
        s_off[k] = run;
        run += n;
    }
    if (tid == kPanelThreads - 1) s_off[nKl] = s_scan[kPanelThreads - 1];
    __syncthreads();
    int *bp = p.bucket_ptr + (size_t)T * (nK + 1) + k_lo;
    for (int k = tid; k <= nKl; k += kPanelThreads) {
        if (k < nKl || sp == nS - 1) bp[k] = base + s_off[k];   // bp[k_hi] belongs to the next split
        if (k < nKl) s_cur[k] = s_off[k];
    }
    __syncthreads();
    // ---- pass 2: emit packed entries; partial row scalars on the way
    for (int r = warp; r < nrows; r += NW) {
        const int64_t beg = s_qlo[r], end = s_qhi[r];
        float rs = 1.f;
        if (p.angular) {   // vmf.hh:255: row L2 norm of log1p(x) over the WHOLE row, F::normalize eps 1e-12
            const int64_t c = p.cells ? (int64_t)p.cells[r0 + r] : (int64_t)(r0 + r);
            float sq = 0.f;
            for (int64_t q = p.indptr[c] + lane, e = p.indptr[c + 1]; q < e; q += 32) {
                const float l = log1pf(p.vals[q]);
                sq = fmaf(l, l, sq);
            }
            sq = warp_sum(sq);
            const float nrm = fmaxf(sqrtf(sq), 1e-12f);
            rs = 1.f / nrm;
            if (lane == 0 && sp == 0 && p.rownorm) p.rownorm[r0 + r] = nrm;
        }
        float depth = 0.f, hn0 = 0.f;
        const bool h_inline = p.H == 1;
        for (int64_t q0 = beg; q0 < end; q0 += 32 * kUn) {
            int j[kUn];
            float x[kUn], it[kUn], wd[kUn], wn[kUn];
#pragma unroll
            for (int u = 0; u < kUn; ++u) {
                const int64_t q = q0 + lane + 32 * u;
                j[u] = q < end ? p.indices[q] : -1;
                x[u] = q < end ? p.vals[q] : 0.f;
                const int k = (j[u] >> 6) - k_lo;
                if (k < 0 || k >= nKl) j[u] = -1;      // (unsorted rows) another split's entry
            }
#pragma unroll
            for (int u = 0; u < kUn; ++u) {
                // (tried, r02g A/B: this CTA's gene range of the three arrays staged in shared memory — no gain, 0.106 vs
                // 0.103 ms: the long-scoreboard stalls are the entry loads, not these L2-resident gathers)
                const int jj = max(j[u], 0);
                it[u] = p.inv_t[jj];
                wd[u] = p.wdepth ? p.wdepth[jj] : 0.f;
                wn[u] = h_inline ? p.WnuT[jj] : 0.f;
            }
#pragma unroll
            for (int u = 0; u < kUn; ++u) {
                if (j[u] < 0) continue;
                const float v = log1pf(x[u]) * rs * it[u];
                const int pos = base + atomicAdd(&s_cur[(j[u] >> 6) - k_lo], 1);
                const uint32_t bits = (uint32_t)__bfloat16_as_ushort(__float2bfloat16(v));
                p.ent[pos] = ((uint32_t)r << 26) | ((uint32_t)(j[u] & 63) << 20) | bits;
                if (p.ent_lo) {
                    const float lo = v - __bfloat162float(__ushort_as_bfloat16((unsigned short)bits));
                    p.ent_lo[pos] = ((uint32_t)r << 26) | ((uint32_t)(j[u] & 63) << 20) |
                                    (uint32_t)__bfloat16_as_ushort(__float2bfloat16(lo));
                }
                p.xraw[pos] = x[u];
                depth = fmaf(x[u], wd[u], depth);
                hn0 = fmaf(x[u], wn[u], hn0);
            }
        }
        // (Hnu / dpre are zero-filled by the launcher; the first split's CTA brings the bias)
        if (h_inline) {   // nb.hh:448 (raw counts), default H = 1
            hn0 = warp_sum(hn0);
            if (lane == 0) atomicAdd(p.Hnu + r0 + r, hn0 + (sp == 0 ? p.bnu[0] : 0.f));
        }
        if (p.wdepth) {   // nb.hh:498 (raw counts)
            depth = warp_sum(depth);
            if (lane == 0) atomicAdd(p.dpre + r0 + r, depth + (sp == 0 ? p.bdepth[0] : 0.f));
        }
        for (int h = 0; h < p.H && !h_inline; ++h) {   // nb.hh:448 (raw counts), general H
            float s = 0.f;
            for (int64_t q = beg + lane; q < end; q += 32) {
                const int j = p.indices[q], k = (j >> 6) - k_lo;
                if (k >= 0 && k < nKl) s = fmaf(p.vals[q], p.WnuT[(size_t)j * p.H + h], s);
            }
            s = warp_sum(s);
            if (lane == 0) atomicAdd(p.Hnu + (size_t)(r0 + r) * p.H + h, s + (sp == 0 ? p.bnu[h] : 0.f));
        }
    }
}

} // namespace

int
launch_panel_build(const panel_args_t &p, cudaStream_t st)
{
    const int nT = panel_tiles(p.B), nK = panel_chunks(p.D);
    // gene splits per cell tile: as many as fit ONE wave of two 512-thread CTAs per SM (64 registers per thread).  Rounding
    // up instead (5 splits x 64 tiles = 320 CTAs on 296 slots at B = 4096) leaves a second, nearly empty wave behind.
    const int slots = 2 * std::max(p.num_sms, 1);
    int nS = nT <= slots ? slots / nT : 1;
    if (const char *e = getenv("MMV_PANEL_SPLITS")) nS = atoi(e);
    nS = std::max(1, std::min(std::min(nS, 8), nK));
    const int nKl = (nK + nS - 1) / nS + 1;
    const size_t smem = sizeof(int) * (2 * (size_t)nKl + 2);
    MMV_CHECK(smem <= 200 * 1024, "panel_build: %d genes exceed the shared-memory bucket table", p.D);
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(panel_build_kernel), 200 * 1024));
    if (p.wdepth && p.dpre) MMV_CUDA(cudaMemsetAsync(p.dpre, 0, sizeof(float) * p.B, st));
    if (p.H > 0 && p.Hnu) MMV_CUDA(cudaMemsetAsync(p.Hnu, 0, sizeof(float) * (size_t)p.B * p.H, st));
    MMV_CUDA(launch_k(panel_build_kernel, dim3(nT, nS), dim3(kPanelThreads), smem, st, p, nK, nS));
    ::mmv::count_launch();
    return 0;
}


// ===========================================================================
// densify-in-smem tcgen05 GEMM:  C[128 x BN] += A_panel[128 x 64] * Dense[64 x BN]  per stage
//   forward : rows = cells (2 tiles of 64), K = genes  (one chunk per stage),  Dense = W1eff^T [D x h1]
//   backward: rows = genes (2 chunks of 64), K = cells (one tile per stage),   Dense = delta    [B x h1]
// Warp roles: 0 = TMA producer (dense operand), 1 = MMA issuer, 2 = TMEM allocator,
//             4-11 = panel densifiers (two groups alternating stages) during the main loop, epilogue afterwards.
// ===========================================================================
namespace {

using namespace tc;

constexpr int BM = 128, BK = 64, UK = 16, STG = 4, TCT = 384;
constexpr int RT = 256;                // run-table entries cached in shared memory per densifier group
constexpr int A_TILE = BM * BK * 2;   // 16 KB
constexpr int PF = 4;                 // entries prefetched per densifier thread per stage (fast path)

template <int BN>
struct tc_cfg {
    static constexpr int B_TILE = BN * BK * 2;
    static constexpr int STAGE = A_TILE + B_TILE;
    static constexpr int SMEM = STG * STAGE + 1024 + 256 + 2 * RT * 16;
};

struct run2_t {   // the two bucket runs of one stage
    int b0, n0, b1, n1;
};

template <bool BWD>
__device__ __forceinline__ run2_t
stage_runs(const int *__restrict__ bp, int kb, int tile, int nT, int nK)
{
    run2_t r;
    if (!BWD) {   // cell tiles 2*tile, 2*tile+1 ; gene chunk kb
        const int T0 = 2 * tile, T1 = T0 + 1;
        const int *p0 = bp + (size_t)T0 * (nK + 1) + kb;
        r.b0 = p0[0]; r.n0 = p0[1] - r.b0;
        if (T1 < nT) {
            const int *p1 = bp + (size_t)T1 * (nK + 1) + kb;
            r.b1 = p1[0]; r.n1 = p1[1] - r.b1;
        } else { r.b1 = 0; r.n1 = 0; }
    } else {      // cell tile kb ; gene chunks 2*tile, 2*tile+1 (adjacent runs)
        const int *p0 = bp + (size_t)kb * (nK + 1) + 2 * tile;
        r.b0 = p0[0]; r.n0 = p0[1] - r.b0;
        r.b1 = p0[1];
        r.n1 = (2 * tile + 1 < nK) ? p0[2] - r.b1 : 0;
    }
    return r;
}

template <bool BWD>
__device__ __forceinline__ void
scatter_entry(uint8_t *tileA, uint32_t e, int rowbase)
{
    const int r6 = (int)(e >> 26), c6 = (int)((e >> 20) & 63);
    const int trow = rowbase + (BWD ? c6 : r6), tcol = BWD ? r6 : c6;
    // K-major SWIZZLE_128B: row = 128 B, 16-byte chunk index XOR (row & 7)
    const int off = trow * 128 + ((((tcol >> 3) ^ (trow & 7))) << 4) + (tcol & 7) * 2;
    *reinterpret_cast<uint16_t *>(tileA + off) = (uint16_t)(e & 0xFFFFu);
}

// SEG: fp32-tolerance mode (three K segments over hi / lo operand copies); compile-time so that the bf16 mode's loops
// carry none of the segment arithmetic
template <int BN, bool BWD, bool SEG>
__global__ void __launch_bounds__(TCT, 1)
senc_tc_kernel(const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmS,
               const __grid_constant__ CUtensorMap tmDlo, senc_tc_args_t a, int nT, int nK)
{
    using cfg = tc_cfg<BN>;
    extern __shared__ uint8_t smem_raw[];
    uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STG * cfg::STAGE);
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 3 * STG + 1);
    const uint32_t fulla0 = smem_u32(bars), fullb0 = smem_u32(bars + STG), empty0 = smem_u32(bars + 2 * STG),
                   tfull = smem_u32(bars + 3 * STG);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.y + (BWD ? a.tile0 : 0), n0 = blockIdx.x * BN;
    // K axis = base_kb pipeline stages (gene chunks forward, cell tiles backward), walked once — or, in the fp32-tolerance
    // mode, three times with different operand copies (segment 0: panel hi x dense lo, 1: panel lo x dense hi, 2: hi x hi)
    const int base_kb = BWD ? nT : nK;
    const int num_kb = SEG ? 3 * base_kb : base_kb;

    if (warp == 0 && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmD) : "memory");
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STG; ++s) {
            mbar_init(fulla0 + 8 * s, 128);   // one densifier group per stage
            mbar_init(fullb0 + 8 * s, 1);
            mbar_init(empty0 + 8 * s, 1);
        }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_u32(tmem_slot)), "r"(BN) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_slot;
    pdl_sync();   // (barrier init and the TMEM allocation above overlap the previous kernel's tail)

    if (warp == 0) {
        // ================= TMA producer: dense operand, [64 k x 64 n] boxes =================
        if (lane == 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STG;
                mbar_wait(empty0 + 8 * s, ((kb / STG) & 1) ^ 1);
                const uint32_t sb = smem_u32(smem + s * cfg::STAGE + A_TILE);
                const uint32_t fb = fullb0 + 8 * s;
                mbar_expect_tx(fb, cfg::B_TILE);
                const int seg = SEG ? kb / base_kb : 0, kk = SEG ? kb - seg * base_kb : kb;
                const CUtensorMap *tm = (SEG && seg == 0) ? &tmDlo : &tmD;
#pragma unroll
                for (int i = 0; i < BN / 64; ++i) tma_load_2d(sb + i * 8192, tm, fb, n0 + 64 * i, kk * BK);
            }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            constexpr uint32_t idesc = make_idesc(false, true, BM, BN);
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % STG;
                const uint32_t ph = (kb / STG) & 1;
                mbar_wait(fullb0 + 8 * s, ph);
                mbar_wait(fulla0 + 8 * s, ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t sa = smem_u32(smem + s * cfg::STAGE), sb = sa + A_TILE;
#pragma unroll
                for (int k = 0; k < BK / UK; ++k) {
                    const uint64_t ad = make_desc(sa + k * 32, 16, 1024);        // K-major, SW128
                    const uint64_t bd = make_desc(sb + k * 2048, 8192, 1024);    // MN-major, SW128
                    umma_bf16(tmem_base, ad, bd, idesc, (kb | k) != 0);
                }
                umma_commit(empty0 + 8 * s);
            }
            umma_commit(tfull);
        }
    } else if (warp >= 4) {
        // ================= densifiers: two groups of 4 warps, alternating stages =================
        const int grp = (warp - 4) >> 2;
        const int t = threadIdx.x - 128 - 128 * grp;   // 0..127 inside the group
        const int *__restrict__ bp = a.bucket_ptr;
        int4 *rtab = reinterpret_cast<int4 *>(smem + STG * cfg::STAGE + 256) + grp * RT;   // run table of this group
        const int n_own = (num_kb - grp + 1) / 2;      // stages grp, grp + 2, ...

        auto refill = [&](int i0) {                    // own-stage indices [i0, i0 + RT)
            asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
            for (int i = t; i < RT; i += 128)
                if (i0 + i < n_own) {
                    const run2_t r = stage_runs<BWD>(bp, SEG ? (grp + 2 * (i0 + i)) % base_kb : grp + 2 * (i0 + i), tile, nT, nK);
                    rtab[i] = make_int4(r.b0, r.n0, r.b1, r.n1);
                }
            asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
        };
        // the panel copy a stage reads: the residual values in segment 1 of the fp32-tolerance mode
        auto ent_of = [&](int i) -> const uint32_t * {
            return (SEG && (grp + 2 * i) / base_kb == 1) ? a.ent_lo : a.ent;
        };
        auto load = [&](uint32_t (&e)[PF], int i) {
            const int4 r = rtab[i % RT];
            const uint32_t *__restrict__ ent = ent_of(i);
#pragma unroll
            for (int q = 0; q < PF; ++q) {
                const int x = t + 128 * q;
                e[q] = x < r.y ? ent[r.x + x] : (x < r.y + r.w ? ent[r.z + x - r.y] : 0u);
            }
        };
        auto build = [&](const uint32_t (&e)[PF], int i) {
            const int kb = grp + 2 * i, s = kb % STG;
            const int4 r = rtab[i % RT];
            const int n = r.y + r.w;
            const uint32_t *__restrict__ ent = ent_of(i);
            mbar_wait(empty0 + 8 * s, ((kb / STG) & 1) ^ 1);
            uint8_t *tileA = smem + s * cfg::STAGE;
            const uint4 z = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
            for (int q = 0; q < A_TILE / (128 * 16); ++q) reinterpret_cast<uint4 *>(tileA)[t + 128 * q] = z;
            asm volatile("bar.sync %0, 128;" ::"r"(1 + grp) : "memory");
#pragma unroll
            for (int q = 0; q < PF; ++q) {
                const int x = t + 128 * q;
                if (x < n) scatter_entry<BWD>(tileA, e[q], x < r.y ? 0 : 64);
            }
            for (int x = t + 128 * PF; x < n; x += 128) {   // dense buckets: slow path
                const uint32_t v = x < r.y ? ent[r.x + x] : ent[r.z + x - r.y];
                scatter_entry<BWD>(tileA, v, x < r.y ? 0 : 64);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(fulla0 + 8 * s) : "memory");
        };

        // pull the runs of a later stage into L2 (backward walks the panel tile by tile: every stage is a
        // first touch of DRAM otherwise, and the register prefetch is only one own-stage deep)
        auto l2_prefetch = [&](int i, int i1) {
            if (i >= i1) return;
            const int4 r = rtab[i % RT];
            const uint32_t *ent = ent_of(i);
            if (t * 32 < r.y) asm volatile("prefetch.global.L2 [%0];" ::"l"(ent + r.x + t * 32));
            if (t * 32 < r.w) asm volatile("prefetch.global.L2 [%0];" ::"l"(ent + r.z + t * 32));
        };
        constexpr int L2D = 4;                           // own-stages ahead (8 pipeline stages)

        uint32_t ea[PF], eb[PF];
        for (int i0 = 0; i0 < n_own; i0 += RT) {
            refill(i0);
            const int i1 = min(n_own, i0 + RT);
            for (int i = i0 + 1; i < i0 + L2D; ++i) l2_prefetch(i, i1);
            load(ea, i0);
            for (int i = i0; i < i1; i += 2) {          // ping-pong register sets: no moves of in-flight loads
                l2_prefetch(i + L2D, i1);
                if (i + 1 < i1) load(eb, i + 1);
                build(ea, i);
                if (i + 1 >= i1) break;
                l2_prefetch(i + 1 + L2D, i1);
                if (i + 2 < i1) load(ea, i + 2);
                build(eb, i + 1);
            }
        }

        // ================= epilogue =================
        mbar_wait(tfull, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int e4 = (warp - 4) & 3;                 // TMEM lanes [32 e4, 32 e4 + 32) (= warp % 4)
        const int trow = 32 * e4 + lane;
        const int c_beg = grp * (BN / 64), c_end = c_beg + BN / 64;   // each group drains half of the columns
        if (!BWD) {
            const int row = tile * BM + trow;
            for (int c = c_beg; c < c_end; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(32 * e4) << 16) + (uint32_t)(32 * c), v);
                const int col0 = n0 + 32 * c;
                if (row < a.B) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        if (col0 + j < a.h1) {   // h1 % 8 == 0
                            float f[8];
#pragma unroll
                            for (int q = 0; q < 8; ++q) {
                                f[q] = __uint_as_float(v[j + q]) + a.bias_eff[col0 + j + q];
                                if (a.relu) f[q] = fmaxf(f[q], 0.f);
                            }
                            float4 *dst = reinterpret_cast<float4 *>(a.H1 + (size_t)row * a.h1 + col0 + j);
                            dst[0] = make_float4(f[0], f[1], f[2], f[3]);
                            dst[1] = make_float4(f[4], f[5], f[6], f[7]);
                            if (a.H1_bf16) {
                                __nv_bfloat162 b0 = __floats2bfloat162_rn(f[0], f[1]), b1 = __floats2bfloat162_rn(f[2], f[3]);
                                __nv_bfloat162 b2 = __floats2bfloat162_rn(f[4], f[5]), b3 = __floats2bfloat162_rn(f[6], f[7]);
                                uint4 pk;
                                pk.x = *reinterpret_cast<uint32_t *>(&b0); pk.y = *reinterpret_cast<uint32_t *>(&b1);
                                pk.z = *reinterpret_cast<uint32_t *>(&b2); pk.w = *reinterpret_cast<uint32_t *>(&b3);
                                *reinterpret_cast<uint4 *>(reinterpret_cast<__nv_bfloat16 *>(a.H1_bf16) +
                                                           (size_t)row * a.h1 + col0 + j) = pk;
                            }
                        }
                    }
                }
            }
        } else {
            // S' = A^T delta tile -> [32 x 32] fp32 boxes in shared memory (SWIZZLE_128B) -> TMA tile store into
            // the S buffer.  The rank-1 correction, the per-gene dots and the x_mean / ln_x_sd gradients are
            // a coalesced row pass afterwards (senc_row_pass_kernel): a thread-per-row epilogue here would
            // touch 32 different lines per request.
            uint8_t *boxes = smem + (size_t)(warp - 4) * (BN / 64) * 4096;   // operand stages are free now
            for (int c = c_beg; c < c_end; ++c) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(32 * e4) << 16) + (uint32_t)(32 * c), v);
                uint8_t *box = boxes + (size_t)(c - c_beg) * 4096;
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    *reinterpret_cast<uint4 *>(box + lane * 128 + ((j ^ (lane & 7)) << 4)) =
                        make_uint4(v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) {
                for (int c = c_beg; c < c_end; ++c) {
                    const int c0 = n0 + 32 * c, c1 = tile * BM + 32 * e4;
                    if (c0 >= a.h1 || c1 >= a.D) continue;
                    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];"
                                 ::"l"(&tmS), "r"(c0), "r"(c1), "r"(smem_u32(boxes + (size_t)(c - c_beg) * 4096)) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
            __syncwarp();
            (void)trow;
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    }

    __syncthreads();
    if (warp == 2) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(BN) : "memory");
    }
}

} // namespace

namespace {

template <int BN, bool BWD>
int
launch_tc(const senc_tc_args_t &a, cudaStream_t st)
{
    using cfg = tc_cfg<BN>;
    const int nT = panel_tiles(a.B), nK = panel_chunks(a.D);
    CUtensorMap tm;
    // dense operand: rows = K index (genes fwd / cells bwd), h1 contiguous
    MMV_TRY(make_tma_map_bf16(&tm, a.dense_bf16, (uint64_t)a.h1, (uint64_t)(BWD ? a.B : a.D), (uint64_t)a.h1, 64, BK));
    CUtensorMap ts = tm, tlo = tm;
    if (BWD) MMV_TRY(make_tma_map_f32(&ts, a.S, (uint64_t)a.h1, (uint64_t)a.D, (uint64_t)a.h1, 32, 32));
    MMV_CHECK((a.ent_lo == nullptr) == (a.dense_lo == nullptr), "senc_tc: the fp32-tolerance mode needs both residual operands");
    if (a.dense_lo) MMV_TRY(make_tma_map_bf16(&tlo, a.dense_lo, (uint64_t)a.h1, (uint64_t)(BWD ? a.B : a.D), (uint64_t)a.h1, 64, BK));
    auto kern = a.ent_lo ? senc_tc_kernel<BN, BWD, true> : senc_tc_kernel<BN, BWD, false>;
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(kern), cfg::SMEM));
    dim3 grid(ceil_div(a.h1, BN), BWD ? (a.ntiles > 0 ? a.ntiles : ceil_div(a.D, BM)) : ceil_div(a.B, BM));
    MMV_CUDA(launch_k(kern, grid, dim3(TCT), cfg::SMEM, st, tm, ts, tlo, a, nT, nK));
    ::mmv::count_launch();
    return 0;
}

// Row pass after the backward GEMM: one warp per gene j, coalesced over the h1 columns.
//   S'[j, :] = sum_i (l_ij / t_j) delta_i   (the panel values carry 1 / t_j)
//   dW1eff[j, :] (+)= S'[j, :] - (x_mean_j / t_j) Delta          nb.hh:406-411 backward, see sparse_encoder.cu
//   d x_mean_j   += -(W_eff[j, :] . Delta) / t_j
//   d t_j         = -(W_eff[j, :] . S' - (x_mean_j / t_j) W_eff[j, :] . Delta) / t_j ; d ln_x_sd_j += d t_j sigmoid(ln_x_sd_j)
__global__ void __launch_bounds__(256)
senc_row_pass_kernel(senc_tc_args_t a, const float *ln_x_sd, float *dx_mean, float *dln_x_sd)
{
    __shared__ float s_ss[8];
    pdl_sync();
    const int lane = threadIdx.x & 31;
    const int j = a.tile0 * 128 + ((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int j_end = a.ntiles > 0 ? min(a.D, (a.tile0 + a.ntiles) * 128) : a.D;
    float ss = 0.f;
    if (j < j_end) {
    const float it = a.inv_t[j], xm = a.x_mean[j], xmit = xm * it;
    const float4 *S4 = reinterpret_cast<const float4 *>(a.S + (size_t)j * a.h1);
    const float4 *W4 = reinterpret_cast<const float4 *>(a.W1T + (size_t)j * a.h1);
    const float4 *D4 = reinterpret_cast<const float4 *>(a.Delta);
    const float4 *C4 = reinterpret_cast<const float4 *>(a.col_scale);
    float4 *O4 = a.dW1T ? reinterpret_cast<float4 *>(a.dW1T + (size_t)j * a.h1) : nullptr;
    float d0 = 0.f, d1 = 0.f;
    for (int n = lane; n < a.h1 / 4; n += 32) {      // h1 % 8 == 0
        const float4 s4 = S4[n], dl = D4[n];
        float4 w = W4[n];
        if (a.angular) {                              // angular.hh:38-40
            const float4 cs = C4[n];
            w.x = (fmaxf(w.x, 0.f) + 1e-4f) * cs.x; w.y = (fmaxf(w.y, 0.f) + 1e-4f) * cs.y;
            w.z = (fmaxf(w.z, 0.f) + 1e-4f) * cs.z; w.w = (fmaxf(w.w, 0.f) + 1e-4f) * cs.w;
        }
        d0 += w.x * s4.x + w.y * s4.y + w.z * s4.z + w.w * s4.w;
        d1 += w.x * dl.x + w.y * dl.y + w.z * dl.z + w.w * dl.w;
        if (O4) {
            float4 o = make_float4(fmaf(-xmit, dl.x, s4.x), fmaf(-xmit, dl.y, s4.y), fmaf(-xmit, dl.z, s4.z),
                                   fmaf(-xmit, dl.w, s4.w));
            if (a.accumulate) {
                const float4 old = O4[n];
                o.x += old.x; o.y += old.y; o.z += old.z; o.w += old.w;
            }
            O4[n] = o;
            ss += o.x * o.x + o.y * o.y + o.z * o.z + o.w * o.w;
        }
    }
    d0 = warp_sum(d0);
    d1 = warp_sum(d1);
    if (lane == 0) {
        dx_mean[j] += -d1 * it;
        const float dt = -(d0 - xmit * d1) * it;
        dln_x_sd[j] += dt * sigmoid_f(ln_x_sd[j]);   // d softplus
    }
    }
    if (a.sumsq) {   // global-norm clip: this gradient's share of ||g||^2, one atomic per CTA
        ss = warp_sum(ss);
        if (lane == 0) s_ss[threadIdx.x >> 5] = ss;
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = 0.f;
            for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_ss[w];
            atomicAdd(a.sumsq, (double)t);
        }
    }
}

// dWnuT[j, h] += sum_e x_e dHnu[row_e, h] ; dwdepth[j] += sum_e x_e ddpre[row_e].
// Grid = gene chunks x kAuxSplit slices of the cell tiles; a CTA accumulates its 64 genes in shared
// memory (one warp per cell tile, lanes over the bucket) and adds them to the gradient once.
constexpr int kAuxSplit = 8;

__global__ void __launch_bounds__(256)
panel_aux_bwd_kernel(const int *bp, const uint32_t *ent, const float *xraw, int nT, int nK, int H,
                     const float *dHnu, const float *ddpre, float *dWnuT, float *dwdepth, int D)
{
    extern __shared__ float s_acc[];   // [64 x (H + 1)]
    const int k = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = H + 1;
    for (int i = tid; i < 64 * W; i += 256) s_acc[i] = 0.f;
    __syncthreads();
    for (int T = blockIdx.y * 8 + warp; T < nT; T += 8 * kAuxSplit) {
        const int *p = bp + (size_t)T * (nK + 1) + k;
        const int b = p[0], e = p[1];
        for (int q = b + lane; q < e; q += 32) {
            const uint32_t v = ent[q];
            const float x = xraw[q];
            const int r = T * 64 + (int)(v >> 26), c = (int)((v >> 20) & 63);
            atomicAdd(&s_acc[c * W + H], x * ddpre[r]);
            for (int h = 0; h < H; ++h) atomicAdd(&s_acc[c * W + h], x * dHnu[(size_t)r * H + h]);
        }
    }
    __syncthreads();
    for (int i = tid; i < 64 * W; i += 256) {
        const int c = i / W, h = i % W, j = k * 64 + c;
        if (j >= D || s_acc[i] == 0.f) continue;
        if (h == H) atomicAdd(dwdepth + j, s_acc[i]);
        else atomicAdd(dWnuT + (size_t)j * H + h, s_acc[i]);
    }
}

// H = 1 (the default nu encoder): no shared-memory float atomics (they compile to compare-and-swap loops, 74 us at cfg 4).
// Every warp owns a private [64 x 2] accumulator; lanes that hold the same gene in one trip are merged with
// match_any + shuffles (the lowest lane of a group collects the others), so a slot has exactly one writer per trip and
// the update is a plain load / add / store; the eight copies are added up once at the end.
__global__ void __launch_bounds__(256)
panel_aux_bwd_h1_kernel(const int *bp, const uint32_t *ent, const float *xraw, int nT, int nK, const float *dHnu,
                        const float *ddpre, float *dWnuT, float *dwdepth, int D)
{
    __shared__ float s_priv[8][64 * 2];
    const int k = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    float *priv = s_priv[warp];
    for (int i = lane; i < 128; i += 32) priv[i] = 0.f;
    __syncwarp();
    pdl_sync();
    for (int T = blockIdx.y * 8 + warp; T < nT; T += 8 * kAuxSplit) {
        const int *p = bp + (size_t)T * (nK + 1) + k;
        const int b = p[0], e = p[1];
        for (int q0 = b; q0 < e; q0 += 32) {
            const int q = q0 + lane;
            const bool on = q < e;
            const unsigned act = __ballot_sync(0xffffffffu, on);
            if (on) {
                const uint32_t v = ent[q];
                const float x = xraw[q];
                const int r = T * 64 + (int)(v >> 26), c = (int)((v >> 20) & 63);
                const float o0 = x * ddpre[r], o1 = x * dHnu[r];      // this lane's own terms (what the shuffles hand out)
                float a0 = o0, a1 = o1;
                const unsigned peers = __match_any_sync(act, c);
                unsigned todo = peers & ~((2u << lane) - 1u);          // the lanes of my group above me
                while (__any_sync(act, todo != 0u)) {
                    const int src = todo ? __ffs(todo) - 1 : lane;
                    const float t0 = __shfl_sync(act, o0, src), t1 = __shfl_sync(act, o1, src);
                    if (todo) { a0 += t0; a1 += t1; todo &= todo - 1u; }
                }
                if ((peers & ((1u << lane) - 1u)) == 0u) {             // lowest lane of the group: the slot's only writer
                    priv[c * 2 + 0] += a0;
                    priv[c * 2 + 1] += a1;
                }
            }
            __syncwarp();
        }
    }
    __syncthreads();
    if (tid < 128) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s_priv[w][tid];
        const int c = tid >> 1, j = k * 64 + c;
        if (j < D && t != 0.f) {
            if (tid & 1) atomicAdd(dWnuT + j, t);
            else atomicAdd(dwdepth + j, t);
        }
    }
}

__global__ void
w1eff_bf16_kernel(const float *W1T, const float *col_scale, __nv_bfloat16 *out, size_t n, int h1)
{
    pdl_sync();
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = __float2bfloat16((fmaxf(W1T[i], 0.f) + 1e-4f) * col_scale[i % h1]);   // angular.hh:38-40
}

} // namespace

bool
senc_tc_supported(int h1)
{
    return h1 >= 16 && h1 % 8 == 0;
}

int
launch_senc_tc_fwd(const senc_tc_args_t &a, cudaStream_t st)
{
    MMV_CHECK(senc_tc_supported(a.h1), "senc_tc: first-layer width %d must be a multiple of 8", a.h1);
    return a.h1 > 128 ? launch_tc<256, false>(a, st) : launch_tc<128, false>(a, st);
}

int
launch_senc_tc_bwd(const senc_tc_args_t &a, cudaStream_t st)
{
    MMV_CHECK(senc_tc_supported(a.h1), "senc_tc: first-layer width %d must be a multiple of 8", a.h1);
    return a.h1 > 128 ? launch_tc<256, true>(a, st) : launch_tc<128, true>(a, st);
}

int
launch_senc_row_pass(const senc_tc_args_t &a, const float *ln_x_sd, float *dx_mean, float *dln_x_sd, cudaStream_t st)
{
    const int64_t genes = a.ntiles > 0 ? std::min<int64_t>(a.D - (int64_t)a.tile0 * 128, (int64_t)a.ntiles * 128) : a.D;
    MMV_CUDA(launch_k(senc_row_pass_kernel, dim3(ceil_div(genes * 32, 256)), dim3(256), 0, st, a, ln_x_sd, dx_mean, dln_x_sd));
    ::mmv::count_launch();
    return 0;
}

int
launch_panel_aux_bwd(const int *bucket_ptr, const uint32_t *ent, const float *xraw, int B, int D, int H,
                     const float *dHnu, const float *ddpre, float *dWnuT, float *dwdepth, cudaStream_t st)
{
    if (H == 1 && dwdepth != nullptr) {
        MMV_CUDA(launch_k(panel_aux_bwd_h1_kernel, dim3(panel_chunks(D), kAuxSplit), dim3(256), 0, st, bucket_ptr, ent, xraw,
                          panel_tiles(B), panel_chunks(D), dHnu, ddpre, dWnuT, dwdepth, D));
        ::mmv::count_launch();
        return 0;
    }
    const size_t smem = sizeof(float) * 64 * (size_t)(H + 1);
    MMV_CHECK(smem <= 48 * 1024, "panel_aux_bwd: nu hidden width %d too large", H);
    panel_aux_bwd_kernel<<<dim3(panel_chunks(D), kAuxSplit), 256, smem, st>>>(bucket_ptr, ent, xraw, panel_tiles(B), panel_chunks(D), H,
                                                             dHnu, ddpre, dWnuT, dwdepth, D);
    ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_w1eff_bf16(const float *W1T, const float *col_scale, void *out, int D, int h1, cudaStream_t st)
{
    const size_t n = (size_t)D * h1;
    MMV_CUDA(launch_k(w1eff_bf16_kernel, dim3((unsigned)min_sz((n + 255) / 256, 148 * 16)), dim3(256), 0, st, W1T, col_scale,
                      reinterpret_cast<__nv_bfloat16 *>(out), n, h1));
    ::mmv::count_launch();
    return 0;
}

} // namespace mmv
