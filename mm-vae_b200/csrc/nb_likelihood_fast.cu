// K4, tuned path (default overdispersion latent Rnu = 1, all-ones covariate) — same math as
// nb_likelihood.cu (nb.hh:433-460,511-531 + autograd), organised around what the ncu captures
// under profiles/ showed.  Round 1 (r01 step 1-10): cluster.sync stalls, serial global-load chains,
// a recomputing patch pass; then — with those gone — 1 390 instructions per warp and row at 48 %
// issue utilisation, bound by the serial phases of a row: two reductions with a cluster hop each,
// four CTA barriers, the softmax pass holding every logit in registers.  Round 2 (this file):
//   * the softmax statistics are NOT computed here: the decoder-last GEMM adds the per-gene bias in
//     its epilogue and leaves per-(row, 256-column tile) max / sum-of-exp2 partials (gemm_tc.cu),
//     a tiny merge kernel turns them into (max, 1 / sum, depth, z_nu) per row (nb_row_meta).  K4 then
//     needs ONE pass over a row and ONE cluster reduction (s = sum g p, d z_nu);
//   * that reduction is software-pipelined: row i's hop travels while row i + 1 is swept, and row i's
//     dL/dlogit is written after row i + 1's sweep (its p and g wait in shared memory) — no thread
//     ever waits for a hop, and a row costs two CTA barriers;
//     (tried and measured slower: no CTA barrier at all — mbarrier arrive / wait + "last warp sends" 0.58 ms; the
//     output pass moved to the end of the iteration for a full iteration of slack on the hop, 0.75 ms);
//   * a thread-block cluster owns whole rows, each CTA a fixed gene slice for the entire kernel:
//     per-gene constants staged in shared memory once, per-gene column sums of the nu gradient in
//     registers; two CTAs (of different clusters) share an SM;
//   * rows arrive through a three-deep cp.async.bulk + mbarrier ring (1.5 rows of lead time); the
//     landing buffer is overwritten in place with p, so the ring is also the store of the deferred
//     output pass;
//   * the exchange crosses the cluster with st.async into every peer's slot + a remote mbarrier
//     complete_tx — no cluster barrier, no memory fence, no L1 invalidate;
//   * a pre-kernel bisects every cell's CSR run by gene slice, so the nonzero fix-up touches only
//     its own entries; its indices/values are prefetched one row ahead; it overwrites g in shared
//     memory before the single coalesced dL/dlogit store (bf16x2 or float2 per thread);
//   * dense sweep on packed fp32 pairs, branch-free (padded slices), 5 MUFU per element (softmax ex2,
//     ex2 + lg2 for softplus, ONE rcp for 1 / (1 + t) and 1 / (nu (m + nu)), lg2 for log1p(m / nu));
//     log1p polynomials on the FMA pipe; small integer counts use product forms of lgamma / digamma.
#include "common.cuh"
#include "kernels.hh"
#include "rowstream.cuh"

#include <cooperative_groups.h>
#include <type_traits>
#include <cuda_bf16.h>

namespace mmv {

namespace cg = cooperative_groups;

namespace {

using namespace rowstream;

#ifndef MMV_K4_XWAIT_NS
#define MMV_K4_XWAIT_NS 0      // suspend-time hint (ns) of the wait for the previous row's cluster hop; 0 = plain polling
#endif
#ifndef MMV_K4_LWAIT_NS
#define MMV_K4_LWAIT_NS 0      // the same for the wait on the row's TMA landing
#endif

constexpr int NZP = 2;                                         // nonzeros prefetched per thread per row
constexpr float kEps = 1e-4f, kNuMin = 1e-4f, kNuMax = 1e4f;   // nb.hh:514,252,459

// log(1 + y), y >= 0: Taylor polynomial on the FMA pipe below 1/16 (where log1p(y) - y / (1 + y)
// cancels and lg2.approx's absolute error near 1 would dominate), MUFU lg2 above
__device__ __forceinline__ float
log1p_fast(float y)
{
    const float poly = y * (1.f - y * (0.5f - y * (0.33333334f - y * (0.25f - y * 0.2f))));
    const float big = lg2_f(1.f + y) * kLn2;
    return y < 0.0625f ? poly : big;
}

struct elem_t {
    float m, nu, l1p, inv_mnu, sig_mask;
};

// Dense sweep outputs of one adjacent gene pair (x = 0): p (softmax), g = dl/dm = nu / (m + nu), nl = nu log1p(m / nu),
// du = (log1p(m / nu) - m / (m + nu)) dnu/du (NOT yet divided by B: the row / column sums are scaled once at the end).
struct pair_t {
    f2_t p, g, nl, du;
};

// rounding-error compensation of log1p(t) = log(w) + (t - (w - 1)) / w, w = fl(1 + t): the term only matters
// for small t (it is <= 2^-24 absolute), where 1 / w = 1 - t to the accuracy that matters; beyond t = 1 it is dropped
__device__ __forceinline__ f2_t
softplus_comp(f2_t t, f2_t w)
{
    const f2_t one = bc2(1.f), neg1 = bc2(-1.f);
    float t0, t1;
    up2(t, t0, t1);
    const f2_t rwa = fma2(pk2(fminf(t0, 1.f), fminf(t1, 1.f)), neg1, one);   // max(1 - t, 0)
    return mul2(fma2(add2(w, neg1), neg1, t), rwa);
}

// Dense sweep of one adjacent gene pair assuming x = 0.  v = softmax numerators, u = (zn wn + ub) log2(e).
// One reciprocal per element: q = 1 / (w nu (m + nu)) gives 1 / w = q nu (m + nu) (the sigmoid) and
// 1 / (nu (m + nu)) = q w (range: w <= e^20 + 1, nu <= 1e4, m + nu <= depth + 1e4 — far inside fp32).
// CLAMPS = false: the caller has shown that every u of the row lies inside (log(expm1(1e-4)), 20), so the
// softplus threshold, the clamp of nb.hh:459 and its gradient mask cannot act.
template <bool CLAMPS>
__device__ __forceinline__ pair_t
pair_v4(f2_t v, f2_t scale, f2_t depth, f2_t u)
{
    const f2_t one = bc2(1.f), neg1 = bc2(-1.f), ln2 = bc2(kLn2), eps = bc2(kEps);
    constexpr float c20 = 20.f * kLog2e;                              // softplus threshold 20 (nb.hh:458)
    pair_t o;
    o.p = mul2(v, scale);
    const f2_t m = fma2(o.p, depth, eps);                             // nb.hh:519
    float u0, u1;
    up2(u, u0, u1);
    const f2_t t = pk2(ex2_f(CLAMPS ? fminf(u0, c20) : u0), ex2_f(CLAMPS ? fminf(u1, c20) : u1));
    const f2_t w = add2(t, one);
    float w0, w1;
    up2(w, w0, w1);
    f2_t sp = fma2(pk2(lg2_f(w0), lg2_f(w1)), ln2, softplus_comp(t, w));
    f2_t nu;
    bool in0 = true, in1 = true;
    if (CLAMPS) {
        // beyond the threshold softplus(u) = u = log1p(e^20) + (u - 20)
        const f2_t over = add2(u, bc2(-c20));
        float v0, v1;
        up2(over, v0, v1);
        sp = fma2(pk2(fmaxf(v0, 0.f), fmaxf(v1, 0.f)), ln2, sp);
        float s0, s1;
        up2(sp, s0, s1);
        const float r0 = fminf(fmaxf(s0, kNuMin), kNuMax), r1 = fminf(fmaxf(s1, kNuMin), kNuMax);   // nb.hh:459
        nu = add2(pk2(r0, r1), eps);                                  // nb.hh:518
        in0 = r0 == s0; in1 = r1 == s1;                               // clamp passes the gradient inside [min, max]
    } else {
        nu = add2(sp, eps);
    }
    const f2_t mnu = add2(m, nu);
    const f2_t nm = mul2(nu, mnu);
    float q0, q1;
    up2(mul2(nm, w), q0, q1);
    const f2_t q = pk2(rcp_f(q0), rcp_f(q1));                         // the element's only reciprocal
    f2_t sig = mul2(t, mul2(q, nm));                                  // d softplus / du = t / w (-> 1 beyond the threshold)
    if (CLAMPS) {
        float g0, g1;
        up2(sig, g0, g1);
        sig = pk2(in0 ? g0 : 0.f, in1 ? g1 : 0.f);
    }
    const f2_t qn = mul2(q, w);                                       // 1 / (nu (m + nu))
    const f2_t minv = mul2(m, mul2(qn, nu));                          // m / (m + nu)
    const f2_t y = mul2(m, mul2(qn, mnu));                            // m / nu
    // log1p(y): Taylor polynomial below 1/16, MUFU lg2 above (see log1p_fast)
    const f2_t ny = mul2(y, neg1);
    f2_t poly = fma2(ny, bc2(0.2f), bc2(0.25f));
    poly = fma2(ny, poly, bc2(0.33333334f));
    poly = fma2(ny, poly, bc2(0.5f));
    poly = fma2(ny, poly, one);
    poly = mul2(y, poly);
    float y0, y1, z0, z1, p0, p1, b0, b1;
    up2(y, y0, y1);
    up2(add2(y, one), z0, z1);
    up2(poly, p0, p1);
    up2(mul2(pk2(lg2_f(z0), lg2_f(z1)), ln2), b0, b1);
    const f2_t l1p = pk2(y0 < 0.0625f ? p0 : b0, y1 < 0.0625f ? p1 : b1);
    o.g = fma2(minv, neg1, one);                                      // nu / (m + nu)
    o.nl = mul2(nu, l1p);
    o.du = mul2(fma2(minv, neg1, l1p), sig);
    return o;
}
// u2 = u * log2(e): the per-gene nu weights are staged pre-scaled so that exp(u) is one ex2.
// 6 MUFU: ex2, rcp(1 + t), lg2(1 + t), rcp(nu (m + nu)), lg2(1 + m / nu) (+ the softmax ex2 outside).
__device__ __forceinline__ elem_t
elem_fast(float p, float depth, float u2)
{
    elem_t e;
    e.m = fmaf(p, depth, kEps);                                  // nb.hh:519
    const bool big = u2 > 20.f * kLog2e;                         // softplus threshold 20 (nb.hh:458)
    const float t = ex2_f(fminf(u2, 20.f * kLog2e));
    const float w = 1.f + t, rw = rcp_f(w);
    // log1p(t) = log(w) + (t - (w - 1)) / w : the second term restores what rounding 1 + t lost
    float sp = fmaf(lg2_f(w), kLn2, (t - (w - 1.f)) * rw);
    sp = big ? u2 * kLn2 : sp;
    const float r = fminf(fmaxf(sp, kNuMin), kNuMax);            // nb.hh:459
    e.nu = r + kEps;                                             // nb.hh:518
    const float sig = big ? 1.f : t * rw;                        // d softplus / du
    e.sig_mask = r == sp ? sig : 0.f;                            // clamp passes the gradient inside [min, max]
    const float mnu = e.m + e.nu;
    const float q = rcp_f(e.nu * mnu);                           // one MUFU for both reciprocals
    e.inv_mnu = q * e.nu;
    e.l1p = log1p_fast(e.m * (q * mnu));                         // log(m + nu) - log(nu)
    return e;
}

// lgamma(nu) - lgamma(nu + x) + lgamma(x + 1) and digamma(nu) - digamma(nu + x) for a count x.
// Small integer counts (the bulk of single-cell data) use the exact product forms
//   lgamma(nu + x) - lgamma(nu) = log prod_{k<x} (nu + k),  digamma(nu + x) - digamma(nu) = P'(nu) / P(nu)
// on the FMA pipe with one lg2 and one rcp; anything else falls back to libdevice.
__constant__ float kLogFact[9] = { 0.f, 0.f, 0.6931471806f, 1.7917594692f, 3.1780538303f, 4.7874917428f,
                                   6.5792512120f, 8.5251613611f, 10.6046029027f };

// lgamma(z) and digamma(z) for z >= 8 by Stirling's series (error < 3e-8 / 2e-9 relative to the leading terms):
//   lgamma(z) = (z - 1/2) ln z - z + ln(2 pi)/2 + 1/(12 z) - 1/(360 z^3) + 1/(1260 z^5)
//   psi(z)    = ln z - 1/(2 z) - 1/(12 z^2) + 1/(120 z^4) - 1/(252 z^6)
__device__ __forceinline__ void
stirling(float z, float &lg, float &dg)
{
    const float lz = logf(z), r = rcp_f(z), r2 = r * r;   // accurate log: (z - 1/2) ln z is ~1e4 for counts in the thousands
    lg = fmaf(z - 0.5f, lz, -z) + 0.9189385332f + r * (0.0833333333f - r2 * (0.00277777778f - r2 * 0.000793650794f));
    dg = lz - 0.5f * r - r2 * (0.0833333333f - r2 * (0.00833333333f - r2 * 0.00396825397f));
}

__device__ __forceinline__ void
count_terms(float nu, float x, float &lg_terms, float &dig_diff)
{
    const int xi = (int)x;
    if ((float)xi == x && xi >= 1 && xi <= 8) {
        float P = nu, dP = 1.f;
        for (int k = 1; k < xi; ++k) {
            const float f = nu + (float)k;
            dP = fmaf(dP, f, P);
            P *= f;
        }
        lg_terms = kLogFact[xi] - __logf(P);
        dig_diff = -dP * rcp_f(P);
    } else if (x > 8.f) {
        // larger counts (a warp meets one almost every row): nu is shifted up by 8 with the product form
        // (lgamma(nu) = lgamma(nu + 8) - ln prod_{k<8}(nu + k), same for digamma), nu + x and x + 1 are >= 9,
        // so all three lgammas and both digammas are Stirling series: ~70 instructions instead of five libdevice calls
        float P = 1.f, dP = 0.f, z = nu;
        if (nu < 8.f) {
            P = nu; dP = 1.f;
#pragma unroll
            for (int k = 1; k < 8; ++k) {
                const float f = nu + (float)k;
                dP = fmaf(dP, f, P);
                P *= f;
            }
            z = nu + 8.f;
        }
        float lg_n, dg_n, lg_nx, dg_nx, lg_x1, dg_x1;
        stirling(z, lg_n, dg_n);
        stirling(nu + x, lg_nx, dg_nx);
        stirling(x + 1.f, lg_x1, dg_x1);
        lg_terms = (lg_n - lg2_f(P) * kLn2) - lg_nx + lg_x1;
        dig_diff = (dg_n - dP * rcp_f(P)) - dg_nx;
    } else {   // non-integer value below 8 (the reference's reader accepts decimals): libdevice
        lg_terms = lgammaf(nu) + lgammaf(x + 1.f) - lgammaf(nu + x);
        dig_diff = digamma_f(nu) - digamma_f(nu + x);
    }
}
// ---------------------------------------------------------------------------
// pre-kernel: for every minibatch row, the absolute CSR offsets at which each gene slice begins
// (bounds[row * (cs + 1) + c]); one warp per row.  Unsorted rows get the whole run for every slice
// (the kernel's range test filters).
// ---------------------------------------------------------------------------
__device__ __forceinline__ void
slice_bounds_body(const int64_t *indptr, const int32_t *indices, const int32_t *cells, const int *unsorted, int B, int cs,
                  int Ds, int64_t *bounds, int t)
{
    // one thread per (row, boundary): lower_bound of c * Ds in the cell's sorted gene ids
    const int row = t / (cs + 1), c = t % (cs + 1);
    if (row >= B) return;
    const int64_t cell = cells ? (int64_t)cells[row] : (int64_t)row;
    const int64_t beg = indptr[cell], end = indptr[cell + 1];
    const bool sorted = unsorted != nullptr && *unsorted == 0;
    int64_t v = c == 0 ? beg : end;                          // unsorted: every slice scans the whole run
    if (sorted && c > 0 && c < cs) {
        const int key = c * Ds;
        int64_t lo = beg, hi = end;
        while (lo < hi) {
            const int64_t mid = (lo + hi) >> 1;
            if (indices[mid] < key) lo = mid + 1; else hi = mid;
        }
        v = lo;
    }
    bounds[(size_t)row * (cs + 1) + c] = v;
}

__global__ void
slice_bounds_kernel(const int64_t *indptr, const int32_t *indices, const int32_t *cells, const int *unsorted, int B,
                    int cs, int Ds, int64_t *bounds)
{
    slice_bounds_body(indptr, indices, cells, unsorted, B, cs, Ds, bounds, blockIdx.x * blockDim.x + threadIdx.x);
}

// ---------------------------------------------------------------------------
// per-row metadata of the tuned kernel: (max of the logits in the log2 domain, 1 / sum of exp2, depth, z_nu)
//   tile_stats != nullptr: merge the [n_tiles x B] partials the decoder-last GEMM's epilogue left
//   tile_stats == nullptr: one pass over G (fp32 mode, or a GEMM variant without the statistics epilogue)
// ---------------------------------------------------------------------------
__device__ __forceinline__ void
row_meta_merge(const float2 *tile_stats, int n_tiles, const float *dpre, const float *znu, int B, float4 *meta, int row)
{
    if (row >= B) return;
    float gm = -INFINITY, gs = 0.f;
    for (int t0 = 0; t0 < n_tiles; t0 += 8) {    // coalesced over rows, 8 partials in flight; online merge in the log2 domain
        float2 p[8];
#pragma unroll
        for (int u = 0; u < 8; ++u)
            p[u] = t0 + u < n_tiles ? __ldg(&tile_stats[(size_t)(t0 + u) * B + row]) : make_float2(-INFINITY, 0.f);
        float cm = p[0].x;
#pragma unroll
        for (int u = 1; u < 8; ++u) cm = fmaxf(cm, p[u].x);
        if (cm > -INFINITY) {
            const float nm = fmaxf(gm, cm);
            float add = 0.f;
#pragma unroll
            for (int u = 0; u < 8; ++u) add += p[u].x > -INFINITY ? p[u].y * ex2_f(p[u].x - nm) : 0.f;
            gs = gs * (gm > -INFINITY ? ex2_f(gm - nm) : 0.f) + add;
            gm = nm;
        }
    }
    const float dp = dpre[row];
    meta[row] = make_float4(gm, 1.f / gs, dp > 20.f ? dp : log1pf(expf(dp)), znu[row]);   // nb.hh:498 (exact, once per row)
}

__global__ void
nb_row_meta_merge_kernel(const float2 *tile_stats, int n_tiles, const float *dpre, const float *znu, int B, float4 *meta)
{
    row_meta_merge(tile_stats, n_tiles, dpre, znu, B, meta, blockIdx.x * blockDim.x + threadIdx.x);
}

// both pre-kernels of the tuned kernel in one launch: blocks [0, meta_blocks) merge the GEMM epilogue's partials into the
// row metadata, the others bisect the CSR runs by gene slice (slice_bounds_kernel's body)
__global__ void
nb_row_prep_kernel(const float2 *tile_stats, int n_tiles, const float *dpre, const float *znu, float4 *meta, int meta_blocks,
                   const int64_t *indptr, const int32_t *indices, const int32_t *cells, const int *unsorted, int B, int cs,
                   int Ds, int64_t *bounds)
{
    pdl_sync();
    if ((int)blockIdx.x < meta_blocks) {
        row_meta_merge(tile_stats, n_tiles, dpre, znu, B, meta, blockIdx.x * blockDim.x + threadIdx.x);
        return;
    }
    slice_bounds_body(indptr, indices, cells, unsorted, B, cs, Ds, bounds, (blockIdx.x - meta_blocks) * blockDim.x + threadIdx.x);
}

__global__ void __launch_bounds__(256)
nb_row_meta_scan_kernel(const float *G, int D, const float *dpre, const float *znu, float4 *meta)
{
    __shared__ float sm[8], ss[8];
    const int row = blockIdx.x, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const float4 *g4 = reinterpret_cast<const float4 *>(G + (size_t)row * D);   // D % 4 == 0, rows 16-byte aligned
    float m = -INFINITY, s = 0.f;
    for (int i = threadIdx.x; i < D / 4; i += 256) {
        const float4 v = g4[i];
        const float cm = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w)) * kLog2e;
        const float nm = fmaxf(m, cm);
        s = s * (m > -INFINITY ? ex2_f(m - nm) : 0.f) + ex2_f(fmaf(v.x, kLog2e, -nm)) + ex2_f(fmaf(v.y, kLog2e, -nm)) +
            ex2_f(fmaf(v.z, kLog2e, -nm)) + ex2_f(fmaf(v.w, kLog2e, -nm));
        m = nm;
    }
    const float wm = warp_max(m);
    const float wsum = warp_sum(m > -INFINITY ? s * ex2_f(m - wm) : 0.f);
    if (lane == 0) { sm[warp] = wm; ss[warp] = wsum; }
    __syncthreads();
    if (warp == 0) {
        const float bm = lane < 8 ? sm[lane] : -INFINITY;
        const float gm = warp_max(bm);
        const float gs = warp_sum(lane < 8 && bm > -INFINITY ? ss[lane] * ex2_f(bm - gm) : 0.f);
        if (lane == 0) {
            const float dp = dpre[row];
            meta[row] = make_float4(gm, 1.f / gs, dp > 20.f ? dp : log1pf(expf(dp)), znu[row]);
        }
    }
}

// LO: the gradient leaves as a bf16 pair hi + lo (fp32-tolerance mode on the tensor cores); a separate instantiation so that
// the bf16 mode's register allocation is untouched
template <int EPT, int T, bool LO = false>
__global__ void __launch_bounds__(T, 2)
nb_likelihood_fast_kernel(nb_lik_args_t a, int cs, int Ds)
{
    static_assert(EPT % 2 == 0 && T % 32 == 0 && T <= 512, "threads own adjacent gene pairs");
    constexpr int kT = T, NW = T / 32;
    constexpr int HP = EPT / 2;                  // pairs per thread
    constexpr int Dp = EPT * kT;                 // padded slice length (>= Ds)
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = cs > 1 ? (int)cluster.block_rank() : 0;
    const int cluster_id = blockIdx.x / cs, n_clusters = gridDim.x / cs;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, D = a.D;
    const int j0 = rank * Ds, nj = max(0, min(D - j0, Ds));
    const bool grad = a.dlogits != nullptr || a.dlogits_bf16 != nullptr;
    const bool sorted = a.unsorted != nullptr && *a.unsorted == 0;

    extern __shared__ __align__(16) float smem[];
    float *s_land = smem;                        // [3][Dp] TMA landing ring (complete logits); swept in place into p
    float *s_gb = s_land + 3 * Dp;               // [2][Dp] g = dl/dm of a row (dense value, nonzeros overwritten), by row parity
    float *s_ub = s_gb + 2 * Dp;                 // (nu_decoding.bias - nu_bias) log2(e)
    float *s_wn = s_ub + Dp;                     // nu_decoding.weight[:, 0] log2(e)
    float *wstat = s_wn + Dp;                    // [2 parity][NW][2] per-warp partials of the row reduction
    float *xslot = wstat + 2 * NW * 2;           // [2 parity][8 ranks][2]
    uint64_t *bars = reinterpret_cast<uint64_t *>(xslot + 32);   // [3] landing ring + [2 parity] exchange
    const uint32_t lbar0 = smem_addr(bars), xbar0 = lbar0 + 24;

    if (tid == 0) {
        for (int i = 0; i < 5; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(lbar0 + 8 * i));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    // extremes of the slice's nu-decoder constants: a row whose u = zn wn + ub provably stays inside
    // (log(expm1(1e-4)), 20) takes the sweep without softplus threshold / clamp / gradient mask
    float wlo = INFINITY, whi = -INFINITY, blo = INFINITY, bhi = -INFINITY;
    for (int jj = tid; jj < Dp; jj += kT) {
        const int j = j0 + jj;
        // padding: logit -inf (softmax numerator 0: the landing buffers' tails are never written, neither by TMA nor by
        // the sweep), u -> -inf (nu clamped, gradient masked), weight 0
        float ub = -1e30f, wn = 0.f;
        if (jj < nj) {
            ub = a.ubias[j] * kLog2e;   // pre-scaled: u * log2(e) = zn * wn' + ub'
            wn = a.Wnd[j] * kLog2e;
            wlo = fminf(wlo, wn); whi = fmaxf(whi, wn);
            blo = fminf(blo, ub); bhi = fmaxf(bhi, ub);
        }
        s_ub[jj] = ub;
        s_wn[jj] = wn;
        if (jj >= nj) s_land[jj] = s_land[Dp + jj] = s_land[2 * Dp + jj] = -INFINITY;
    }
    wlo = -warp_max(-wlo); whi = warp_max(whi); blo = -warp_max(-blo); bhi = warp_max(bhi);
    float *ext = s_gb;                           // (scratch before the row loop starts)
    if (lane == 0) { ext[warp * 4 + 0] = wlo; ext[warp * 4 + 1] = whi; ext[warp * 4 + 2] = blo; ext[warp * 4 + 3] = bhi; }
    if (cs > 1) cluster.sync(); else __syncthreads();   // barriers initialised cluster-wide before any remote signal
    for (int w8 = 0; w8 < NW; ++w8) {
        wlo = fminf(wlo, ext[w8 * 4 + 0]); whi = fmaxf(whi, ext[w8 * 4 + 1]);
        blo = fminf(blo, ext[w8 * 4 + 2]); bhi = fmaxf(bhi, ext[w8 * 4 + 3]);
    }
    __syncthreads();                                     // the scratch is the g buffer of row 0

    const uint32_t row_bytes = (uint32_t)nj * 4u;
    const int row0 = cluster_id;
    if (tid == 0 && nj > 0) {                            // rows 0 and 1 of this cluster
        if (row0 < a.B) bulk_load(smem_addr(s_land), a.G + (size_t)row0 * D + j0, row_bytes, lbar0);
        if (row0 + n_clusters < a.B)
            bulk_load(smem_addr(s_land + Dp), a.G + (size_t)(row0 + n_clusters) * D + j0, row_bytes, lbar0 + 8);
    }

    f2_t cdu[HP], cduz[HP];            // per-gene column sums of d nu-logit (x zn), un-normalised (1/B at the end)
#pragma unroll
    for (int h = 0; h < HP; ++h) cdu[h] = cduz[h] = 0ull;
    f2_t nll2 = 0ull;
    float nll_acc = 0.f;

    // per-row metadata, one row ahead: (max, 1 / sum, depth, zn), the slice's CSR range, the first NZP nonzeros per thread
    const float4 *rowmeta = reinterpret_cast<const float4 *>(a.rowmeta);
    const int bstride = cs + 1;
    float4 meta_n = make_float4(0.f, 0.f, 0.f, 0.f);
    int64_t qlo_n = 0, qhi_n = 0;
    int pj[NZP];
    float px[NZP];
    if (row0 < a.B) {
        meta_n = rowmeta[row0];
        qlo_n = a.slice_bounds[(size_t)row0 * bstride + (sorted ? rank : 0)];
        qhi_n = a.slice_bounds[(size_t)row0 * bstride + (sorted ? rank + 1 : cs)];
#pragma unroll
        for (int i = 0; i < NZP; ++i) {
            const int64_t q = qlo_n + tid + i * kT;
            pj[i] = q < qhi_n ? a.indices[q] : -1;
            px[i] = q < qhi_n ? a.vals[q] : 0.f;
        }
    }

    // the row whose dL/dlogit is still to be written (its p and g sit in shared memory, its hop is in flight)
    int prev_row = -1, prev_slot = 0;
    float prev_depth = 0.f, prev_coef = 0.f;

    int slot = 0;                                        // it % 3
    for (int it = 0;; ++it) {
        const int row = row0 + it * n_clusters;
        const bool active = row < a.B;
        const int par = it & 1;
        float *sL = s_land + slot * Dp;
        float *sg = s_gb + par * Dp;
        float *ws = wstat + par * NW * 2;
        float zn = 0.f, depth = 0.f, gm2 = 0.f, invS = 0.f, coef = 0.f;
        int64_t q_lo = 0, q_hi = 0;
        f2_t red0 = 0ull, red1 = 0ull;     // s = sum g p ; d znu (un-normalised)

        if (active) {
            gm2 = meta_n.x; invS = meta_n.y; depth = meta_n.z; zn = meta_n.w;
            q_lo = qlo_n; q_hi = qhi_n;
            coef = depth * a.inv_B;
            const int next = row + n_clusters;
            if (next < a.B) {                        // next row's scalars and CSR range (consumed next iteration)
                meta_n = rowmeta[next];
                qlo_n = a.slice_bounds[(size_t)next * bstride + (sorted ? rank : 0)];
                qhi_n = a.slice_bounds[(size_t)next * bstride + (sorted ? rank + 1 : cs)];
            }
#if MMV_K4_LWAIT_NS > 0
            if (nj > 0) bar_wait_hint(lbar0 + 8 * slot, (uint32_t)(it / 3) & 1u, MMV_K4_LWAIT_NS);
#else
            if (nj > 0) bar_wait(lbar0 + 8 * slot, (uint32_t)(it / 3) & 1u);
#endif

            // ---- the sweep: softmax from the row statistics, dense NB terms assuming x = 0 (lgammas cancel, no
            //      digamma).  Branch-free: lanes on the padding compute on the constants and are masked at the sums.
            const f2_t l2e = bc2(kLog2e), ngm = bc2(-gm2), scale2 = bc2(invS), depth2 = bc2(depth), zn2 = bc2(zn);
            // u' = u log2(e) over the slice lies in [ulo, uhi] (interval arithmetic on zn wn' + ub')
            const float ulo = fminf(zn * wlo, zn * whi) + blo, uhi = fmaxf(zn * wlo, zn * whi) + bhi;
            const bool plain = ulo > -13.f && uhi < 28.5f;    // log2(e) * (-9.21, 20) = (-13.29, 28.85), with margin
            auto sweep = [&](auto clamps) {
#pragma unroll
                for (int h = 0; h < HP; ++h) {
                    // warp-uniform skip of a warp that sits wholly on the padding (only the last two pair slots can:
                    // the launcher picks the slice class so)
                    if (h >= HP - 2 && h * 2 * kT + warp * 64 >= nj) continue;
                    const int jj = 2 * tid + h * 2 * kT;
                    float y0, y1;
                    up2(fma2(*reinterpret_cast<const f2_t *>(sL + jj), l2e, ngm), y0, y1);
                    const f2_t v = pk2(ex2_f(y0), ex2_f(y1));           // ex2(-inf) = 0 on the padding
                    const f2_t wn2 = *reinterpret_cast<const f2_t *>(s_wn + jj);
                    const f2_t ub2 = *reinterpret_cast<const f2_t *>(s_ub + jj);
                    const pair_t e = pair_v4<decltype(clamps)::value>(v, scale2, depth2, fma2(zn2, wn2, ub2));
                    const bool valid = h < HP - 2 || jj < nj;
                    nll2 = fma2(e.nl, bc2(valid ? 1.f : 0.f), nll2);
                    red0 = fma2(e.g, e.p, red0);
                    cdu[h] = add2(cdu[h], e.du);
                    cduz[h] = fma2(e.du, zn2, cduz[h]);
                    red1 = fma2(e.du, wn2, red1);
                    if (valid) *reinterpret_cast<f2_t *>(sL + jj) = e.p;   // the padding keeps its -inf
                    *reinterpret_cast<f2_t *>(sg + jj) = e.g;
                }
            };
            if (plain) sweep(std::false_type{}); else sweep(std::true_type{});
        }

        // ---- deferred output of the previous row: its hop left a whole sweep ago
        if (prev_row >= 0 && grad) {
            const int pp = par ^ 1;
            const float *xs = xslot + pp * 16;
#if MMV_K4_XWAIT_NS > 0
            bar_wait_hint(xbar0 + 8 * pp, (uint32_t)((it - 1) >> 1) & 1u, MMV_K4_XWAIT_NS);
#else
            bar_wait(xbar0 + 8 * pp, (uint32_t)((it - 1) >> 1) & 1u);
#endif
            float s = 0.f, dz = 0.f;
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r < cs) { s += xs[r * 2]; dz += xs[r * 2 + 1]; }
            const float *sLp = s_land + prev_slot * Dp;
            const float *sgp = s_gb + pp * Dp;
            // dL/dlogit_j = (d_i / B) p_j (g_j - s_i), one coalesced store per pair
            const f2_t ns2 = bc2(-s), coef2 = bc2(prev_coef);
            if (a.dlogits_bf16) {
                __nv_bfloat16 *orow = reinterpret_cast<__nv_bfloat16 *>(a.dlogits_bf16) + (size_t)prev_row * D + j0 + 2 * tid;
#pragma unroll
                for (int h = 0; h < HP; ++h) {
                    const int jj = 2 * tid + h * 2 * kT;
                    if (h >= HP - 2 && jj >= nj) continue;
                    float o0, o1;
                    up2(mul2(mul2(*reinterpret_cast<const f2_t *>(sLp + jj), coef2), add2(*reinterpret_cast<const f2_t *>(sgp + jj), ns2)), o0, o1);
                    const __nv_bfloat162 hi2 = __floats2bfloat162_rn(o0, o1);
                    *reinterpret_cast<__nv_bfloat162 *>(orow + h * 2 * kT) = hi2;
                    if (LO) {
                        __nv_bfloat16 *lrow = reinterpret_cast<__nv_bfloat16 *>(a.dlogits_lo) + (size_t)prev_row * D + j0 + 2 * tid;
                        *reinterpret_cast<__nv_bfloat162 *>(lrow + h * 2 * kT) =
                            __floats2bfloat162_rn(o0 - __low2float(hi2), o1 - __high2float(hi2));
                    }
                }
            } else {
                float *orow = a.dlogits + (size_t)prev_row * D + j0 + 2 * tid;
#pragma unroll
                for (int h = 0; h < HP; ++h) {
                    const int jj = 2 * tid + h * 2 * kT;
                    if (h >= HP - 2 && jj >= nj) continue;
                    float o0, o1;
                    up2(mul2(mul2(*reinterpret_cast<const f2_t *>(sLp + jj), coef2), add2(*reinterpret_cast<const f2_t *>(sgp + jj), ns2)), o0, o1);
                    *reinterpret_cast<float2 *>(orow + h * 2 * kT) = make_float2(o0, o1);
                }
            }
            if (rank == 0 && tid == 0) {
                // d softplus(dp) / d dp = sigmoid(dp) = 1 - exp(-softplus(dp)): from the depth the row metadata carries — the
                // pre-activation itself would be one more global load per row that EVERY warp waits for (6 % of the stall
                // samples in profiles/r02_step9_ncu_full_sass_nb_likelihood_fast.txt); beyond the threshold it rounds to 1
                a.ddpre[prev_row] = s * a.inv_B * -expm1f(-prev_depth);
                a.dznu[prev_row] = dz;
            }
        }
        if (!active) break;

        // (the sweep wrote the ring slot through the generic proxy; the TMA unit refills it later through the async one)
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();    // A: g / p of this row complete (the fix-up overwrites entries); every thread is done with the
                            //    previous row's output, so its ring slot can take the row after next
        if (tid == 0 && nj > 0) {
            const int r2 = row + 2 * n_clusters;
            const int s2 = slot == 0 ? 2 : slot - 1;     // (it + 2) % 3
            if (r2 < a.B) bulk_load(smem_addr(s_land + s2 * Dp), a.G + (size_t)r2 * D + j0, row_bytes, lbar0 + 8 * s2);
        }

        // ---- this cell's nonzeros that fall in the slice (prefetched; overflow from global)
        float red[2] = { 0.f, 0.f };
        auto fix = [&](int j, float x) {
            const int jj = j - j0;
            if (jj < 0 || jj >= nj || x == 0.f) return;
            const float p = sL[jj];
            const float wn = s_wn[jj];
            const elem_t e = elem_fast(p, depth, fmaf(zn, wn, s_ub[jj]));
            const float L = __logf(e.m + e.nu);
            float lg_terms, dig_diff;
            count_terms(e.nu, x, lg_terms, dig_diff);
            nll_acc += lg_terms + x * (L - __logf(e.m));   // nb.hh:522-528 minus the x = 0 value already counted
            if (grad) {
                const float gx = (e.nu + x) * e.inv_mnu - x * rcp_f(e.m);
                red[0] = fmaf(gx - e.nu * e.inv_mnu, p, red[0]);
                sg[jj] = gx;
                const float ddu = (dig_diff + x * e.inv_mnu) * e.sig_mask;   // dnu_x - dnu_0 (un-normalised)
                atomicAdd(a.col_du + j, ddu * a.inv_B);
                atomicAdd(a.col_duz + j, ddu * zn * a.inv_B);
                red[1] = fmaf(ddu, wn, red[1]);
            }
        };
#pragma unroll
        for (int i = 0; i < NZP; ++i)
            if (pj[i] >= 0) fix(pj[i], px[i]);
        for (int64_t q = q_lo + tid + NZP * kT; q < q_hi; q += kT) fix(a.indices[q], a.vals[q]);
        if (row + n_clusters < a.B) {          // prefetch the next row's nonzeros (its range arrived long ago)
#pragma unroll
            for (int i = 0; i < NZP; ++i) {
                const int64_t q = qlo_n + tid + i * kT;
                pj[i] = q < qhi_n ? a.indices[q] : -1;
                px[i] = q < qhi_n ? a.vals[q] : 0.f;
            }
        }

        if (grad) {
            // ---- the row reduction: s = sum g p, d znu; warp 0 finishes it and sends the hop, nobody waits
            {
                float a0, a1, b0, b1;
                up2(red0, a0, a1);
                up2(red1, b0, b1);
                red[0] = warp_sum(red[0] + a0 + a1);
                red[1] = warp_sum(red[1] + b0 + b1);
            }
            if (lane == 0) { ws[warp * 2 + 0] = red[0]; ws[warp * 2 + 1] = red[1]; }
            // B: partials in.  Only warp 0 (which adds them up and sends the hop) blocks; the others arrive and go on to
            // the next row.  (Hardware barriers on purpose: warps spinning on an mbarrier inside the CTA take issue slots
            // from the CTA that shares the SM — measured, 0.52 -> 0.58 ms with mbarrier arrive / wait in their place;
            // this arrive / sync split: 0.527 -> 0.520 ms, profiles/r02_step3_k4_ab.txt.)
            // The output pass sees every fix of g through the hop: our share of it leaves after this barrier completes.
            if (warp != 0) {
                __syncwarp();
                asm volatile("bar.arrive 1, %0;" ::"r"(kT) : "memory");
            } else {
                asm volatile("bar.sync 1, %0;" ::"r"(kT) : "memory");
                float s = lane < NW ? ws[lane * 2 + 0] : 0.f, dz = lane < NW ? ws[lane * 2 + 1] : 0.f;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    s += __shfl_xor_sync(0xffffffffu, s, o);
                    dz += __shfl_xor_sync(0xffffffffu, dz, o);
                }
                // red[1] was accumulated against wn * log2(e) and without the 1 / B
                dz *= kLn2 * a.inv_B;
                if (lane == 0) {
                    const uint32_t xb = xbar0 + 8 * par;
                    if (cs > 1) {
                        xchg_send(smem_addr(xslot + par * 16 + rank * 2), xb, cs, s, dz);
                    } else {
                        xslot[par * 16 + 0] = s; xslot[par * 16 + 1] = dz;
                        asm volatile("mbarrier.arrive.release.cta.shared::cta.b64 _, [%0];" ::"r"(xb) : "memory");
                    }
                }
            }
        } else {
            __syncthreads();
        }
        prev_row = row; prev_slot = slot; prev_depth = depth; prev_coef = coef;
        slot = slot == 2 ? 0 : slot + 1;
        if (a.row_lse && rank == 0 && tid == 0) a.row_lse[row] = gm2 * kLn2 - logf(invS);
    }

    {
        float n0, n1;
        up2(nll2, n0, n1);
        nll_acc = warp_sum(nll_acc + n0 + n1);
    }
    if (lane == 0 && nll_acc != 0.f) atomicAdd(a.loss_acc, (double)nll_acc);
    if (grad) {
#pragma unroll
        for (int h = 0; h < HP; ++h) {
            const int jj = 2 * tid + h * 2 * kT;
            if (jj < nj) {
                float c0, c1, z0, z1;
                up2(cdu[h], c0, c1);
                up2(cduz[h], z0, z1);
                atomicAdd(a.col_du + j0 + jj, c0 * a.inv_B);
                atomicAdd(a.col_du + j0 + jj + 1, c1 * a.inv_B);
                atomicAdd(a.col_duz + j0 + jj, z0 * a.inv_B);
                atomicAdd(a.col_duz + j0 + jj + 1, z1 * a.inv_B);
            }
        }
    }
    if (cs > 1) cluster.sync();   // no CTA exits while a peer may still signal its barriers
}

size_t
fast_smem_bytes(int ept, int T)
{
    return sizeof(float) * ((size_t)7 * ept * T + 2 * (T / 32) * 2 + 32) + 5 * 8 + 16;
}

// One instantiation of the tuned kernel and the slice layout it gives D genes.  The kernel checks the padding only in
// the last two pair slots of a thread, so the shortest slice (the last rank's) must reach into them.
// Measured at cfg 4 (D = 30 000, 8 CTAs of 3 752 genes): <16, 256> 0.50 ms, <10, 384> 0.62 ms, <8, 512> 0.80 ms (the
// last two spill) — fewer, fatter threads win, as in round 1.
struct k4_variant_t {
    int ept, T;
    void (*kern)(nb_lik_args_t, int, int);
    void (*kern_lo)(nb_lik_args_t, int, int);
};
const k4_variant_t kVariants[] = {
    { 4, 256, nb_likelihood_fast_kernel<4, 256>, nb_likelihood_fast_kernel<4, 256, true> },
    { 8, 256, nb_likelihood_fast_kernel<8, 256>, nb_likelihood_fast_kernel<8, 256, true> },
    { 12, 256, nb_likelihood_fast_kernel<12, 256>, nb_likelihood_fast_kernel<12, 256, true> },
    { 16, 256, nb_likelihood_fast_kernel<16, 256>, nb_likelihood_fast_kernel<16, 256, true> },
};

bool
variant_fits(const k4_variant_t &v, int D, int &cs, int &Ds)
{
    const size_t budget = 113 * 1024;   // two CTAs per SM
    if (fast_smem_bytes(v.ept, v.T) > budget) return false;
    for (cs = 1; cs <= 8; cs *= 2) {
        Ds = ((D + cs - 1) / cs + 3) & ~3;
        if (Ds > v.ept * v.T) continue;
        const int nj_last = D - (cs - 1) * Ds;
        if (nj_last <= 0) return false;
        if (v.ept > 4 && nj_last <= (v.ept / 2 - 2) * 2 * v.T) return false;   // a larger class (more padding) has to take it
        return true;
    }
    return false;
}

const k4_variant_t *
fast_layout(int D, int &cs, int &Ds)
{
    for (const k4_variant_t &v : kVariants)
        if (variant_fits(v, D, cs, Ds)) return &v;
    return nullptr;
}

} // namespace

int
launch_slice_bounds(const int64_t *indptr, const int32_t *indices, const int32_t *cells, const int *unsorted, int B, int cs,
                    int Ds, int64_t *bounds, cudaStream_t stream)
{
    slice_bounds_kernel<<<ceil_div((int64_t)B * (cs + 1), 256), 256, 0, stream>>>(indptr, indices, cells, unsorted, B, cs, Ds, bounds);
    ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_nb_row_meta(const float *G, const void *tile_stats, int n_tiles, const float *dpre, const float *znu, int B, int D,
                   void *meta, cudaStream_t stream)
{
    if (tile_stats)
        nb_row_meta_merge_kernel<<<ceil_div(B, 128), 128, 0, stream>>>(reinterpret_cast<const float2 *>(tile_stats), n_tiles, dpre,
                                                                      znu, B, reinterpret_cast<float4 *>(meta));
    else
        nb_row_meta_scan_kernel<<<B, 256, 0, stream>>>(G, D, dpre, znu, reinterpret_cast<float4 *>(meta));
    ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

// the tuned kernel reads COMPLETE logits (the caller had the per-gene bias added by the producing GEMM) and the
// per-row metadata of launch_nb_row_meta
bool
nb_likelihood_fast_eligible(const nb_lik_args_t &a)
{
    int cs, Ds;
    return a.Rnu == 1 && a.cov == nullptr && a.D % 4 == 0 && a.recon_mu == nullptr && a.recon_nu == nullptr &&
        a.slice_bounds != nullptr && (reinterpret_cast<uintptr_t>(a.G) & 15) == 0 &&
        (a.dlogits == nullptr || (reinterpret_cast<uintptr_t>(a.dlogits) & 7) == 0) && fast_layout(a.D, cs, Ds) != nullptr;
}

int
launch_nb_likelihood_fast(const nb_lik_args_t &a, int num_sms, cudaStream_t stream)
{
    MMV_CHECK(a.rowmeta != nullptr && a.abias == nullptr,
              "nb_likelihood (tuned): needs the row metadata and logits with the bias already added");
    int cs = 1, Ds = 0;
    const k4_variant_t *v = fast_layout(a.D, cs, Ds);
    MMV_CHECK(v != nullptr, "nb_likelihood: D=%d does not fit a cluster of 8 CTAs", a.D);
    if (a.tile_stats) {   // row metadata from the GEMM epilogue's partials + slice bounds: one pre-kernel
        const int meta_blocks = ceil_div(a.B, 256);
        MMV_CUDA(launch_k(nb_row_prep_kernel, dim3(meta_blocks + ceil_div((int64_t)a.B * (cs + 1), 256)), dim3(256), 0, stream,
                          reinterpret_cast<const float2 *>(a.tile_stats), a.n_tiles, a.dpre, a.znu,
                          reinterpret_cast<float4 *>(const_cast<void *>(a.rowmeta)), meta_blocks, a.indptr, a.indices, a.cells,
                          a.unsorted, a.B, cs, Ds, a.slice_bounds));
        ::mmv::count_launch();
    } else {              // (the caller ran launch_nb_row_meta's scan over the complete logits)
        MMV_TRY(launch_slice_bounds(a.indptr, a.indices, a.cells, a.unsorted, a.B, cs, Ds, a.slice_bounds, stream));
    }
    MMV_CHECK(!a.dlogits_lo || a.dlogits_bf16, "nb_likelihood (tuned): the residual gradient goes with the bf16 one");
    void (*kern)(nb_lik_args_t, int, int) = a.dlogits_lo ? v->kern_lo : v->kern;
    const size_t smem = fast_smem_bytes(v->ept, v->T);
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(kern), (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(num_sms / cs * cs);
    cfg.blockDim = dim3(v->T);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    // (no programmatic dependent launch here, unlike the kernels around it: launched early, while the pre-kernel's CTAs
    // still hold SMs, the 33 persistent clusters no longer all become resident at once and the kernel takes 0.89 ms
    // instead of 0.50 — profiles/r02_step6_pdl_ab.txt)
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // persistent clusters: exactly one wave (a second wave would double the kernel time)
    int n_clusters = 2 * num_sms / cs;
    if (cs > 1) {
        int fit = 0;
        MMV_CUDA(cudaOccupancyMaxActiveClusters(&fit, kern, &cfg));
        if (fit > 0 && fit < n_clusters) n_clusters = fit;
    }
    if (n_clusters > a.B) n_clusters = a.B;
    if (n_clusters < 1) n_clusters = 1;
    cfg.gridDim = dim3(n_clusters * cs);
    MMV_CUDA(cudaLaunchKernelEx(&cfg, kern, a, cs, Ds));
    ::mmv::count_launch();
    return 0;
}

} // namespace mmv
