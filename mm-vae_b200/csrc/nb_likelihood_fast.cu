// K4, tuned path (default overdispersion latent Rnu = 1, all-ones covariate) — same math
// and cluster mapping as nb_likelihood.cu, reorganised after the first ncu capture
// (profiles/r01_step1_*: 41 % long-scoreboard stalls on one in-flight LDG per thread,
// 16 warps/SM, ~180 libdevice instructions per element):
//   * the row slice of the decoder output is brought in by the TMA unit
//     (cp.async.bulk 1D + mbarrier tx-count), double buffered: row r+1 streams from HBM
//     while row r is being processed, so no warp ever waits on a global load;
//   * the per-gene constants of the slice (bias, nu bias, nu weight) are staged in shared
//     memory once per CTA instead of being re-read from L2 for every row;
//   * p, g and the per-gene column sums of the nu gradient live in REGISTERS
//     (each thread owns genes tid, tid+512, ... of the slice; EPT is a template constant);
//   * MUFU ex2/lg2/rcp with FMA-pipe polynomials where the MUFU form loses accuracy
//     (log1p of small arguments); 7 MUFU per element;
//   * the nonzeros are handled by a compact second loop (all lanes busy) that corrects
//     the row sums, sends its column-sum deltas straight to red.global.add and, after the
//     dense store, patches its dL/dlogit entries in place.
#include "cluster_reduce.cuh"
#include "kernels.hh"

#include <cuda_bf16.h>

namespace mmv {

namespace {

constexpr int kT = 512;
constexpr float kEps = 1e-4f, kNuMin = 1e-4f, kNuMax = 1e4f;   // nb.hh:514,252,459

__device__ __forceinline__ float
rcp_f(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// log(1 + y), y >= 0: Taylor polynomial on the FMA pipe below 1/16, MUFU lg2 above
__device__ __forceinline__ float
log1p_fast(float y)
{
    const float poly = y * (1.f - y * (0.5f - y * (0.33333334f - y * (0.25f - y * (0.2f - y * 0.16666667f)))));
    const float big = __logf(1.f + y);
    return y < 0.0625f ? poly : big;
}

struct elem_t {
    float m, nu, l1p, inv_mnu, sig_mask;
};

__device__ __forceinline__ elem_t
elem_fast(float p, float depth, float u)
{
    elem_t e;
    e.m = fmaf(p, depth, kEps);                                  // nb.hh:519
    const float uc = fminf(u, 20.f);
    const float t = __expf(uc);
    const float sp = u > 20.f ? u : log1p_fast(t);               // softplus, threshold 20 (nb.hh:458)
    const float r = fminf(fmaxf(sp, kNuMin), kNuMax);            // nb.hh:459
    e.nu = r + kEps;                                             // nb.hh:518
    const float sig = u > 20.f ? 1.f : 1.f - rcp_f(1.f + t);     // d softplus / du
    e.sig_mask = (sp >= kNuMin && sp <= kNuMax) ? sig : 0.f;     // clamp passes the gradient inside
    e.l1p = log1p_fast(e.m * rcp_f(e.nu));                       // log(m + nu) - log(nu)
    e.inv_mnu = rcp_f(e.m + e.nu);
    return e;
}

__device__ __forceinline__ uint32_t
smem_addr(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}

__device__ __forceinline__ void
bulk_load(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

__device__ __forceinline__ void
bar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "NBW_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra NBW_DONE;\n"
        "bra NBW_LOOP;\n"
        "NBW_DONE:\n"
        "}\n" ::"r"(bar), "r"(parity) : "memory");
}

// lgamma(nu) - lgamma(nu + x) + lgamma(x + 1) and digamma(nu) - digamma(nu + x) for a count x.
// Small integer counts (the bulk of single-cell data) use the exact product forms
//   lgamma(nu + x) - lgamma(nu) = log prod_{k<x} (nu + k),  digamma(nu + x) - digamma(nu) = P'(nu) / P(nu)
// on the FMA pipe with one lg2 and one rcp; anything else falls back to libdevice.
__constant__ float kLogFact[9] = { 0.f, 0.f, 0.6931471806f, 1.7917594692f, 3.1780538303f, 4.7874917428f,
                                   6.5792512120f, 8.5251613611f, 10.6046029027f };

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
    } else {
        lg_terms = lgammaf(nu) + lgammaf(x + 1.f) - lgammaf(nu + x);
        dig_diff = digamma_f(nu) - digamma_f(nu + x);
    }
}

template <int EPT>
__global__ void __launch_bounds__(kT, 1)
nb_likelihood_fast_kernel(nb_lik_args_t a, int cs, int Ds)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = cs > 1 ? (int)cluster.block_rank() : 0;
    const int cluster_id = blockIdx.x / cs, n_clusters = gridDim.x / cs;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, D = a.D;
    constexpr int NW = kT / 32;
    const int j0 = rank * Ds, nj = max(0, min(D - j0, Ds));
    const bool grad = a.dlogits != nullptr || a.dlogits_bf16 != nullptr;
    const bool sorted = a.unsorted != nullptr && *a.unsorted == 0;

    extern __shared__ __align__(16) float smem[];
    float *sG0 = smem, *sG1 = sG0 + Ds;          // TMA landing buffers (raw decoder output)
    float *s_ab = sG1 + Ds;                      // bias: mu_decoding.bias + covar_decoding.bias + mu_bias + sum_c Wcd
    float *s_ub = s_ab + Ds;                     // nu_decoding.bias - nu_bias
    float *s_wn = s_ub + Ds;                     // nu_decoding.weight[:, 0]
    float *scratch = s_wn + Ds;                  // [2 * 32]
    float *slots = scratch + 64;                 // [2][2 + 2] cluster exchange, double buffered by row parity
    float *wstat = slots + 8;                    // [2][NW][4] per-warp (max, sumexp, n_lo, n_hi)
    uint64_t *bars = reinterpret_cast<uint64_t *>(wstat + 2 * NW * 4);   // [2], 8-byte aligned (Ds % 4 == 0)
    const uint32_t bar0 = smem_addr(bars), bar1 = bar0 + 8;

    if (tid == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar1));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int jj = tid; jj < nj; jj += kT) {
        const int j = j0 + jj;
        float ab = a.abias[j];
        for (int c = 0; c < a.C; ++c) ab += a.Wcd[(size_t)j * a.C + c];   // covariates are all ones on this path
        s_ab[jj] = ab;
        s_ub[jj] = a.ubias[j];
        s_wn[jj] = a.Wnd[j];
    }
    __syncthreads();

    const uint32_t row_bytes = (uint32_t)nj * 4u;
    int row = cluster_id;
    if (tid == 0 && row < a.B && nj > 0) bulk_load(smem_addr(sG0), a.G + (size_t)row * D + j0, row_bytes, bar0);

    float cdu[EPT], cduz[EPT];
#pragma unroll
    for (int k = 0; k < EPT; ++k) cdu[k] = cduz[k] = 0.f;
    float nll_acc = 0.f;

    for (int it = 0; row < a.B; row += n_clusters, ++it) {
        const int cur = it & 1;
        const float *sG = cur ? sG1 : sG0;
        float *slot = slots + cur * 4;
        float *ws = wstat + cur * NW * 4;
        const int next = row + n_clusters;
        if (tid == 0 && next < a.B && nj > 0)   // stream the next row while this one is processed
            bulk_load(smem_addr(cur ? sG0 : sG1), a.G + (size_t)next * D + j0, row_bytes, cur ? bar0 : bar1);

        const int64_t cell = a.cells ? (int64_t)a.cells[row] : (int64_t)row;
        const float dp = a.dpre[row];
        const float depth = dp > 20.f ? dp : log1pf(expf(dp));       // nb.hh:498 (once per row: exact)
        const float zn = a.znu[row];
        const int64_t nz_beg = a.indptr[cell], nz_end = a.indptr[cell + 1];

        // position of this CTA's gene slice inside the (sorted) nonzeros of the cell
        int n_lo = 0, n_hi = 0;
        if (sorted)
            for (int64_t q = nz_beg + tid; q < nz_end; q += kT) {
                const int j = a.indices[q];
                n_lo += j < j0;
                n_hi += j < j0 + nj;
            }

        if (nj > 0) bar_wait(cur ? bar1 : bar0, (uint32_t)(it >> 1) & 1u);

        // ---- pass 1: logits into registers; softmax numerators against the WARP max, so one
        //      block hop + one cluster hop give max and sum together (online-softmax merge)
        float v[EPT];
        float wm = -INFINITY;
#pragma unroll
        for (int k = 0; k < EPT; ++k) {
            const int jj = tid + k * kT;
            v[k] = jj < nj ? sG[jj] + s_ab[jj] : -INFINITY;
            wm = fmaxf(wm, v[k]);
        }
        wm = warp_max(wm);
        float wsum = 0.f;
        if (wm > -INFINITY) {
#pragma unroll
            for (int k = 0; k < EPT; ++k) {
                v[k] = __expf(v[k] - wm);          // exp(-inf) = 0 for the padding lanes
                wsum += v[k];
            }
        } else {
#pragma unroll
            for (int k = 0; k < EPT; ++k) v[k] = 0.f;
        }
        wsum = warp_sum(wsum);
        n_lo = __reduce_add_sync(0xffffffffu, n_lo);
        n_hi = __reduce_add_sync(0xffffffffu, n_hi);
        if (lane == 0) {
            ws[warp * 4 + 0] = wm;
            ws[warp * 4 + 1] = wsum;
            ws[warp * 4 + 2] = __int_as_float(n_lo);
            ws[warp * 4 + 3] = __int_as_float(n_hi);
        }
        __syncthreads();
        float bm = lane < NW ? ws[lane * 4 + 0] : -INFINITY;
        const float bs_in = lane < NW ? ws[lane * 4 + 1] : 0.f;
        int c_lo = lane < NW ? __float_as_int(ws[lane * 4 + 2]) : 0;
        int c_hi = lane < NW ? __float_as_int(ws[lane * 4 + 3]) : 0;
        const float bmax = warp_max(bm);
        float bsum = warp_sum(bm > -INFINITY ? bs_in * __expf(bm - bmax) : 0.f);
        c_lo = __reduce_add_sync(0xffffffffu, c_lo);
        c_hi = __reduce_add_sync(0xffffffffu, c_hi);
        float gmax = bmax, gsum = bsum;
        if (cs > 1) {
            if (tid == 0) { slot[0] = bmax; slot[1] = bsum; }
            cluster.sync();
            gmax = -INFINITY;
            float rm[8], rs[8];
#pragma unroll
            for (int r = 0; r < 8; ++r) {
                rm[r] = -INFINITY; rs[r] = 0.f;
                if (r < cs) {
                    const float *remote = cluster.map_shared_rank(slot, r);
                    rm[r] = remote[0]; rs[r] = remote[1];
                }
                gmax = fmaxf(gmax, rm[r]);
            }
            gsum = 0.f;
#pragma unroll
            for (int r = 0; r < 8; ++r) gsum += rm[r] > -INFINITY ? rs[r] * __expf(rm[r] - gmax) : 0.f;
        }
        const float inv_sum = 1.f / gsum;
        const float scale = wm > -INFINITY ? __expf(wm - gmax) * inv_sum : 0.f;
        if (a.row_lse && rank == 0 && tid == 0) a.row_lse[row] = gmax + logf(gsum);
        const int64_t q_lo = sorted ? nz_beg + c_lo : nz_beg, q_hi = sorted ? nz_beg + c_hi : nz_end;

        // ---- pass 2: dense sweep assuming x = 0 (lgammas cancel, no digamma)
        float g[EPT];
        float red[2] = { 0.f, 0.f };           // s = sum g p ; d znu
#pragma unroll
        for (int k = 0; k < EPT; ++k) {
            const int jj = tid + k * kT;
            if (jj < nj) {
                const float p = v[k] * scale;
                const float wn = s_wn[jj];
                const elem_t e = elem_fast(p, depth, fmaf(zn, wn, s_ub[jj]));
                nll_acc = fmaf(e.nu, e.l1p, nll_acc);
                v[k] = p;
                g[k] = e.nu * e.inv_mnu;
                red[0] = fmaf(g[k], p, red[0]);
                const float du = (e.l1p - e.m * e.inv_mnu) * e.sig_mask * a.inv_B;
                cdu[k] += du;
                cduz[k] = fmaf(du, zn, cduz[k]);
                red[1] = fmaf(du, wn, red[1]);
            } else {
                g[k] = 0.f;
            }
        }

        // ---- pass 2b: compact loop over this cell's nonzeros that fall in the slice
        for (int64_t q = q_lo + tid; q < q_hi; q += kT) {
            const int j = a.indices[q];
            const int jj = j - j0;
            const float x = a.vals[q];
            if (jj < 0 || jj >= nj || x == 0.f) continue;
            const float p = __expf(sG[jj] + s_ab[jj] - gmax) * inv_sum;
            const float wn = s_wn[jj];
            const elem_t e = elem_fast(p, depth, fmaf(zn, wn, s_ub[jj]));
            const float L = __logf(e.m + e.nu);
            float lg_terms, dig_diff;
            count_terms(e.nu, x, lg_terms, dig_diff);
            // nb.hh:522-528 minus the x = 0 value already counted
            nll_acc += lg_terms + x * (L - __logf(e.m));
            if (grad) {
                const float gx = (e.nu + x) * e.inv_mnu - x * rcp_f(e.m);
                red[0] = fmaf(gx - e.nu * e.inv_mnu, p, red[0]);
                const float ddnu = dig_diff + x * e.inv_mnu;   // dnu_x - dnu_0
                const float ddu = ddnu * e.sig_mask * a.inv_B;
                atomicAdd(a.col_du + j, ddu);
                atomicAdd(a.col_duz + j, ddu * zn);
                red[1] = fmaf(ddu, wn, red[1]);
            }
        }

        if (grad) {
            cluster_sum<2>(red, slot + 2, scratch, cluster, cs);
            const float s = red[0];
            const float coef = depth * a.inv_B;
            // ---- pass 3: dL/dlogit_j = (d_i / B) p_j (g_j - s_i), coalesced
            if (a.dlogits_bf16) {
                __nv_bfloat16 *drow = reinterpret_cast<__nv_bfloat16 *>(a.dlogits_bf16) + (size_t)row * D + j0;
#pragma unroll
                for (int k = 0; k < EPT; ++k) {
                    const int jj = tid + k * kT;
                    if (jj < nj) drow[jj] = __float2bfloat16(coef * v[k] * (g[k] - s));
                }
            } else {
                float *drow = a.dlogits + (size_t)row * D + j0;
#pragma unroll
                for (int k = 0; k < EPT; ++k) {
                    const int jj = tid + k * kT;
                    if (jj < nj) drow[jj] = coef * v[k] * (g[k] - s);
                }
            }
            if (rank == 0 && tid == 0) {
                a.ddpre[row] = s * a.inv_B * (dp > 20.f ? 1.f : 1.f / (1.f + expf(-dp)));
                a.dznu[row] = red[1];
            }
            __syncthreads();   // dense stores of this block are visible before the patches below
            // ---- pass 3b: patch the entries of the nonzeros
            for (int64_t q = q_lo + tid; q < q_hi; q += kT) {
                const int j = a.indices[q];
                const int jj = j - j0;
                const float x = a.vals[q];
                if (jj < 0 || jj >= nj || x == 0.f) continue;
                const float p = __expf(sG[jj] + s_ab[jj] - gmax) * inv_sum;
                const elem_t e = elem_fast(p, depth, fmaf(zn, s_wn[jj], s_ub[jj]));
                const float gx = (e.nu + x) * e.inv_mnu - x * rcp_f(e.m);
                const float dl = coef * p * (gx - s);
                if (a.dlogits_bf16)
                    reinterpret_cast<__nv_bfloat16 *>(a.dlogits_bf16)[(size_t)row * D + j] = __float2bfloat16(dl);
                else
                    a.dlogits[(size_t)row * D + j] = dl;
            }
        }
        __syncthreads();       // everyone is done with sG[cur] before the TMA unit refills it
    }

    float nl[1] = { nll_acc };
    block_sum<1>(nl, scratch);
    if (tid == 0 && nl[0] != 0.f) atomicAdd(a.loss_acc, (double)nl[0]);
    if (grad) {
#pragma unroll
        for (int k = 0; k < EPT; ++k) {
            const int jj = tid + k * kT;
            if (jj < nj) {
                atomicAdd(a.col_du + j0 + jj, cdu[k]);
                atomicAdd(a.col_duz + j0 + jj, cduz[k]);
            }
        }
    }
    if (cs > 1) cluster.sync();
}

size_t
fast_smem_bytes(int Ds)
{
    return sizeof(float) * ((size_t)5 * Ds + 64 + 8 + 2 * (kT / 32) * 4) + 16;
}

} // namespace

bool
nb_likelihood_fast_eligible(const nb_lik_args_t &a)
{
    return a.Rnu == 1 && a.cov == nullptr && a.D % 4 == 0 && a.recon_mu == nullptr && a.recon_nu == nullptr &&
        (reinterpret_cast<uintptr_t>(a.G) & 15) == 0;
}

int
launch_nb_likelihood_fast(const nb_lik_args_t &a, int num_sms, cudaStream_t stream)
{
    const size_t budget = 200 * 1024;
    int cs = 1, Ds = 0, ept = 0;
    for (;; cs *= 2) {
        Ds = ((a.D + cs - 1) / cs + 3) & ~3;
        ept = (Ds + kT - 1) / kT;
        if (fast_smem_bytes(Ds) <= budget && ept <= 16) break;
        MMV_CHECK(cs < 8, "nb_likelihood: D=%d does not fit a cluster of 8 CTAs", a.D);
    }
    void (*kern)(nb_lik_args_t, int, int) = ept <= 4 ? nb_likelihood_fast_kernel<4>
        : ept <= 8 ? nb_likelihood_fast_kernel<8> : ept <= 12 ? nb_likelihood_fast_kernel<12> : nb_likelihood_fast_kernel<16>;
    const size_t smem = fast_smem_bytes(Ds);
    MMV_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(num_sms / cs * cs);
    cfg.blockDim = dim3(kT);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    // persistent clusters: exactly one wave.  A GPC holds fewer whole clusters than SMs / cs
    // (148 SMs take 33 clusters of 4, not 37), and a second wave would double the kernel time.
    int n_clusters = num_sms / cs;
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
