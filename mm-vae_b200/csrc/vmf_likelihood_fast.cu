// K5, tuned path (all-ones covariate) — same math as vmf_likelihood.cu (vmf.hh:283-290,419-440 +
// autograd), organised like K4's tuned kernel (nb_likelihood_fast.cu):
//   * a thread-block cluster owns whole rows, each CTA a fixed gene slice for the entire kernel: the
//     per-gene constants (decoding bias pre-scaled by log2 e, covariate term) staged in shared memory
//     once, the per-gene column sums of dL/dq in registers; three CTAs share an SM;
//   * rows arrive through a double-buffered cp.async.bulk + mbarrier ring;
//   * ONE cluster hop per row carries the four row sums (sum q^2, sum q, sum_nnz l q, ||yobs||^2 part)
//     with st.async + a remote mbarrier complete_tx;
//   * the dense sweep (x = 0 for 95 % of the entries: yobs_ij = e / Nobs_i) runs on packed fp32 pairs, one
//     MUFU per element, and stores dL/dpre straight to HBM (bf16x2 or float2); the cell's nonzeros inside
//     the slice (bisected by the slice-bounds pre-kernel, prefetched one row ahead) then overwrite their
//     own entries and correct the column sums.
// 6 bytes per element (fp32 logit in, bf16 gradient out): the kernel is bound by HBM.
#include "common.cuh"
#include "kernels.hh"
#include "rowstream.cuh"

#include <cooperative_groups.h>
#include <cuda_bf16.h>

namespace mmv {

namespace cg = cooperative_groups;

namespace {

using namespace rowstream;

constexpr int kT = 256;
constexpr int NW = kT / 32;
constexpr int NZP = 2;   // nonzeros prefetched per thread per row

template <int EPT>
__global__ void __launch_bounds__(kT, 3)
vmf_likelihood_fast_kernel(vmf_lik_args_t a, int cs, int Ds)
{
    static_assert(EPT % 2 == 0, "threads own adjacent gene pairs");
    constexpr int HP = EPT / 2;
    constexpr int Dp = EPT * kT;
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = cs > 1 ? (int)cluster.block_rank() : 0;
    const int cluster_id = blockIdx.x / cs, n_clusters = gridDim.x / cs;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, D = a.D, C = a.C;
    const int j0 = rank * Ds, nj = max(0, min(D - j0, Ds));
    const bool grad = a.dpre != nullptr || a.dpre_bf16 != nullptr;
    const bool sorted = a.unsorted != nullptr && *a.unsorted == 0;
    const float e_obs = a.eps_obs;

    extern __shared__ __align__(16) float smem[];
    float *sG0 = smem, *sG1 = sG0 + Dp;          // TMA landing buffers (last decoder layer output, no bias)
    float *s_b = sG1 + Dp;                       // decoding.bias * log2(e)   (padding: -inf -> exp = 0)
    float *s_cv = s_b + Dp;                      // covar_decoding_ bias + sum_c weight (covariates are all ones)
    float *wstat = s_cv + Dp;                    // [2 parity][NW][4] per-warp partials
    float *xslot = wstat + 2 * NW * 4;           // [2 parity][8 ranks][4]
    uint64_t *bars = reinterpret_cast<uint64_t *>(xslot + 64);   // [2] row ring + [2 parity] exchange
    const uint32_t bar0 = smem_addr(bars), bar1 = bar0 + 8, xbar0 = bar0 + 16;

    if (tid == 0) {
        for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar0 + 8 * i));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int jj = tid; jj < Dp; jj += kT) {
        const int j = j0 + jj;
        float b = -INFINITY, cv = 0.f;
        if (jj < nj) {
            b = a.blast[j] * kLog2e;
            cv = a.bcd[j];
            for (int c = 0; c < C; ++c) cv += a.Wcd[(size_t)j * C + c];
        }
        s_b[jj] = b;
        s_cv[jj] = cv;
        if (jj >= nj) sG0[jj] = sG1[jj] = 0.f;   // the padding of the landing buffers is never written by TMA
    }
    if (cs > 1) cluster.sync(); else __syncthreads();

    const uint32_t row_bytes = (uint32_t)nj * 4u;
    int row = cluster_id;
    if (tid == 0 && row < a.B && nj > 0) bulk_load(smem_addr(sG0), a.G + (size_t)row * D + j0, row_bytes, bar0);

    f2_t cdq[HP];                                // per-gene column sums of dL/dq over this cluster's rows
#pragma unroll
    for (int h = 0; h < HP; ++h) cdq[h] = 0ull;
    const float kappa = *a.kappa;
    float t_acc = 0.f;

    const int bstride = cs + 1;
    int64_t qlo_n = 0, qhi_n = 0, qlo_p = 0, qhi_p = 0;   // CSR ranges: row about to be processed / the one after it
    int pj[NZP];
    float px[NZP];
    if (row < a.B) {
        qlo_n = a.slice_bounds[(size_t)row * bstride + (sorted ? rank : 0)];
        qhi_n = a.slice_bounds[(size_t)row * bstride + (sorted ? rank + 1 : cs)];
#pragma unroll
        for (int i = 0; i < NZP; ++i) {
            const int64_t q = qlo_n + tid + i * kT;
            pj[i] = q < qhi_n ? a.indices[q] : -1;
            px[i] = q < qhi_n ? a.vals[q] : 0.f;
        }
        if (row + n_clusters < a.B) {
            qlo_p = a.slice_bounds[(size_t)(row + n_clusters) * bstride + (sorted ? rank : 0)];
            qhi_p = a.slice_bounds[(size_t)(row + n_clusters) * bstride + (sorted ? rank + 1 : cs)];
        }
    }

    for (int it = 0; row < a.B; row += n_clusters, ++it) {
        const int cur = it & 1;
        const uint32_t ph = (uint32_t)(it >> 1) & 1u;
        const float *sG = cur ? sG1 : sG0;
        float *ws = wstat + cur * NW * 4;
        float *xs = xslot + cur * 32;
        const uint32_t xb = xbar0 + cur * 8;
        const int next = row + n_clusters;
        if (tid == 0 && next < a.B && nj > 0)   // stream the next row while this one is processed
            bulk_load(smem_addr(cur ? sG0 : sG1), a.G + (size_t)next * D + j0, row_bytes, cur ? bar0 : bar1);
        const int64_t q_lo = qlo_n, q_hi = qhi_n;
        qlo_n = qlo_p; qhi_n = qhi_p;          // the next row's CSR range was read one iteration ago
        // pull the next row's nonzeros into L2 now, a whole iteration before the register prefetch at the bottom
        // asks for them: with a shard of 10^9 nonzeros every run is a cold HBM + TLB miss otherwise
        // (K5 0.51 ms at 1 M resident cells vs 0.34 ms at 16 k)
        if (tid < 128) {
            const int64_t q = qlo_n + (int64_t)(tid & 63) * 32;
            if (q < qhi_n) {
                const void *p = (tid >> 6) ? (const void *)(a.vals + q) : (const void *)(a.indices + q);
                asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
            }
        }
        qlo_p = qhi_p = 0;
        if (next + n_clusters < a.B) {
            qlo_p = a.slice_bounds[(size_t)(next + n_clusters) * bstride + (sorted ? rank : 0)];
            qhi_p = a.slice_bounds[(size_t)(next + n_clusters) * bstride + (sorted ? rank + 1 : cs)];
        }
        if (nj > 0) bar_wait(cur ? bar1 : bar0, ph);

        // ---- pass 1: q = exp(pre + bias) + covariate term (vmf.hh:287-289) into registers; sum q^2, sum q
        f2_t qp[HP];
        f2_t r0 = 0ull, r1 = 0ull;
        {
            const f2_t l2e = bc2(kLog2e);
#pragma unroll
            for (int h = 0; h < HP; ++h) {
                const int jj = 2 * tid + h * 2 * kT;
                float y0, y1;
                up2(fma2(*reinterpret_cast<const f2_t *>(sG + jj), l2e, *reinterpret_cast<const f2_t *>(s_b + jj)), y0, y1);
                qp[h] = add2(pk2(ex2_f(y0), ex2_f(y1)), *reinterpret_cast<const f2_t *>(s_cv + jj));
                r0 = fma2(qp[h], qp[h], r0);
                r1 = add2(r1, qp[h]);
            }
        }
        float red[4];
        {
            float a0, a1, b0, b1;
            up2(r0, a0, a1);
            up2(r1, b0, b1);
            red[0] = a0 + a1; red[1] = b0 + b1; red[2] = 0.f; red[3] = 0.f;
        }
        // ---- the cell's nonzeros inside the slice: sparse part of <yobs, q> and of ||yobs||^2 (vmf.hh:424-426)
        float nl[NZP], nq[NZP];
        auto probe = [&](int j, float x, float &l, float &qv) {
            const int jj = j - j0;
            l = 0.f; qv = 0.f;
            if (jj < 0 || jj >= nj) return false;
            l = log1pf(fmaxf(x, 0.f));
            qv = ex2_f(fmaf(sG[jj], kLog2e, s_b[jj])) + s_cv[jj];
            red[2] = fmaf(l, qv, red[2]);
            red[3] += (l + e_obs) * (l + e_obs) - e_obs * e_obs;
            return true;
        };
#pragma unroll
        for (int i = 0; i < NZP; ++i) {
            nl[i] = nq[i] = 0.f;
            if (pj[i] >= 0 && !probe(pj[i], px[i], nl[i], nq[i])) pj[i] = -1;
        }
        for (int64_t q = q_lo + tid + NZP * kT; q < q_hi; q += kT) {
            float l, qv;
            probe(a.indices[q], a.vals[q], l, qv);
        }
        // ---- one reduction of the four sums: warp -> CTA -> cluster
#pragma unroll
        for (int k = 0; k < 4; ++k) red[k] = warp_sum(red[k]);
        if (lane == 0) *reinterpret_cast<float4 *>(ws + warp * 4) = make_float4(red[0], red[1], red[2], red[3]);
        __syncthreads();
        {
            float4 v = lane < NW ? *reinterpret_cast<const float4 *>(ws + lane * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
            red[0] = warp_sum(v.x); red[1] = warp_sum(v.y); red[2] = warp_sum(v.z); red[3] = warp_sum(v.w);
        }
        if (cs > 1) {
            if (tid == 0) {
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(xb), "r"(16u * (uint32_t)cs) : "memory");
                const uint32_t slot_local = smem_addr(xs + rank * 4);
                for (int r = 0; r < cs; ++r) {
                    uint32_t rslot, rbar;
                    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rslot) : "r"(slot_local), "r"(r));
                    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rbar) : "r"(xb), "r"(r));
                    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];"
                                 ::"r"(rslot), "f"(red[0]), "f"(red[1]), "f"(red[2]), "f"(red[3]), "r"(rbar) : "memory");
                }
            }
            bar_wait(xb, ph);
            red[0] = red[1] = red[2] = red[3] = 0.f;
#pragma unroll
            for (int r = 0; r < 8; ++r)
                if (r < cs) {
                    const float4 v = *reinterpret_cast<const float4 *>(xs + r * 4);
                    red[0] += v.x; red[1] += v.y; red[2] += v.z; red[3] += v.w;
                }
        }
        const float qn = fmaxf(sqrtf(red[0]), 1e-12f);                                   // F::normalize eps
        const float nobs = fmaxf(sqrtf(red[3] + (float)D * e_obs * e_obs), 1e-12f);
        const float t = (red[2] + e_obs * red[1]) / (nobs * qn);
        if (rank == 0 && tid == 0) {
            t_acc += t;
            if (a.row_t) a.row_t[row] = t;
        }

        if (grad) {
            // ---- pass 2: dL/dq = -kappa (yobs - recon t) / (B ||q||) with yobs = e / Nobs on the dense sweep;
            //      dL/dpre = dL/dq exp(pre) = dL/dq (q - covariate term); straight to HBM
            const float coef = -kappa * a.inv_B / qn, inv_nobs = 1.f / nobs;
            const f2_t cA = bc2(-coef * t / qn), cB = bc2(coef * e_obs * inv_nobs), neg1 = bc2(-1.f);
            if (a.dpre_bf16) {
                __nv_bfloat16 *orow = reinterpret_cast<__nv_bfloat16 *>(a.dpre_bf16) + (size_t)row * D + j0 + 2 * tid;
#pragma unroll
                for (int h = 0; h < HP; ++h) {
                    const int jj = 2 * tid + h * 2 * kT;
                    const f2_t dq = fma2(qp[h], cA, cB);
                    cdq[h] = add2(cdq[h], dq);
                    float o0, o1;
                    up2(mul2(dq, fma2(*reinterpret_cast<const f2_t *>(s_cv + jj), neg1, qp[h])), o0, o1);
                    if (jj < nj) *reinterpret_cast<__nv_bfloat162 *>(orow + h * 2 * kT) = __floats2bfloat162_rn(o0, o1);
                }
            } else {
                float *orow = a.dpre + (size_t)row * D + j0 + 2 * tid;
#pragma unroll
                for (int h = 0; h < HP; ++h) {
                    const int jj = 2 * tid + h * 2 * kT;
                    const f2_t dq = fma2(qp[h], cA, cB);
                    cdq[h] = add2(cdq[h], dq);
                    float o0, o1;
                    up2(mul2(dq, fma2(*reinterpret_cast<const f2_t *>(s_cv + jj), neg1, qp[h])), o0, o1);
                    if (jj < nj) *reinterpret_cast<float2 *>(orow + h * 2 * kT) = make_float2(o0, o1);
                }
            }
            __syncthreads();   // the dense stores of this CTA precede the per-entry overwrites below
            // ---- nonzeros: yobs = (l + e) / Nobs -> dL/dq gains coef l / Nobs; overwrite the entry, fix the column sums
            auto fix = [&](int j, float l, float qv) {
                const int jj = j - j0;
                const float ddq = coef * l * inv_nobs;
                const float dq = fmaf(qv, -coef * t / qn, coef * e_obs * inv_nobs) + ddq;
                const float o = dq * (qv - s_cv[jj]);
                if (a.dpre_bf16) reinterpret_cast<__nv_bfloat16 *>(a.dpre_bf16)[(size_t)row * D + j] = __float2bfloat16(o);
                else a.dpre[(size_t)row * D + j] = o;
                atomicAdd(a.col_dq + j, ddq);
                for (int c = 0; c < C; ++c) atomicAdd(a.col_dqc + (size_t)j * C + c, ddq);
            };
#pragma unroll
            for (int i = 0; i < NZP; ++i)
                if (pj[i] >= 0) fix(pj[i], nl[i], nq[i]);
            for (int64_t q = q_lo + tid + NZP * kT; q < q_hi; q += kT) {
                const int j = a.indices[q], jj = j - j0;
                if (jj < 0 || jj >= nj) continue;
                fix(j, log1pf(fmaxf(a.vals[q], 0.f)), ex2_f(fmaf(sG[jj], kLog2e, s_b[jj])) + s_cv[jj]);
            }
        }
        if (next < a.B) {   // prefetch the next row's nonzeros
#pragma unroll
            for (int i = 0; i < NZP; ++i) {
                const int64_t q = qlo_n + tid + i * kT;
                pj[i] = q < qhi_n ? a.indices[q] : -1;
                px[i] = q < qhi_n ? a.vals[q] : 0.f;
            }
        }
        __syncthreads();   // everyone is done with sG[cur] before the TMA unit refills it
    }

    if (rank == 0 && tid == 0 && a.t_acc) atomicAdd(a.t_acc, (double)t_acc);
    if (grad) {
#pragma unroll
        for (int h = 0; h < HP; ++h) {
            const int jj = 2 * tid + h * 2 * kT;
            if (jj < nj) {
                float c0, c1;
                up2(cdq[h], c0, c1);
                atomicAdd(a.col_dq + j0 + jj, c0);
                atomicAdd(a.col_dq + j0 + jj + 1, c1);
                for (int c = 0; c < C; ++c) {
                    atomicAdd(a.col_dqc + (size_t)(j0 + jj) * C + c, c0);
                    atomicAdd(a.col_dqc + (size_t)(j0 + jj + 1) * C + c, c1);
                }
            }
        }
    }
    if (cs > 1) cluster.sync();   // no CTA exits while a peer may still signal its barriers
}

size_t
fast_smem_bytes(int ept)
{
    return sizeof(float) * ((size_t)4 * ept * kT + 2 * NW * 4 + 64) + 4 * 8 + 16;
}

int
ept_class(int ept)
{
    return ept <= 4 ? 4 : ept <= 8 ? 8 : ept <= 12 ? 12 : 16;
}

} // namespace

bool
vmf_likelihood_fast_eligible(const vmf_lik_args_t &a)
{
    return a.cov == nullptr && a.recon == nullptr && a.D % 4 == 0 && a.D <= 8 * 16 * kT - 32 && a.slice_bounds != nullptr &&
        (reinterpret_cast<uintptr_t>(a.G) & 15) == 0 && (a.dpre == nullptr || (reinterpret_cast<uintptr_t>(a.dpre) & 7) == 0) &&
        a.C >= 1 && a.C <= 8;
}

int
launch_vmf_likelihood_fast(const vmf_lik_args_t &a, int num_sms, cudaStream_t stream)
{
    const size_t budget = 72 * 1024;   // three CTAs per SM
    int cs = 1, Ds = 0, ept = 0;
    for (;; cs *= 2) {
        Ds = ((a.D + cs - 1) / cs + 3) & ~3;
        ept = (Ds + kT - 1) / kT;
        if (ept <= 16 && fast_smem_bytes(ept_class(ept)) <= budget) break;
        MMV_CHECK(cs < 8, "vmf_likelihood: D=%d does not fit a cluster of 8 CTAs", a.D);
    }
    MMV_TRY(launch_slice_bounds(a.indptr, a.indices, a.cells, a.unsorted, a.B, cs, Ds, a.slice_bounds, stream));
    void (*kern)(vmf_lik_args_t, int, int) = ept <= 4 ? vmf_likelihood_fast_kernel<4>
        : ept <= 8 ? vmf_likelihood_fast_kernel<8> : ept <= 12 ? vmf_likelihood_fast_kernel<12> : vmf_likelihood_fast_kernel<16>;
    const size_t smem = fast_smem_bytes(ept_class(ept));
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(kern), (int)smem));
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
    // persistent clusters: exactly one wave
    int n_clusters = 3 * num_sms / cs;
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
