// K5 — fused exp + covariate add + row-L2 normalise + sparse dot + von
// Mises-Fisher log-likelihood + gradient.
//
// Replaces vmf_vae_tImpl::decode (vmf.hh:283-290), the data side of
// vmf_vae_loss (vmf.hh:419-440) and their autograd backward:
//   q_ij = exp(pre_ij) + cov_ij ;  recon_i = q_i / max(||q_i||, 1e-12)
//   yobs_ij = (log1p(relu x_ij) + e) / Nobs_i , e = 1e-2 / D   (dense in the reference)
//   t_i = <yobs_i, recon_i> = (sum_nnz l_ij q_ij + e sum_j q_ij) / (Nobs_i ||q_i||)
//   dL/dq_ij = -kappa (yobs_ij - recon_ij t_i) / (B ||q_i||) ; dL/dpre_ij = dL/dq_ij exp(pre_ij)
// Same mapping as K4: one thread-block cluster per cell, gene slices in shared
// memory, warp-shuffle + DSMEM row reductions, per-gene column sums of dL/dq
// (covar_decoding_ gradients) accumulated in shared memory and flushed once.
#include "common.cuh"
#include "kernels.hh"

#include <cooperative_groups.h>

namespace cg = cooperative_groups;

namespace mmv {

constexpr int kVmfThreads = 512;
constexpr int kVmfCMax = 8;

template <int N>
__device__ __forceinline__ void
vmf_cluster_sum(float (&v)[N], float *slot, float *scratch, cg::cluster_group &cluster, int cs)
{
    block_sum<N>(v, scratch);
    if (cs == 1) return;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int k = 0; k < N; ++k) slot[k] = v[k];
    }
    cluster.sync();
#pragma unroll
    for (int k = 0; k < N; ++k) v[k] = 0.f;
    for (int r = 0; r < cs; ++r) {
        const float *remote = cluster.map_shared_rank(slot, r);
#pragma unroll
        for (int k = 0; k < N; ++k) v[k] += remote[k];
    }
}

__global__ void __launch_bounds__(kVmfThreads, 1)
vmf_likelihood_kernel(vmf_lik_args_t a, int cs, int Ds)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = cs > 1 ? (int)cluster.block_rank() : 0;
    const int cluster_id = blockIdx.x / cs, n_clusters = gridDim.x / cs;
    const int tid = threadIdx.x, D = a.D, C = a.C;
    const int j0 = rank * Ds, nj = max(0, min(D - j0, Ds));
    const bool grad = a.dpre != nullptr;
    const float e_obs = a.eps_obs;

    extern __shared__ __align__(16) float smem[];
    float *sq = smem;                    // [Ds] q
    float *sl = sq + Ds;                 // [Ds] log1p(relu x), dense (zero-filled)
    float *sc_dq = sl + Ds;              // [Ds] column sum of dL/dq
    float *sc_dqc = sc_dq + Ds;          // [C][Ds] column sum of dL/dq * c
    float *scratch = sc_dqc + (size_t)C * Ds;   // [4*32]
    float *slots = scratch + 4 * 32;            // [2][4]

    for (int jj = tid; jj < nj; jj += kVmfThreads) {
        sc_dq[jj] = 0.f;
        for (int c = 0; c < C; ++c) sc_dqc[c * Ds + jj] = 0.f;
    }
    __syncthreads();

    const float kappa = *a.kappa;
    float t_acc = 0.f;
    int parity = 0;

    for (int row = cluster_id; row < a.B; row += n_clusters, parity ^= 1) {
        float *slot = slots + parity * 4;
        const int64_t cell = a.cells ? (int64_t)a.cells[row] : (int64_t)row;
        const float *Grow = a.G + (size_t)row * D + j0;
        float cv[kVmfCMax];
#pragma unroll
        for (int c = 0; c < kVmfCMax; ++c) cv[c] = c < C ? (a.cov ? a.cov[cell * C + c] : 1.f) : 0.f;

        // ---- pass 1: q_j, sum q^2, sum q
        float red[4] = { 0.f, 0.f, 0.f, 0.f };   // sum q^2, sum q, sum_nnz l q, sum_nnz (l+e)^2 - e^2
        for (int jj = tid; jj < nj; jj += kVmfThreads) {
            const int j = j0 + jj;
            float cvj = a.bcd[j];
#pragma unroll
            for (int c = 0; c < kVmfCMax; ++c)
                if (c < C) cvj = fmaf(cv[c], a.Wcd[(size_t)j * C + c], cvj);
            const float q = expf(Grow[jj] + a.blast[j]) + cvj;    // vmf.hh:287-289
            sq[jj] = q;
            sl[jj] = 0.f;
            red[0] = fmaf(q, q, red[0]);
            red[1] += q;
        }
        __syncthreads();
        // ---- sparse part of <yobs, q> and ||yobs||
        {
            const int64_t beg = a.indptr[cell], end = a.indptr[cell + 1];
            for (int64_t k = beg + tid; k < end; k += kVmfThreads) {
                const int jj = a.indices[k] - j0;
                if (jj < 0 || jj >= nj) continue;
                const float l = log1pf(fmaxf(a.vals[k], 0.f));    // vmf.hh:424
                sl[jj] = l;
                red[2] = fmaf(l, sq[jj], red[2]);
                red[3] += (l + e_obs) * (l + e_obs) - e_obs * e_obs;
            }
        }
        vmf_cluster_sum<4>(red, slot, scratch, cluster, cs);
        const float qn = fmaxf(sqrtf(red[0]), 1e-12f);
        const float nobs = fmaxf(sqrtf(red[3] + (float)D * e_obs * e_obs), 1e-12f);
        const float t = (red[2] + e_obs * red[1]) / (nobs * qn);
        if (rank == 0 && tid == 0) {
            t_acc += t;
            if (a.row_t) a.row_t[row] = t;
        }

        // ---- pass 2: gradient
        if (grad || a.recon) {
            const float coef = -kappa * a.inv_B / qn;
            const float inv_qn = 1.f / qn, inv_nobs = 1.f / nobs;
            for (int jj = tid; jj < nj; jj += kVmfThreads) {
                const int j = j0 + jj;
                const float q = sq[jj];
                const float rec = q * inv_qn;
                if (a.recon) a.recon[(size_t)row * D + j] = rec;
                if (grad) {
                    const float y = (sl[jj] + e_obs) * inv_nobs;
                    const float dq = coef * (y - rec * t);
                    float cvj = a.bcd[j];
#pragma unroll
                    for (int c = 0; c < kVmfCMax; ++c)
                        if (c < C) cvj = fmaf(cv[c], a.Wcd[(size_t)j * C + c], cvj);
                    a.dpre[(size_t)row * D + j] = dq * (q - cvj);   // * exp(pre)
                    sc_dq[jj] += dq;
#pragma unroll
                    for (int c = 0; c < kVmfCMax; ++c)
                        if (c < C) sc_dqc[c * Ds + jj] = fmaf(dq, cv[c], sc_dqc[c * Ds + jj]);
                }
            }
        }
        __syncthreads();
    }

    if (rank == 0 && tid == 0 && a.t_acc) atomicAdd(a.t_acc, (double)t_acc);
    if (grad) {
        for (int jj = tid; jj < nj; jj += kVmfThreads) {
            const int j = j0 + jj;
            atomicAdd(a.col_dq + j, sc_dq[jj]);
            for (int c = 0; c < C; ++c) atomicAdd(a.col_dqc + (size_t)j * C + c, sc_dqc[c * Ds + jj]);
        }
    }
    if (cs > 1) cluster.sync();
}

static size_t
vmf_smem_bytes(int Ds, int C)
{
    return sizeof(float) * ((size_t)(3 + C) * Ds + 4 * 32 + 8);
}

int
launch_vmf_likelihood(const vmf_lik_args_t &a, int num_sms, cudaStream_t stream)
{
    MMV_CHECK(a.C >= 1 && a.C <= kVmfCMax, "vmf_likelihood: covariate dim %d not in [1,%d]", a.C, kVmfCMax);
    if (a.allow_fast && vmf_likelihood_fast_eligible(a)) return launch_vmf_likelihood_fast(a, num_sms, stream);
    MMV_CHECK(a.dpre_bf16 == nullptr, "vmf_likelihood: the general kernel writes fp32 gradients only");
    const size_t budget = 200 * 1024;
    int cs = 1, Ds = 0;
    for (;; cs *= 2) {
        Ds = ((a.D + cs - 1) / cs + 3) & ~3;
        if (vmf_smem_bytes(Ds, a.C) <= budget) break;
        MMV_CHECK(cs < 8, "vmf_likelihood: D=%d does not fit a cluster of 8 CTAs", a.D);
    }
    const size_t smem = vmf_smem_bytes(Ds, a.C);
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(vmf_likelihood_kernel), (int)smem));
    int n_clusters = num_sms / cs;
    if (n_clusters > a.B) n_clusters = a.B;
    if (n_clusters < 1) n_clusters = 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_clusters * cs);
    cfg.blockDim = dim3(kVmfThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MMV_CUDA(cudaLaunchKernelEx(&cfg, vmf_likelihood_kernel, a, cs, Ds));
    ::mmv::count_launch();
    return 0;
}

// ---------------------------------------------------------------- scalar parts
// kappa = clamp(exp(ln_kappa), kmin, kmax)                         vmf.hh:301
__global__ void
vmf_kappa_kernel(const float *ln_kappa, float kmin, float kmax, float *kappa)
{
    // double exp, rounded once: at init ln_kappa = log(kappa_min) sits on the clamp edge and the
    // reference's CPU expf lands just below kappa_min (grad masked)
    const float e = (float)exp((double)ln_kappa[0]);
    kappa[0] = fminf(fmaxf(e, kmin), kmax);
}

// Loss constant + d ln_kappa.  lbessel forward/backward: operators.hh:13-101.
//   llik_sum = kappa sum_i t_i + B (df log kappa - lbessel(kappa, df) - 0.5 D fasterlog(2 pi))
//   dL/dkappa = -(1/B) sum_i t_i - df / kappa + G(kappa)   (G ignores the upstream gradient)
__global__ void
vmf_scalar_kernel(vmf_scalar_args_t s)
{
    const float kappa = s.kappa[0];
    const float df = s.df;
    const float eta = s.eta;
    const float logk = logf(kappa);
    float lb;
    if (kappa <= df) lb = df * logk + eta * kappa - s.c_eta_nu_log2 - s.c_fasterlgamma;
    else lb = kappa - 0.5f * logk - s.c_half_log_2pi;
    const double tsum = *s.t_acc;
    const float per_cell = df * logk - lb - s.c_half_D_fasterlog_2pi;
    s.llik_sum[0] = (double)kappa * tsum + (double)s.B * (double)per_cell;
    if (s.dln_kappa) {
        // -df/kappa and G(kappa) ~ df/kappa cancel to O(1): evaluate in double (the reference's fp32
        // autograd sum of these terms carries ~1e-3 relative noise at D = 20k; see tests)
        const double kd = kappa, dfd = df;
        const double lbq = sqrt(kd * kd * dfd / (dfd + 1.) + dfd * dfd);
        const double ubq = sqrt(kd * kd + dfd * dfd);
        const double G = 0.5 * (lbq + ubq) / kd;
        // the data term is a per-cell sum (summed over ranks by the allreduce); the two
        // scalar terms are not, so each rank contributes its 1/nranks share
        const float dk = (float)(-tsum * (double)s.inv_B_global + (-dfd / kd + G) * (double)s.inv_ranks);
        const float ek = (float)exp((double)s.ln_kappa[0]);
        const bool inside = ek >= s.kmin && ek <= s.kmax;
        s.dln_kappa[0] += inside ? dk * ek : 0.f;
    }
}

int
launch_vmf_kappa(const float *ln_kappa, float kmin, float kmax, float *kappa, cudaStream_t st)
{
    vmf_kappa_kernel<<<1, 1, 0, st>>>(ln_kappa, kmin, kmax, kappa); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

int
launch_vmf_scalar(const vmf_scalar_args_t &s, cudaStream_t st)
{
    vmf_scalar_kernel<<<1, 1, 0, st>>>(s); ::mmv::count_launch();
    MMV_CUDA(cudaGetLastError());
    return 0;
}

} // namespace mmv
