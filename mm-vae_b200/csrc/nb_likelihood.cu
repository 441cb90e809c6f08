// K4 — fused softmax + negative-binomial log-likelihood + gradient.
//
// Replaces, in ONE pass over the B x D decoder output, the reference's
//   decode_mu   nb.hh:433-442  (covar_dec + mu_bias + log_softmax + exp)
//   decode_nu   nb.hh:453-460  (rank-Rnu outer product, softplus, clamp)
//   nllik_loss  nb.hh:511-531  (3 lgamma + logs over B x D)
//   and the autograd backward of all of them.
//
// Mapping: one thread-block CLUSTER per cell (row); the D genes of the row are
// split across the CTAs of the cluster, each CTA keeps its slice of the row
// (softmax numerators -> p, and g = dl/dm) in shared memory, and the three row
// reductions (max, sum-exp, s = sum g*p) go through warp shuffles, one shared
// memory hop, and DSMEM across the cluster.  Per-gene column sums of the
// nu-path gradient are accumulated in shared memory across all the rows a CTA
// processes and flushed once with red.global.add at the end.
// x = 0 entries (95 %) need no lgamma/digamma: the dense sweep assumes x = 0
// and a sparse fix-up over the row's CSR entries corrects the nonzeros.
#include "common.cuh"
#include "kernels.hh"

#include "cluster_reduce.cuh"

#include <cuda_bf16.h>

namespace mmv {

constexpr int kNbThreads = 512;
constexpr float kEps = 1e-4f;  // nb.hh:514
constexpr float kNuMin = 1e-4f, kNuMax = 1e4f;  // nb.hh:252,459

struct nb_elem_t {
    float m, nu, l1p, sig_mask, inv_mnu;
};

// Element quantities shared by the dense (x = 0) sweep and the sparse fix-up.
__device__ __forceinline__ nb_elem_t
nb_elem(float p, float depth, float u)
{
    nb_elem_t e;
    e.m = p * depth + kEps;                               // nb.hh:519
    const float sp = softplus_f(u);                       // nb.hh:458
    const float r = fminf(fmaxf(sp, kNuMin), kNuMax);     // nb.hh:459
    e.nu = r + kEps;                                      // nb.hh:518
    e.l1p = log1pf(e.m / e.nu);                           // log(m+nu) - log(nu)
    e.inv_mnu = 1.f / (e.m + e.nu);
    // d softplus/du = sigmoid(u) (1 beyond the threshold); clamp passes grad inside [min,max]
    const bool inside = (sp >= kNuMin) && (sp <= kNuMax);
    e.sig_mask = inside ? (u > 20.f ? 1.f : sigmoid_f(u)) : 0.f;
    return e;
}

template <int RMAX>
__global__ void __launch_bounds__(kNbThreads, 1)
nb_likelihood_kernel(nb_lik_args_t a, int cs, int Ds)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = cs > 1 ? (int)cluster.block_rank() : 0;
    const int cluster_id = blockIdx.x / cs;
    const int n_clusters = gridDim.x / cs;
    const int tid = threadIdx.x;
    const int D = a.D, C = a.C, R = a.Rnu;
    const int j0 = rank * Ds;
    const int nj = max(0, min(D - j0, Ds));
    const bool grad = a.dlogits != nullptr || a.dlogits_bf16 != nullptr;

    extern __shared__ __align__(16) float smem[];
    float *sa = smem;                    // [Ds] logits -> exp -> p
    float *sg = sa + Ds;                 // [Ds] g = dl/dm
    float *sc_du = sg + Ds;              // [Ds] column sum of du
    float *sc_duz = sc_du + Ds;          // [R][Ds] column sum of du * znu_r
    float *scratch = sc_duz + (size_t)RMAX * Ds;  // [(1+RMAX)*32]
    float *slots = scratch + (1 + RMAX) * 32;     // [2][3][1+RMAX]

    if (grad) {
        for (int jj = tid; jj < nj; jj += kNbThreads) {
            sc_du[jj] = 0.f;
            for (int r = 0; r < R; ++r) sc_duz[r * Ds + jj] = 0.f;
        }
    }
    __syncthreads();

    float nll_acc = 0.f;
    int parity = 0;

    for (int row = cluster_id; row < a.B; row += n_clusters, parity ^= 1) {
        float *slot_max = slots + (parity * 3 + 0) * (1 + RMAX);
        float *slot_sum = slots + (parity * 3 + 1) * (1 + RMAX);
        float *slot_s = slots + (parity * 3 + 2) * (1 + RMAX);

        const int64_t cell = a.cells ? (int64_t)a.cells[row] : (int64_t)row;
        const float *Grow = a.G + (size_t)row * D + j0;
        const float depth = softplus_f(a.dpre[row]);          // nb.hh:498
        float zn[RMAX];
#pragma unroll
        for (int r = 0; r < RMAX; ++r) zn[r] = r < R ? a.znu[(size_t)row * R + r] : 0.f;

        // ---- pass 1: logits a_j = G_ij + (b3+b_cd+mu_bias)_j + c_i . Wcd_j ; row max
        float mx = -INFINITY;
        for (int jj = tid; jj < nj; jj += kNbThreads) {
            const int j = j0 + jj;
            float v = Grow[jj] + a.abias[j];
            for (int c = 0; c < C; ++c) {
                const float cv = a.cov ? a.cov[cell * C + c] : 1.f;
                v = fmaf(cv, a.Wcd[(size_t)j * C + c], v);
            }
            sa[jj] = v;
            mx = fmaxf(mx, v);
        }
        mx = cluster_max(mx, slot_max, scratch, cluster, cs);

        // ---- pass 2: softmax numerators
        float se[1] = { 0.f };
        for (int jj = tid; jj < nj; jj += kNbThreads) {
            const float e = expf(sa[jj] - mx);
            sa[jj] = e;
            se[0] += e;
        }
        cluster_sum<1>(se, slot_sum, scratch, cluster, cs);
        const float inv_sum = 1.f / se[0];
        if (a.row_lse && rank == 0 && tid == 0) a.row_lse[row] = mx + logf(se[0]);

        // ---- pass 3: dense sweep assuming x = 0
        float red[1 + RMAX];
#pragma unroll
        for (int k = 0; k < 1 + RMAX; ++k) red[k] = 0.f;
        for (int jj = tid; jj < nj; jj += kNbThreads) {
            const int j = j0 + jj;
            const float p = sa[jj] * inv_sum;
            float u = a.ubias[j];
#pragma unroll
            for (int r = 0; r < RMAX; ++r)
                if (r < R) u = fmaf(zn[r], a.Wnd[(size_t)j * R + r], u);
            const nb_elem_t e = nb_elem(p, depth, u);
            nll_acc += e.nu * e.l1p;                       // x = 0: lgammas cancel
            sa[jj] = p;
            if (a.recon_mu) a.recon_mu[(size_t)row * D + j] = p;
            if (a.recon_nu) a.recon_nu[(size_t)row * D + j] = e.nu - kEps;
            if (grad) {
                const float g = e.nu * e.inv_mnu;          // (nu+0)/(m+nu) - 0/m
                sg[jj] = g;
                red[0] = fmaf(g, p, red[0]);
                const float dnu = e.l1p - e.m * e.inv_mnu;
                const float du = dnu * e.sig_mask * a.inv_B;
                sc_du[jj] += du;
#pragma unroll
                for (int r = 0; r < RMAX; ++r)
                    if (r < R) {
                        sc_duz[r * Ds + jj] = fmaf(du, zn[r], sc_duz[r * Ds + jj]);
                        red[1 + r] = fmaf(du, a.Wnd[(size_t)j * R + r], red[1 + r]);
                    }
            }
        }
        __syncthreads();

        // ---- pass 3b: sparse fix-up for the nonzeros of this cell
        {
            const int64_t beg = a.indptr[cell], end = a.indptr[cell + 1];
            for (int64_t k = beg + tid; k < end; k += kNbThreads) {
                const int j = a.indices[k];
                const int jj = j - j0;
                if (jj < 0 || jj >= nj) continue;
                const float x = a.vals[k];
                if (x == 0.f) continue;
                const float p = sa[jj];
                float u = a.ubias[j];
#pragma unroll
                for (int r = 0; r < RMAX; ++r)
                    if (r < R) u = fmaf(zn[r], a.Wnd[(size_t)j * R + r], u);
                const nb_elem_t e = nb_elem(p, depth, u);
                const float L = logf(e.m + e.nu);
                // nb.hh:522-528
                const float ell = lgammaf(e.nu) + lgammaf(x + 1.f) - lgammaf(e.nu + x) +
                    x * (L - logf(e.m)) + e.nu * e.l1p;
                nll_acc += ell - e.nu * e.l1p;
                if (grad) {
                    const float g0 = e.nu * e.inv_mnu;
                    const float gx = (e.nu + x) * e.inv_mnu - x / e.m;
                    sg[jj] = gx;
                    red[0] = fmaf(gx - g0, p, red[0]);
                    const float dnu0 = e.l1p - e.m * e.inv_mnu;
                    const float dnux = digamma_f(e.nu) - digamma_f(e.nu + x) + e.l1p +
                        (x - e.m) * e.inv_mnu;
                    const float ddu = (dnux - dnu0) * e.sig_mask * a.inv_B;
                    sc_du[jj] += ddu;
#pragma unroll
                    for (int r = 0; r < RMAX; ++r)
                        if (r < R) {
                            sc_duz[r * Ds + jj] = fmaf(ddu, zn[r], sc_duz[r * Ds + jj]);
                            red[1 + r] = fmaf(ddu, a.Wnd[(size_t)j * R + r], red[1 + r]);
                        }
                }
            }
        }

        if (grad) {
            cluster_sum<1 + RMAX>(red, slot_s, scratch, cluster, cs);
            const float s = red[0];
            // ---- pass 4: dL/dlogit_j = (d_i / B) p_j (g_j - s_i)
            const float coef = depth * a.inv_B;
            if (a.dlogits_bf16) {
                __nv_bfloat16 *drow = reinterpret_cast<__nv_bfloat16 *>(a.dlogits_bf16) + (size_t)row * D + j0;
                for (int jj = tid; jj < nj; jj += kNbThreads)
                    drow[jj] = __float2bfloat16(coef * sa[jj] * (sg[jj] - s));
            } else {
                float *drow = a.dlogits + (size_t)row * D + j0;
                for (int jj = tid; jj < nj; jj += kNbThreads)
                    drow[jj] = coef * sa[jj] * (sg[jj] - s);
            }
            if (rank == 0 && tid == 0) {
                const float dp = a.dpre[row];
                a.ddpre[row] = s * a.inv_B * (dp > 20.f ? 1.f : sigmoid_f(dp));
                for (int r = 0; r < R; ++r) a.dznu[(size_t)row * R + r] = red[1 + r];
            }
        } else {
            __syncthreads();
        }
    }

    // ---- flush: loss partial and per-gene column sums
    float nl[1] = { nll_acc };
    block_sum<1>(nl, scratch);
    if (tid == 0 && nl[0] != 0.f) atomicAdd(a.loss_acc, (double)nl[0]);
    if (grad) {
        for (int jj = tid; jj < nj; jj += kNbThreads) {
            const int j = j0 + jj;
            atomicAdd(a.col_du + j, sc_du[jj]);
            for (int r = 0; r < R; ++r) atomicAdd(a.col_duz + (size_t)j * R + r, sc_duz[r * Ds + jj]);
        }
    }
    if (cs > 1) cluster.sync();  // keep DSMEM alive until every peer is done
}

static size_t
nb_smem_bytes(int Ds, int rmax)
{
    return sizeof(float) * ((size_t)(3 + rmax) * Ds + (1 + rmax) * 32 + 2 * 3 * (1 + rmax));
}

int
launch_nb_likelihood(const nb_lik_args_t &a, int num_sms, cudaStream_t stream)
{
    if (a.rowmeta != nullptr) return launch_nb_likelihood_fast(a, num_sms, stream);   // the caller prepared the tuned path
    MMV_CHECK(a.Rnu >= 1 && a.Rnu <= 8, "nb_likelihood: overdispersion latent dim %d not in [1,8]", a.Rnu);
    const int rmax = a.Rnu == 1 ? 1 : (a.Rnu <= 4 ? 4 : 8);
    const size_t budget = 200 * 1024;
    int cs = 1, Ds = 0;
    for (;; cs *= 2) {
        Ds = ((a.D + cs - 1) / cs + 3) & ~3;
        if (nb_smem_bytes(Ds, rmax) <= budget) break;
        MMV_CHECK(cs < 8, "nb_likelihood: D=%d does not fit a cluster of 8 CTAs", a.D);
    }
    const size_t smem = nb_smem_bytes(Ds, rmax);
    void (*kern)(nb_lik_args_t, int, int) =
        rmax == 1 ? nb_likelihood_kernel<1> : (rmax == 4 ? nb_likelihood_kernel<4> : nb_likelihood_kernel<8>);
    MMV_TRY(dev_set_smem(reinterpret_cast<const void *>(kern), (int)smem));
    int n_clusters = num_sms / cs;
    if (n_clusters > a.B) n_clusters = a.B;
    if (n_clusters < 1) n_clusters = 1;

    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(n_clusters * cs);
    cfg.blockDim = dim3(kNbThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    MMV_CUDA(cudaLaunchKernelEx(&cfg, kern, a, cs, Ds));
    ::mmv::count_launch();
    return 0;
}

} // namespace mmv
