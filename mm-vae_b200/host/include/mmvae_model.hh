// Model handles with the reference's constructor arguments, over the C ABI (include/mmvae_b200.h):
//   mmvae::nb::nbvae_t    <- nbvae_tImpl ctor   include/models/nb.hh:214-221,299-401
//   mmvae::vmf::vmf_vae_t <- vmf_vae_tImpl ctor include/models/vmf.hh:200-207,307-389
// and the recorder visitors (update_on_batch / update_on_epoch, nb.hh:569-662, vmf.hh:457-551) reduced to
// what the hot path needs: the per-cell latent mean / log-variance table through mmv_encode (eval mode,
// no covariates, nb.hh:627-628) written as gz text at the recording epochs.
#ifndef MMVAE_B200_MODEL_HH_
#define MMVAE_B200_MODEL_HH_

#include "../../../include/mmvae_b200.h"
#include "mmvae_io.hh"
#include "mmvae_options.hh"

#include <condition_variable>
#include <cstdio>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>
#include <zlib.h>

namespace mmvae {

// What the host threads of one run share (--ngpu G: one thread per GPU): a reusable barrier, the NCCL unique id,
// and the recorder's latent tables [N x R], which every rank fills for the data points of its own shard.
struct shared_run_t {
    explicit shared_run_t(int n) : nthreads(n) { std::memset(uid, 0, sizeof(uid)); }
    void barrier()
    {
        if (nthreads <= 1) return;
        std::unique_lock<std::mutex> lk(mu);
        const uint64_t gen = generation;
        if (++waiting == nthreads) {
            waiting = 0;
            ++generation;
            cv.notify_all();
        } else {
            cv.wait(lk, [&] { return generation != gen; });
        }
    }
    int nthreads;
    unsigned char uid[128];
    std::vector<float> mean, lnvar;
    std::mutex mu;
    std::condition_variable cv;
    int waiting = 0;
    uint64_t generation = 0;
};

struct device_t {   // one context + the resident data (one shard) of a run
    mmv_ctx *ctx = nullptr;
    mmv_csr *csr = nullptr;
    int rank = 0, nranks = 1;
    shared_run_t *shared = nullptr;
    ~device_t()
    {
        if (csr) mmv_csr_destroy(csr);
        if (ctx) mmv_ctx_destroy(ctx);
    }
    void barrier() { if (shared) shared->barrier(); }
    // upload this block's counts (cells x genes CSR, shard-local cell ids) and covariates (dense [n x C], or none = all ones)
    int upload(int device, const mtx_data_block_t &data, const std::vector<float> *covar, int64_t c_dim)
    {
        MMV_OK(mmv_ctx_create(device, &ctx));
        MMV_OK(mmv_csr_upload(ctx, data.nlocal(), data.nfeature(), data.indptr.data(), data.indices.data(),
                              data.values.data(), covar ? covar->data() : nullptr, c_dim, &csr));
        return EXIT_SUCCESS;
    }
};

class model_base_t {
public:
    ~model_base_t() { if (m_) mmv_model_destroy(m_); }
    mmv_model *get() const { return m_; }
    int configure(const training_options_t &opt)
    {
        MMV_OK(mmv_model_set_option(m_, "lr", opt.lr));
        MMV_OK(mmv_model_set_option(m_, "grad_clip", opt.grad_clip));
        MMV_OK(mmv_model_set_option(m_, "weight_decay", 1e-4));   // mmvae_alg.hh:236
        MMV_OK(mmv_model_set_option(m_, "dtype", opt.bf16 ? 1 : 0));
        MMV_OK(mmv_model_set_option(m_, "train_all_layers", opt.train_all_layers ? 1 : 0));
        MMV_OK(mmv_model_set_option(m_, "frozen_grads", 0));      // never read on this path: skip the dead work
        return EXIT_SUCCESS;
    }
    void pretty_print(std::ostream &os) const
    {
        int n = 0;
        mmv_model_num_params(m_, &n);
        for (int i = 0; i < n; ++i) {
            char name[MMV_NAME_LEN];
            int64_t r, c;
            int reg;
            mmv_model_param_info(m_, i, name, &r, &c, &reg);
            os << "  " << name << " [" << r << " x " << c << "]" << (reg ? "" : " (not registered: frozen unless --train_all_layers)") << "\n";
        }
    }
    // --init_dir: ${dir}/${parameter key}.f32 (raw float32, libtorch [rows x cols] layout) replaces that parameter
    int load_params(const std::string &dir)
    {
        int n = 0, loaded = 0;
        MMV_OK(mmv_model_num_params(m_, &n));
        std::vector<float> host;
        for (int i = 0; i < n; ++i) {
            char name[MMV_NAME_LEN];
            int64_t r = 0, c = 0;
            int reg = 0;
            MMV_OK(mmv_model_param_info(m_, i, name, &r, &c, &reg));
            FILE *f = std::fopen((dir + "/" + name + ".f32").c_str(), "rb");
            if (!f) continue;
            host.resize((size_t)(r * c));
            const size_t got = std::fread(host.data(), 4, host.size(), f);
            const bool at_end = std::fgetc(f) == EOF;
            std::fclose(f);
            ERR_RET(got != host.size() || !at_end, dir << "/" << name << ".f32 does not hold " << r << " x " << c << " floats");
            MMV_OK(mmv_model_set_param(m_, name, host.data(), r * c));
            ++loaded;
        }
        TLOG("Initial values of " << loaded << " / " << n << " parameters read from " << dir);
        return EXIT_SUCCESS;
    }
    int64_t latent_dim = 0;
    int64_t eps_per_cell = 0;   // floats of reparameterisation noise per cell and forward pass (mu latent [+ nu latent])

protected:
    mmv_model *m_ = nullptr;
};

namespace nb {

class nbvae_t : public model_base_t {
public:
    nbvae_t(mmv_ctx *ctx, int64_t data_dim, int64_t covar_dim, const std::vector<int64_t> &mu_enc_h,
            const std::vector<int64_t> &mu_dec_h, int64_t mu_r_dim, int64_t nu_h_dim, int64_t nu_r_dim, bool do_relu,
            uint64_t seed)
    {
        mmv_nb_dims d;
        std::memset(&d, 0, sizeof(d));
        d.x_dim = data_dim; d.c_dim = covar_dim;
        ASSERT(mu_enc_h.size() <= MMV_MAX_LAYERS && mu_dec_h.size() <= MMV_MAX_LAYERS, "too many layers");
        d.n_enc = (int32_t)mu_enc_h.size();
        for (size_t i = 0; i < mu_enc_h.size(); ++i) d.enc_dims[i] = mu_enc_h[i];
        d.n_dec = (int32_t)mu_dec_h.size();
        for (size_t i = 0; i < mu_dec_h.size(); ++i) d.dec_dims[i] = mu_dec_h[i];
        d.mu_r_dim = mu_r_dim; d.nu_h_dim = nu_h_dim; d.nu_r_dim = nu_r_dim; d.do_relu = do_relu ? 1 : 0;
        MMV_OK_EXIT(mmv_nb_model_create(ctx, &d, seed, &m_));
        latent_dim = mu_r_dim;
        eps_per_cell = mu_r_dim + nu_r_dim;
    }
};

} // namespace nb

namespace vmf {

class vmf_vae_t : public model_base_t {
public:
    vmf_vae_t(mmv_ctx *ctx, int64_t data_dim, int64_t covar_dim, int64_t z_dim, const std::vector<int64_t> &enc_h,
              const std::vector<int64_t> &dec_h, float kappa_min, float kappa_max, bool do_relu, uint64_t seed)
    {
        mmv_vmf_dims d;
        std::memset(&d, 0, sizeof(d));
        d.x_dim = data_dim; d.c_dim = covar_dim; d.z_dim = z_dim;
        ASSERT(enc_h.size() <= MMV_MAX_LAYERS && dec_h.size() <= MMV_MAX_LAYERS, "too many layers");
        d.n_enc = (int32_t)enc_h.size();
        for (size_t i = 0; i < enc_h.size(); ++i) d.enc_dims[i] = enc_h[i];
        d.n_dec = (int32_t)dec_h.size();
        for (size_t i = 0; i < dec_h.size(); ++i) d.dec_dims[i] = dec_h[i];
        d.kappa_min = kappa_min; d.kappa_max = kappa_max; d.do_relu = do_relu ? 1 : 0;
        MMV_OK_EXIT(mmv_vmf_model_create(ctx, &d, seed, &m_));
        latent_dim = z_dim;
        eps_per_cell = z_dim;
    }
};

} // namespace vmf

// VISITOR concept (mmvae_alg.hh:316,332).  Writes ${out}.latent_mean_${epoch}.gz / latent_lnvar at the
// recording epochs: one line per data point, R columns (ostream default precision like io.hh:308-331).
// nbvae_recorder_t (nb.hh:569-662) / vmf_vae_recorder_t (vmf.hh:457-551): per recorded epoch
//   ${out}_${zeropad(epoch)}.mu_mean.gz / .mu_lnvar.gz        (nb)
//   ${out}_${zeropad(epoch)}.latent_mean.gz / .latent_lnvar.gz (vmf)
//   ${out}_${zeropad(epoch)}_${parameter key}.gz              every named parameter, as write_tensor
// (mmvae_io.hh:11-28) writes it: a 2-D tensor one row per line, a 1-D tensor one value per line, values
// space-separated with the ostream default format (%g, 6 significant digits; io.hh:522-534).
// The latent tables come from mmv_encode (eval mode, no covariate term: nb.hh:627-628, vmf.hh:516-517).
class latent_recorder_t {
public:
    // One recorder per rank over the run's SHARED latent tables: a rank writes the rows of its own shard
    // (global data-point id = shard_lo + local id), rank 0 writes the files at the recording epochs.
    latent_recorder_t(const std::string &out, int64_t max_epoch, shared_run_t *shared, int64_t ntot, bool is_vmf = false,
                      bool write_params = true)
        : out_(out), max_epoch_(max_epoch), ntot_(ntot), is_vmf_(is_vmf), write_params_(write_params), mean_(shared->mean),
          lnvar_(shared->lnvar), shared_(shared)
    {
    }

    template <typename MODEL, typename DATA_BLOCK>
    void update_on_batch(MODEL &model, DATA_BLOCK &data, const std::vector<typename DATA_BLOCK::Index> &batch, device_t &dev)
    {
        const int64_t R = model.latent_dim, B = (int64_t)batch.size();
        {
            std::lock_guard<std::mutex> lk(shared_->mu);
            if (mean_.empty()) {
                mean_.assign((size_t)(ntot_ * R), 0.f);
                lnvar_.assign((size_t)(ntot_ * R), 0.f);
            }
        }
        std::vector<int64_t> cells(batch.begin(), batch.end());
        std::vector<float> mu((size_t)(B * R)), lv((size_t)(B * R));
        MMV_OK_EXIT(mmv_encode(model.get(), dev.csr, cells.data(), B, mu.data(), lv.data()));
        const int64_t lo = (int64_t)data.shard_lo();
        for (int64_t j = 0; j < B; ++j) {
            std::copy(mu.begin() + j * R, mu.begin() + (j + 1) * R, mean_.begin() + (lo + cells[(size_t)j]) * R);
            std::copy(lv.begin() + j * R, lv.begin() + (j + 1) * R, lnvar_.begin() + (lo + cells[(size_t)j]) * R);
        }
    }

    template <typename MODEL>
    void update_on_epoch(MODEL &model, int64_t epoch)
    {
        const std::string tag = out_ + "_" + zeropad((int)epoch, (int)max_epoch_);   // nb.hh:590
        const int64_t R = model.latent_dim;
        write_matrix(tag + (is_vmf_ ? ".latent_mean.gz" : ".mu_mean.gz"), mean_.data(), R > 0 ? (int64_t)mean_.size() / R : 0, R);
        write_matrix(tag + (is_vmf_ ? ".latent_lnvar.gz" : ".mu_lnvar.gz"), lnvar_.data(), R > 0 ? (int64_t)lnvar_.size() / R : 0, R);
        if (!write_params_) return;
        int n = 0;
        MMV_OK_EXIT(mmv_model_num_params(model.get(), &n));
        std::vector<float> host;
        for (int i = 0; i < n; ++i) {   // nb.hh:599-615: mu_enc, mu_dec, then every registered parameter
            char name[MMV_NAME_LEN];
            int64_t r = 0, c = 0;
            int reg = 0;
            MMV_OK_EXIT(mmv_model_param_info(model.get(), i, name, &r, &c, &reg));
            host.resize((size_t)(r * c));
            MMV_OK_EXIT(mmv_model_get_param(model.get(), name, host.data(), r * c));
            const std::string key(name);
            const bool one_d = key.size() > 5 && key.compare(key.size() - 5, 5, ".bias") == 0;   // torch 1-D tensor -> Eigen column vector
            if (one_d) write_matrix(tag + "_" + key + ".gz", host.data(), r * c, 1);
            else write_matrix(tag + "_" + key + ".gz", host.data(), r, c);
        }
    }

    // row-major [rows x cols] -> gz text, "%g" (= operator<< of a float at the default precision)
    static void write_matrix(const std::string &path, const float *v, int64_t rows, int64_t cols)
    {
        if (rows <= 0 || cols <= 0) return;
        gzFile g = gzopen(path.c_str(), "wb");
        if (!g) { ELOG("unable to write " << path); return; }
        std::vector<char> buf((size_t)1 << 20);
        size_t used = 0;
        for (int64_t i = 0; i < rows; ++i) {
            for (int64_t j = 0; j < cols; ++j) {
                if (used + 32 > buf.size()) { gzwrite(g, buf.data(), (unsigned)used); used = 0; }
                used += (size_t)std::snprintf(buf.data() + used, 24, "%g", (double)v[i * cols + j]);
                buf[used++] = j + 1 == cols ? '\n' : ' ';
            }
        }
        gzwrite(g, buf.data(), (unsigned)used);
        gzclose(g);
    }

private:
    std::string out_;
    int64_t max_epoch_, ntot_;
    bool is_vmf_, write_params_;
    std::vector<float> &mean_, &lnvar_;
    shared_run_t *shared_;
};

// io.hh:308-331 write_vector_file: one value per line, gz
inline int
write_vector_file(const std::string &path, const std::vector<float> &v)
{
    gzFile g = gzopen(path.c_str(), "wb");
    ERR_RET(!g, "unable to write " << path);
    std::ostringstream os;
    for (float x : v) os << x << "\n";
    const std::string s = os.str();
    gzwrite(g, s.data(), (unsigned)s.size());
    gzclose(g);
    return EXIT_SUCCESS;
}

} // namespace mmvae
#endif
