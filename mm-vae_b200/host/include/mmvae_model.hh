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

// What the reference's LOSS functors (nb_loss_t / vmf_loss_t of the two mains) exchange with the training loop.  There
// the functor gets the minibatch tensor x and the model output y and returns the loss tensor whose backward() the loop
// calls (mmvae_alg.hh:277-279,303-307).  Here x and y never exist on the host: loss / backward run inside mmv_step, and
// the one thing the functor decides — the KL weight of this epoch — travels back in the "tensor" it returns.
struct batch_ref_t { };              // stands for the minibatch x (resident on the device)
struct recon_ref_t { };              // stands for the model output y
struct loss_request_t {              // what mmvae::nb::loss / mmvae::vmf::vmf_vae_loss return
    float kl_weight = 1.f;
};

class model_base_t {
public:
    model_base_t() = default;
    model_base_t(const model_base_t &) = delete;
    model_base_t &operator=(const model_base_t &) = delete;
    virtual ~model_base_t() { if (m_) mmv_model_destroy(m_); }
    mmv_model *get() const { return m_; }
    bool bound() const { return m_ != nullptr; }
    void release() { if (m_) mmv_model_destroy(m_); m_ = nullptr; }
    // reference-style construction (dims only, no device): the training loop binds the model to its context
    virtual int bind(mmv_ctx *, uint64_t /*seed*/) { return EXIT_SUCCESS; }
    int configure(const training_options_t &opt)
    {
        MMV_OK(mmv_model_set_option(m_, "lr", opt.lr));
        MMV_OK(mmv_model_set_option(m_, "grad_clip", opt.grad_clip));
        MMV_OK(mmv_model_set_option(m_, "weight_decay", 1e-4));   // mmvae_alg.hh:236
        MMV_OK(mmv_model_set_option(m_, "dtype", opt.f32x3 ? 2 : (opt.bf16 ? 1 : 0)));
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

// positive sizes / scalars with a name, as the reference passes them to its constructors (nb.hh:24-47, vmf.hh:14-48)
struct pos_size_t {
    explicit pos_size_t(const int64_t v) : val(v) { ASSERT(v > 0, "a dimension must be positive"); }
    const int64_t val;
};
struct pos_scalar_t {
    explicit pos_scalar_t(const float v) : val(v) { ASSERT(v > 0.f, "a scalar must be positive"); }
    const float val;
};
#define MMVAE_DIM(V) struct V : ::mmvae::pos_size_t { explicit V(const int64_t x) : ::mmvae::pos_size_t(x) { } }
#define MMVAE_SC(V) struct V : ::mmvae::pos_scalar_t { explicit V(const float x) : ::mmvae::pos_scalar_t(x) { } }

namespace nb {

MMVAE_DIM(data_dim);
MMVAE_DIM(covar_dim);
MMVAE_DIM(mu_encoder_h_dim);
MMVAE_DIM(mu_decoder_h_dim);
MMVAE_DIM(mu_encoder_r_dim);
MMVAE_DIM(nu_encoder_h_dim);
MMVAE_DIM(nu_encoder_r_dim);

class nbvae_t : public model_base_t {
public:
    nbvae_t(mmv_ctx *ctx, int64_t data_dim, int64_t covar_dim, const std::vector<int64_t> &mu_enc_h,
            const std::vector<int64_t> &mu_dec_h, int64_t mu_r_dim, int64_t nu_h_dim, int64_t nu_r_dim, bool do_relu,
            uint64_t seed)
    {
        set_dims(data_dim, covar_dim, mu_enc_h, mu_dec_h, mu_r_dim, nu_h_dim, nu_r_dim, do_relu);
        MMV_OK_EXIT(mmv_nb_model_create(ctx, &d_, seed, &m_));
    }
    // the reference's constructor (nbvae_tImpl, nb.hh:214-221): dimensions only; train_vae_model binds the device
    nbvae_t(const data_dim xd, const covar_dim cd, const std::vector<mu_encoder_h_dim> &eh, const std::vector<mu_decoder_h_dim> &dh,
            const mu_encoder_r_dim mr, const nu_encoder_h_dim oh, const nu_encoder_r_dim orr, const bool do_relu = true)
    {
        std::vector<int64_t> e, d;
        for (const auto &v : eh) e.push_back(v.val);
        for (const auto &v : dh) d.push_back(v.val);
        set_dims(xd.val, cd.val, e, d, mr.val, oh.val, orr.val, do_relu);
    }
    int bind(mmv_ctx *ctx, uint64_t seed) override
    {
        if (m_) return EXIT_SUCCESS;
        MMV_OK(mmv_nb_model_create(ctx, &d_, seed, &m_));
        return EXIT_SUCCESS;
    }
    nbvae_t *operator->() { return this; }   // the reference's model is a module holder: model->...
    int64_t dim_data() const { return d_.x_dim; }
    int64_t dim_mu_latent() const { return d_.mu_r_dim; }
    int64_t dim_nu_latent() const { return d_.nu_r_dim; }

private:
    void set_dims(int64_t data_dim, int64_t covar_dim, const std::vector<int64_t> &mu_enc_h, const std::vector<int64_t> &mu_dec_h,
                  int64_t mu_r_dim, int64_t nu_h_dim, int64_t nu_r_dim, bool do_relu)
    {
        std::memset(&d_, 0, sizeof(d_));
        d_.x_dim = data_dim; d_.c_dim = covar_dim;
        ASSERT(mu_enc_h.size() <= MMV_MAX_LAYERS && mu_dec_h.size() <= MMV_MAX_LAYERS, "too many layers");
        d_.n_enc = (int32_t)mu_enc_h.size();
        for (size_t i = 0; i < mu_enc_h.size(); ++i) d_.enc_dims[i] = mu_enc_h[i];
        d_.n_dec = (int32_t)mu_dec_h.size();
        for (size_t i = 0; i < mu_dec_h.size(); ++i) d_.dec_dims[i] = mu_dec_h[i];
        d_.mu_r_dim = mu_r_dim; d_.nu_h_dim = nu_h_dim; d_.nu_r_dim = nu_r_dim; d_.do_relu = do_relu ? 1 : 0;
        latent_dim = mu_r_dim;
        eps_per_cell = mu_r_dim + nu_r_dim;
    }
    mmv_nb_dims d_;
};

// nb.hh:540-548 `loss(x, y, kl_weight)`: on this path the ELBO and its backward run inside mmv_step (K4 + K3); what is
// left of the call is the KL weight the functor chose
inline loss_request_t
loss(batch_ref_t, recon_ref_t, float kl_weight = 1.f)
{
    loss_request_t r;
    r.kl_weight = kl_weight;
    return r;
}

} // namespace nb

namespace vmf {

MMVAE_DIM(data_dim);
MMVAE_DIM(covar_dim);
MMVAE_DIM(z_repr_dim);
MMVAE_DIM(z_enc_dim);
MMVAE_DIM(z_dec_dim);
MMVAE_SC(kappa_min);
MMVAE_SC(kappa_max);

class vmf_vae_t : public model_base_t {
public:
    vmf_vae_t(mmv_ctx *ctx, int64_t data_dim, int64_t covar_dim, int64_t z_dim, const std::vector<int64_t> &enc_h,
              const std::vector<int64_t> &dec_h, float kappa_min, float kappa_max, bool do_relu, uint64_t seed)
    {
        set_dims(data_dim, covar_dim, z_dim, enc_h, dec_h, kappa_min, kappa_max, do_relu);
        MMV_OK_EXIT(mmv_vmf_model_create(ctx, &d_, seed, &m_));
    }
    // the reference's constructor (vmf_vae_tImpl, vmf.hh:200-207,307-314): dimensions only
    vmf_vae_t(const data_dim xd, const covar_dim cd, const z_repr_dim zd, const std::vector<z_enc_dim> &eh,
              const std::vector<z_dec_dim> &dh, const kappa_min kmin = kappa_min{ 1. }, const kappa_max kmax = kappa_max{ 100. },
              const bool do_relu = true)
    {
        std::vector<int64_t> e, d;
        for (const auto &v : eh) e.push_back(v.val);
        for (const auto &v : dh) d.push_back(v.val);
        set_dims(xd.val, cd.val, zd.val, e, d, kmin.val, kmax.val, do_relu);
    }
    int bind(mmv_ctx *ctx, uint64_t seed) override
    {
        if (m_) return EXIT_SUCCESS;
        MMV_OK(mmv_vmf_model_create(ctx, &d_, seed, &m_));
        return EXIT_SUCCESS;
    }
    vmf_vae_t *operator->() { return this; }

private:
    void set_dims(int64_t data_dim, int64_t covar_dim, int64_t z_dim, const std::vector<int64_t> &enc_h,
                  const std::vector<int64_t> &dec_h, float kmin, float kmax, bool do_relu)
    {
        std::memset(&d_, 0, sizeof(d_));
        d_.x_dim = data_dim; d_.c_dim = covar_dim; d_.z_dim = z_dim;
        ASSERT(enc_h.size() <= MMV_MAX_LAYERS && dec_h.size() <= MMV_MAX_LAYERS, "too many layers");
        d_.n_enc = (int32_t)enc_h.size();
        for (size_t i = 0; i < enc_h.size(); ++i) d_.enc_dims[i] = enc_h[i];
        d_.n_dec = (int32_t)dec_h.size();
        for (size_t i = 0; i < dec_h.size(); ++i) d_.dec_dims[i] = dec_h[i];
        d_.kappa_min = kmin; d_.kappa_max = kmax; d_.do_relu = do_relu ? 1 : 0;
        latent_dim = z_dim;
        eps_per_cell = z_dim;
    }
    mmv_vmf_dims d_;
};

// vmf.hh:419-440 `vmf_vae_loss(x, y, kl_weight)`: see mmvae::nb::loss above
inline loss_request_t
vmf_vae_loss(batch_ref_t, recon_ref_t, float kl_weight = 1.f)
{
    loss_request_t r;
    r.kl_weight = kl_weight;
    return r;
}

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

// The recorders as the reference's mains construct them (nb.hh:569-580, vmf.hh:457-468): output header + number of
// epochs.  The six-argument train_vae_model (mmvae_alg.hh) turns one into a latent_recorder_t over the run's tables.
struct recorder_spec_t {
    recorder_spec_t(const std::string &hdr, int64_t max_epoch_, bool is_vmf_) : header(hdr), max_epoch(max_epoch_), is_vmf(is_vmf_) { }
    const std::string header;
    const int64_t max_epoch;
    const bool is_vmf;
};
namespace nb {
struct nbvae_recorder_t : recorder_spec_t {
    explicit nbvae_recorder_t(const std::string _hdr, const int64_t _max_epoch) : recorder_spec_t(_hdr, _max_epoch, false) { }
};
} // namespace nb
namespace vmf {
struct vmf_vae_recorder_t : recorder_spec_t {
    explicit vmf_vae_recorder_t(const std::string _hdr, const int64_t _max_epoch) : recorder_spec_t(_hdr, _max_epoch, true) { }
};
} // namespace vmf

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
