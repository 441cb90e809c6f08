// train_vae_model — the reference's epoch / batch / bootstrap schedule (include/mmvae_alg.hh:200-338)
// driving the C ABI.  Per batch: one forward in TRAIN mode (z is sampled, :274-279) whose loss is
// reported, then nboot x (bootstrap resample with replacement :292-301, forward, loss, zero_grad,
// backward, clip_grad_norm_, Adam step :303-310).  The last batch wraps modulo N (:264-266).  What the
// reference does per batch on the host — re-parse text, densify, copy to the device, index_select —
// is gone: a batch is the list of cell ids, the bootstrap is an index composition inside mmv_step.
#ifndef MMVAE_B200_ALG_HH_
#define MMVAE_B200_ALG_HH_

#include "mmvae_model.hh"

#include <iomanip>
#include <random>

// KL anneal shared by nb_loss_t / vmf_loss_t (nb_vae_main.cc:14-37, vmf_vae_main.cc:13-36):
// rate = max(max_rate * exp(-time_discount * epoch), min_rate)
struct kl_anneal_t {
    kl_anneal_t() : time_discount(0.f), max_rate(1.f), min_rate(1.f) { }
    float operator()(const int64_t epoch) const
    {
        const float t = static_cast<float>(epoch);
        const float rate = max_rate * std::exp(-time_discount * t);
        return std::max(rate, min_rate);
    }
    float time_discount, max_rate, min_rate;
};
using nb_loss_t = kl_anneal_t;
using vmf_loss_t = kl_anneal_t;

template <typename MODEL, typename VISITOR, typename DATA_BLOCK, typename LOSS>
std::vector<float>
train_vae_model(MODEL &model, VISITOR &visitor, DATA_BLOCK &data_block, mmvae::device_t &dev, training_options_t &opt,
                LOSS loss_fun)
{
    TLOG("Training on GPU " << opt.device << " (B200, sm_100a)");
    const int64_t ntot = data_block.ntot();
    const int64_t batch_size = data_block.size();
    ASSERT(batch_size > 0 && ntot > 0, "empty data or batch");
    int64_t nbatch = ntot / batch_size;
    if ((nbatch * batch_size) < ntot) ++nbatch;
    std::vector<typename DATA_BLOCK::Index> batch(batch_size);
    TLOG("Batch size = " << batch_size << ", Number of batches = " << nbatch);

    ASSERT(model.configure(opt) == EXIT_SUCCESS, "model options");
    model.pretty_print(std::cerr);

    std::random_device rd;
    std::mt19937 rng(opt.seed ? (unsigned)opt.seed : rd());
    std::uniform_int_distribution<> rboot(0, (int)batch_size - 1);
    std::vector<int64_t> ridx(batch_size);

    std::vector<float> loss_vec;
    loss_vec.reserve(opt.max_epoch);

    for (int64_t epoch = 0; epoch < opt.max_epoch; ++epoch) {
        const float kl_weight = loss_fun(epoch);
        float _loss_epoch = 0.;
        for (int64_t b = 0; b < nbatch; ++b) {
            const int64_t lb = b * batch_size, ub = (b + 1) * batch_size;
            for (int64_t j = 0; j < (ub - lb); ++j) batch[j] = (lb + j) % ntot;
            data_block.read(batch);
            const std::vector<int64_t> &cells = data_block.batch();

            {   // reported loss: forward in train mode.  Deferred: this call hands back the terms of the previous
                // batch's evaluation, so the host never waits for the device inside the epoch (the reference pays
                // one .item() round trip per batch, mmvae_alg.hh:281); the last batch is collected after the loop
                mmv_loss_terms terms;
                MMV_OK_EXIT(mmv_step(model.get(), dev.csr, cells.data(), batch_size, nullptr, nullptr, kl_weight,
                                     MMV_FWD | MMV_EPS_PHILOX | MMV_TERMS_DEFERRED, &terms));
                if (b > 0) {
                    _loss_epoch += terms.loss * batch_size;
                    std::cerr << "\r[" << std::setw(20) << b << "] " << std::setw(20) << terms.loss;
                }
            }
            for (int64_t boot = 0; boot < opt.nboot; ++boot) {
                for (size_t r = 0; r < ridx.size(); ++r) ridx[r] = rboot(rng);
                MMV_OK_EXIT(mmv_step(model.get(), dev.csr, cells.data(), batch_size, ridx.data(), nullptr, kl_weight,
                                     MMV_FWD | MMV_BWD | MMV_OPTIM | MMV_EPS_PHILOX, nullptr));
            }
            if ((epoch + 1) % opt.recording == 0) visitor.update_on_batch(model, data_block, batch);
            data_block.clear();
        }
        {
            mmv_loss_terms terms;
            MMV_OK_EXIT(mmv_step_collect(model.get(), &terms));
            _loss_epoch += terms.loss * batch_size;
        }
        std::cerr << "\r";
        _loss_epoch /= batch_size * nbatch;
        TLOG("[" << std::setw(20) << (epoch + 1) << "] " << std::setw(20) << _loss_epoch);
        loss_vec.emplace_back(_loss_epoch);
        if ((epoch + 1) % opt.recording == 0) visitor.update_on_epoch(model, epoch);
    }
    MMV_OK_EXIT(mmv_ctx_sync(dev.ctx));
    TLOG("Done training");
    return loss_vec;
}

// Shared front half of the two mains (nb_vae_main.cc:39-82, vmf_vae_main.cc:38-80): index, data block,
// covariates.  Without --covar the reference writes a 1 x N all-ones mtx and reads it back; here the
// absent covariate IS the all-ones single covariate (mmv_csr_upload with covar = NULL), nothing is written.
struct run_inputs_t {
    std::unique_ptr<mmvae::mtx_data_block_t> data;
    std::vector<float> covar;   // [N x C] when --covar was given
    int64_t c_dim = 1;
    bool has_covar = false;
};

inline int
load_inputs(const mmvae_options_t &main_options, run_inputs_t &in)
{
    if (!file_exists(main_options.mtx)) return EXIT_FAILURE;
    if (!file_exists(main_options.idx)) CHK(mmutil::index::build_mmutil_index(main_options.mtx, main_options.idx));
    in.data.reset(new mmvae::mtx_data_block_t(main_options.mtx, main_options.idx, main_options.batch_size, main_options.csr_cache));
    if (file_exists(main_options.covar_mtx)) {
        if (!file_exists(main_options.covar_idx))
            CHK(mmutil::index::build_mmutil_index(main_options.covar_mtx, main_options.covar_idx));
        mmvae::mtx_data_block_t covar_block(main_options.covar_mtx, main_options.covar_idx, main_options.batch_size);
        ERR_RET(covar_block.ntot() != in.data->ntot(), "data and covar on the same set of data points");
        covar_block.dense_all(in.covar);
        in.c_dim = covar_block.nfeature();
        in.has_covar = true;
    } else {
        // nb_vae_main.cc:67-73: the reference writes a 1 x N all-ones Matrix Market file next to the outputs,
        // indexes it and reads it back; the same two files are produced here (create_ones_like), while the
        // device path keeps the all-ones covariate implicit (no B x 1 tensor, K4's tuned kernel stays eligible)
        if (main_options.out.size() > 0) {
            const std::string covar_mtx = main_options.out + ".covar.mtx.gz";
            CHK(mmvae::create_ones_like(in.data->ntot(), covar_mtx));
            TLOG("No covariate file is given. So we use this: " << covar_mtx);
            CHK(mmutil::index::build_mmutil_index(covar_mtx, covar_mtx + ".index"));
        }
    }
    return EXIT_SUCCESS;
}

#endif
