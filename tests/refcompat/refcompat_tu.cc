// A translation unit written against the REFERENCE's header names, types and call signatures (mmvae.hh, mmvae_io.hh,
// mmvae_alg.hh, nb.hh / vmf.hh, util.hh, io.hh; strong dimension types; a LOSS functor returning torch::Tensor; the
// six-argument train_vae_model) — compiled against mm-vae_b200/host/include/refcompat by tests/test_refcompat_cpu.py
// (CPU, recording test double of the C ABI) and tests/test_gpu_host_main.py (real library).  -DREFCOMPAT_VMF selects the
// vMF model.  What the reference's own mains do (src/nb_vae_main.cc:39-133, src/vmf_vae_main.cc:38-127) in fewer lines.
#include "mmvae.hh"
#include "mmvae_io.hh"
#include "mmvae_alg.hh"
#ifdef REFCOMPAT_VMF
#include "vmf.hh"
#else
#include "nb.hh"
#endif
#include "util.hh"
#include "std_util.hh"
#include "io.hh"
#include "mmvae_mem.hh"

struct annealed_elbo_t {   // LOSS concept of the reference: (x, y, epoch) -> loss tensor
    float hi = 1.f, lo = 1.f, decay = 0.f;
    template <typename X, typename Y>
    torch::Tensor operator()(X x, Y y, const int64_t epoch)
    {
        const float w = std::max(hi * std::exp(-decay * static_cast<float>(epoch)), lo);
#ifdef REFCOMPAT_VMF
        return mmvae::vmf::vmf_vae_loss(x, y, w);
#else
        return mmvae::nb::loss(x, y, w);
#endif
    }
};

int
main(const int argc, const char *argv[])
{
    mmvae_options_t io_opt;
    training_options_t fit_opt;
    CHK(parse_mmvae_options(argc, argv, io_opt));
    CHK(parse_training_options(argc, argv, fit_opt));
    if (!file_exists(io_opt.mtx)) return EXIT_FAILURE;
    if (!file_exists(io_opt.idx)) CHK(mmutil::index::build_mmutil_index(io_opt.mtx, io_opt.idx));
    mmvae::mtx_data_block_t counts(io_opt.mtx, io_opt.idx, io_opt.batch_size);

    std::string cov_mtx = io_opt.covar_mtx, cov_idx = io_opt.covar_idx;
    if (!file_exists(cov_mtx)) {
        cov_mtx = io_opt.out + ".covar.mtx.gz";
        cov_idx = cov_mtx + ".index";
        create_ones_like(counts, cov_mtx);
        CHK(mmutil::index::build_mmutil_index(cov_mtx, cov_idx));
    } else if (!file_exists(cov_idx)) {
        CHK(mmutil::index::build_mmutil_index(cov_mtx, cov_idx));
    }
    mmvae::mtx_data_block_t covariates(cov_mtx, cov_idx, io_opt.batch_size);

#ifdef REFCOMPAT_VMF
    using namespace mmvae::vmf;
    vmf_options_t arch;
    CHK(parse_vmf_options(argc, argv, arch));
    std::vector<z_enc_dim> enc;
    std::vector<z_dec_dim> dec;
    for (const int64_t d : arch.encoding_layers) enc.emplace_back(d);
    for (const int64_t d : arch.decoding_layers) dec.emplace_back(d);
    vmf_vae_t model(data_dim(counts.nfeature()), covar_dim(covariates.nfeature()), z_repr_dim(arch.latent), enc, dec,
                    kappa_min(arch.kappa_min), kappa_max(arch.kappa_max), arch.do_relu);
    vmf_vae_recorder_t recorder(io_opt.out, fit_opt.max_epoch);
#else
    using namespace mmvae::nb;
    nbvae_options_t arch;
    CHK(parse_nbvae_options(argc, argv, arch));
    std::vector<mu_encoder_h_dim> enc;
    std::vector<mu_decoder_h_dim> dec;
    for (const int64_t d : arch.mean_encoding_layers) enc.emplace_back(d);
    for (const int64_t d : arch.mean_decoding_layers) dec.emplace_back(d);
    nbvae_t model(data_dim(counts.nfeature()), covar_dim(covariates.nfeature()), enc, dec, mu_encoder_r_dim(arch.mean_latent),
                  nu_encoder_h_dim(arch.overdispersion_encoding), nu_encoder_r_dim(arch.overdispersion_latent), arch.do_relu);
    nbvae_recorder_t recorder(io_opt.out, fit_opt.max_epoch);
#endif

    annealed_elbo_t elbo;
    elbo.lo = io_opt.kl_min;
    elbo.hi = io_opt.kl_max;
    elbo.decay = io_opt.kl_discount;
    const auto scores = train_vae_model(model, recorder, counts, covariates, fit_opt, elbo);
    write_vector_file(io_opt.out + ".scores.gz", scores);
    return EXIT_SUCCESS;
}
