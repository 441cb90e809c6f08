// nb_vae — same command line as the reference's src/nb_vae_main.cc:39-133, B200 build (+ --ngpu G: cells sharded
// over G GPUs of the box, one host thread and one context per GPU, see mmvae_alg.hh).
#include "mmvae_alg.hh"

int
main(const int argc, const char *argv[])
{
    mmvae_options_t main_options;
    training_options_t train_opt;
    mmvae::nb::nbvae_options_t nb_opt;

    CHK(parse_mmvae_options(argc, argv, main_options));
    CHK(parse_nbvae_options(argc, argv, nb_opt));
    CHK(parse_training_options(argc, argv, train_opt));

    return run_vae_main(main_options, train_opt, /*is_vmf=*/false, [&](mmv_ctx *ctx, int64_t n_features, int64_t c_dim) {
        return std::unique_ptr<mmvae::nb::nbvae_t>(new mmvae::nb::nbvae_t(
            ctx, n_features, c_dim, nb_opt.mean_encoding_layers, nb_opt.mean_decoding_layers, nb_opt.mean_latent,
            nb_opt.overdispersion_encoding, nb_opt.overdispersion_latent, nb_opt.do_relu, train_opt.seed + 7));
    });
}
