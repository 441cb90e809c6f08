// vmf_vae — same command line as the reference's src/vmf_vae_main.cc:38-127, B200 build (+ --ngpu G, see mmvae_alg.hh).
#include "mmvae_alg.hh"

int
main(const int argc, const char *argv[])
{
    mmvae_options_t main_options;
    training_options_t train_opt;
    mmvae::vmf::vmf_options_t vmf_opt;

    CHK(parse_mmvae_options(argc, argv, main_options));
    CHK(parse_vmf_options(argc, argv, vmf_opt));
    CHK(parse_training_options(argc, argv, train_opt));
    if (vmf_opt.vmf_latent) ERR_RET(vmf_opt.latent < 3, "--vmf_latent needs --latent >= 3");

    return run_vae_main(main_options, train_opt, /*is_vmf=*/true, [&](mmv_ctx *ctx, int64_t n_features, int64_t c_dim) {
        std::unique_ptr<mmvae::vmf::vmf_vae_t> model(new mmvae::vmf::vmf_vae_t(
            ctx, n_features, c_dim, vmf_opt.latent, vmf_opt.encoding_layers, vmf_opt.decoding_layers, vmf_opt.kappa_min,
            vmf_opt.kappa_max, vmf_opt.do_relu, train_opt.seed + 7));
        if (vmf_opt.vmf_latent) MMV_OK_EXIT(mmv_model_set_option(model->get(), "vmf_latent", 1));
        return model;
    });
}
