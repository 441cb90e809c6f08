// vmf_vae — same command line as the reference's src/vmf_vae_main.cc:38-127, B200 build.
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

    run_inputs_t in;
    CHK(load_inputs(main_options, in));

    mmvae::device_t dev;
    CHK(dev.upload(train_opt.device, *in.data, in.has_covar ? &in.covar : nullptr, in.c_dim));

    TLOG("Constructing a model");
    mmvae::vmf::vmf_vae_t model(dev.ctx, in.data->nfeature(), in.c_dim, vmf_opt.latent, vmf_opt.encoding_layers,
                                vmf_opt.decoding_layers, vmf_opt.kappa_min, vmf_opt.kappa_max, vmf_opt.do_relu,
                                train_opt.seed + 7);
    if (vmf_opt.vmf_latent) {
        ASSERT(vmf_opt.latent >= 3, "--vmf_latent needs --latent >= 3");
        MMV_OK_EXIT(mmv_model_set_option(model.get(), "vmf_latent", 1));
    }

    TLOG("Training the model...");
    mmvae::latent_recorder_t recorder(main_options.out, train_opt.max_epoch, dev, /*is_vmf=*/true, train_opt.record_params);
    vmf_loss_t loss;
    loss.min_rate = main_options.kl_min;
    loss.max_rate = main_options.kl_max;
    loss.time_discount = main_options.kl_discount;

    auto scores = train_vae_model(model, recorder, *in.data, dev, train_opt, loss);
    CHK(mmvae::write_vector_file(main_options.out + ".scores.gz", scores));
    TLOG("Done");
    return EXIT_SUCCESS;
}
