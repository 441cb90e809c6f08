// nb_vae — same command line as the reference's src/nb_vae_main.cc:39-133, B200 build.
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

    run_inputs_t in;
    CHK(load_inputs(main_options, in));

    mmvae::device_t dev;
    CHK(dev.upload(train_opt.device, *in.data, in.has_covar ? &in.covar : nullptr, in.c_dim));

    TLOG("Constructing a model");
    mmvae::nb::nbvae_t model(dev.ctx, in.data->nfeature(), in.c_dim, nb_opt.mean_encoding_layers,
                             nb_opt.mean_decoding_layers, nb_opt.mean_latent, nb_opt.overdispersion_encoding,
                             nb_opt.overdispersion_latent, nb_opt.do_relu, train_opt.seed + 7);

    TLOG("Training the model...");
    mmvae::latent_recorder_t recorder(main_options.out, train_opt.max_epoch, dev, /*is_vmf=*/false, train_opt.record_params);
    nb_loss_t loss;
    loss.min_rate = main_options.kl_min;
    loss.max_rate = main_options.kl_max;
    loss.time_discount = main_options.kl_discount;

    auto scores = train_vae_model(model, recorder, *in.data, dev, train_opt, loss);
    CHK(mmvae::write_vector_file(main_options.out + ".scores.gz", scores));
    TLOG("Done");
    return EXIT_SUCCESS;
}
