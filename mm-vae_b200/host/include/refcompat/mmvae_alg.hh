// Reference header name -> this build (include/mmvae_alg.hh: training_options_t, train_vae_model).
#ifndef MMVAE_B200_REFCOMPAT_MMVAE_ALG_HH_
#define MMVAE_B200_REFCOMPAT_MMVAE_ALG_HH_
#include "../mmvae_alg.hh"
#endif
